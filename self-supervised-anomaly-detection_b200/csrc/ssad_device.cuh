// Device-side building blocks of the anomaly generator (sm_100a).
//
// Each block restates one third-party routine the reference's path runs inside Pillow /
// torchvision (CPU model in oracle/pillow_model.py), in integer or
// individually-rounded float32 arithmetic so that results are bit-identical:
//   * magic-number u8 <-> f32 and Image.blend (Blend.c)            -> jitter / brightness
//   * RGB -> L (Convert.c)                                           -> contrast mean, saturation
//   * polygon_generic scan-line fill (Draw.c)                        -> rect2poly mask
//   * affine_fixed 16.16 nearest resampling (Geometry.c)             -> RandomAffine, scar.rotate
//   * k-th set pixel of a bit-packed mask window                     -> coords_map[randint]
//   * Container clamp (dataset_generator.py:104-144)
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/ssad.h"

namespace ssad {

// ------------------------------------------------------------------ magic-number u8 <-> f32
// A u8 value v is carried as the float 2^23 + v, whose bit pattern is 0x4B0000vv: PRMT builds it
// straight from a packed byte (no I2F), `M - 2^23` is the exact float value, and adding 2^23 with
// round-toward-zero truncates a float in [0, 2^23) back into the same form (no F2I).
constexpr float kMagic = 8388608.0f;       // 2^23
constexpr uint32_t kMagicBits = 0x4B000000u;

__device__ __forceinline__ float magic_from_byte(uint32_t word, int byte) {
    return __uint_as_float(__byte_perm(word, kMagicBits, 0x7540 + byte));
}
__device__ __forceinline__ float magic_from_u8(uint32_t v) { return __uint_as_float(kMagicBits | v); }
__device__ __forceinline__ uint32_t magic_to_u8(float m) { return __float_as_uint(m) & 0xFFu; }
// pack the low bytes of four magic floats into one little-endian word
__device__ __forceinline__ uint32_t magic_pack4(float a, float b, float c, float d) {
    uint32_t lo = __byte_perm(__float_as_uint(a), __float_as_uint(b), 0x0040);
    uint32_t hi = __byte_perm(__float_as_uint(c), __float_as_uint(d), 0x0040);
    return __byte_perm(lo, hi, 0x5410);
}
// Blend.c tail: (0<=alpha<=1) ? trunc(t) : clip(trunc(t)) -> magic form.  CLAMP is compile time.
template <bool CLAMP>
__device__ __forceinline__ float magic_quantize(float t) {
    if (CLAMP) t = fminf(fmaxf(t, 0.0f), 255.0f);
    return __fadd_rz(t, kMagic);
}
// Image.blend(A, B, alpha) for one channel: t = f32(A + f32(alpha * f32(B - A))).
// mB, mA magic forms; aF = float value of A.
template <bool CLAMP>
__device__ __forceinline__ float magic_blend(float mB, float mA, float aF, float alpha) {
    float d = __fsub_rn(mB, mA);               // exact integer difference
    float p = __fmul_rn(alpha, d);
    return magic_quantize<CLAMP>(__fadd_rn(aF, p));
}
// ImageEnhance.Brightness: A == 0
template <bool CLAMP>
__device__ __forceinline__ float magic_brightness(float mB, float alpha) {
    float v = __fsub_rn(mB, kMagic);
    float t = __fmul_rn(alpha, v);
    if (CLAMP) t = fminf(t, 255.0f);            // v >= 0 and alpha > 1 here: only the upper clip can bind
    return __fadd_rz(t, kMagic);
}
__device__ __forceinline__ bool blend_needs_clip(float alpha) { return !(alpha >= 0.0f && alpha <= 1.0f); }

// scalar u8 brightness (patch colour under the similarity branch, datasets.py:298-299)
__device__ __forceinline__ uint32_t brightness_u8(uint32_t v, float alpha) {
    float t = __fmul_rn(alpha, (float)v);
    if (blend_needs_clip(alpha)) t = fminf(fmaxf(t, 0.0f), 255.0f);
    return (uint32_t)__float2int_rz(t);
}

// Convert.c rgb2l on magic bit patterns: coefficients sum to 65536, so the 0x4B000000 offsets cancel
// modulo 2^32 and the accumulator equals 19595 R + 38470 G + 7471 B + 0x8000.
__device__ __forceinline__ uint32_t luma_acc(float mr, float mg, float mb) {
    return 19595u * __float_as_uint(mr) + 38470u * __float_as_uint(mg) + 7471u * __float_as_uint(mb) + 0x8000u;
}
__device__ __forceinline__ uint32_t luma_u8(uint32_t r, uint32_t g, uint32_t b) {
    return (19595u * r + 38470u * g + 7471u * b + 0x8000u) >> 16;
}

// ------------------------------------------------------------------ affine_fixed (Geometry.c)
__device__ __forceinline__ void affine_fixed_map(const int32_t* a, int x, int y, int& xin, int& yin) {
    int xx = a[2] + y * a[1] + x * a[0];
    int yy = a[5] + y * a[4] + x * a[3];
    xin = xx >> 16;
    yin = yy >> 16;
}

// ------------------------------------------------------------------ Container clamp
__device__ __forceinline__ void clamp_to_container(const int32_t* box, int pw, int ph, int cx, int cy, int& pl, int& pt) {
    pl = cx - pw / 2;            // int(pw/2): pw > 0 so C division == trunc
    pt = cy - ph / 2;
    int pr = cx + pw / 2, pb = cy + ph / 2;
    if (pr > box[2]) pl = box[2] - pw;
    if (pb > box[3]) pt = box[3] - ph;
    if (pl < box[0]) pl = box[0];
    if (pt < box[1]) pt = box[1];
}

// ------------------------------------------------------------------ bit-packed mask window
// popcount of bits [x0, x0 + w) of a mask row
__device__ __forceinline__ int row_window_popc(const uint32_t* __restrict__ row, int x0, int w) {
    int x1 = x0 + w - 1;
    int w0 = x0 >> 5, w1 = x1 >> 5;
    int c = 0;
    for (int j = w0; j <= w1; ++j) {
        uint32_t v = __ldg(row + j);
        if (j == w0) v &= 0xFFFFFFFFu << (x0 & 31);
        if (j == w1) v &= 0xFFFFFFFFu >> (31 - (x1 & 31));
        c += __popc(v);
    }
    return c;
}
// x (relative to x0) of the k-th (0-based) set bit of bits [x0, x0 + w); k must be < popcount
__device__ __forceinline__ int row_window_select(const uint32_t* __restrict__ row, int x0, int w, int k) {
    int x1 = x0 + w - 1;
    int w0 = x0 >> 5, w1 = x1 >> 5;
    for (int j = w0; j <= w1; ++j) {
        uint32_t v = __ldg(row + j);
        if (j == w0) v &= 0xFFFFFFFFu << (x0 & 31);
        if (j == w1) v &= 0xFFFFFFFFu >> (31 - (x1 & 31));
        int c = __popc(v);
        if (k < c) return (j << 5) + (int)__fns(v, 0, k + 1) - x0;
        k -= c;
    }
    return -1;
}

// ------------------------------------------------------------------ polygon_generic (Draw.c)
struct PolyEdge {
    int xmin, xmax, ymin, ymax, x0, y0;
    float dx;
};

struct Polygon {
    PolyEdge e[8];
    int n;        // edges
    int ymin, ymax;
};

__device__ __forceinline__ float edge_x(const PolyEdge& e, int y) {
    return __fadd_rn(__fmul_rn((float)(y - e.y0), e.dx), (float)e.x0);
}

// edges i -> i+1, closed unless last == first (ImagingDrawPolygon); canvas height h clips [ymin, ymax]
__device__ inline void polygon_setup(Polygon& P, const int16_t (*xy)[2], int nv, int h) {
    P.n = 0;
    int ymin = h - 1, ymax = 0;
    for (int i = 0; i < nv; ++i) {
        int j = i + 1;
        if (j == nv) {
            if (xy[i][0] == xy[0][0] && xy[i][1] == xy[0][1]) break;
            j = 0;
        }
        int x0 = xy[i][0], y0 = xy[i][1], x1 = xy[j][0], y1 = xy[j][1];
        PolyEdge& e = P.e[P.n++];
        e.xmin = min(x0, x1); e.xmax = max(x0, x1);
        e.ymin = min(y0, y1); e.ymax = max(y0, y1);
        e.dx = (y0 == y1) ? 0.0f : __fdiv_rn((float)(x1 - x0), (float)(y1 - y0));
        e.x0 = x0; e.y0 = y0;
        ymin = min(ymin, e.ymin);
        ymax = max(ymax, e.ymax);
    }
    P.ymin = max(ymin, 0);
    P.ymax = min(ymax, h);
}

__device__ __forceinline__ int poly_round_up(float f) {
    return (f >= 0.0f) ? (int)floorf(__fadd_rn(f, 0.5f)) : -(int)floorf(__fadd_rn(fabsf(f), 0.5f));
}
__device__ __forceinline__ int poly_round_down(float f) {
    return (f >= 0.0f) ? (int)ceilf(__fsub_rn(f, 0.5f)) : -(int)ceilf(__fsub_rn(fabsf(f), 0.5f));
}

// set bits [x0, x1] (already clipped by the caller's hline rules) in a row of words
__device__ __forceinline__ void row_set_bits(uint32_t* row, int x0, int x1) {
    for (int j = x0 >> 5; j <= (x1 >> 5); ++j) {
        uint32_t m = 0xFFFFFFFFu;
        if (j == (x0 >> 5)) m &= 0xFFFFFFFFu << (x0 & 31);
        if (j == (x1 >> 5)) m &= 0xFFFFFFFFu >> (31 - (x1 & 31));
        row[j] |= m;
    }
}
// hline8/hline32 of Draw.c restricted to one known row
__device__ __forceinline__ void poly_hline(uint32_t* row, int x0, int x1, int w) {
    if (x0 < 0) x0 = 0; else if (x0 >= w) return;
    if (x1 < 0) return; else if (x1 >= w) x1 = w - 1;
    if (x0 <= x1) row_set_bits(row, x0, x1);
}

// One canvas row y (0 <= y < h) of the fill; `row` holds ceil(w/32) zero-initialised words.
__device__ inline void polygon_fill_row(const Polygon& P, int y, int w, uint32_t* row) {
    // horizontal edges are drawn immediately and never enter the edge table
    for (int i = 0; i < P.n; ++i)
        if (P.e[i].ymin == P.e[i].ymax && P.e[i].ymin == y) poly_hline(row, P.e[i].xmin, P.e[i].xmax, w);
    if (y < P.ymin || y > P.ymax) return;
    float xx[16];
    int tab[8];
    int nt = 0;
    for (int i = 0; i < P.n; ++i)
        if (P.e[i].ymin != P.e[i].ymax) tab[nt++] = i;
    int j = 0;
    for (int i = 0; i < nt; ++i) {
        const PolyEdge& cur = P.e[tab[i]];
        if (y >= cur.ymin && y <= cur.ymax) {
            float xc = edge_x(cur, y);
            xx[j++] = xc;
            if (y == cur.ymax && y < P.ymax) {
                xx[j++] = xc;                       // duplicated endpoint
            } else if (cur.dx != 0.0f && (j & 1) && rintf(xc) == xc) {
                for (int k = 0; k < i; ++k) {        // corner-joining fix-up
                    const PolyEdge& other = P.e[tab[k]];
                    if ((cur.dx > 0.0f && other.dx <= 0.0f) || (cur.dx < 0.0f && other.dx >= 0.0f)) continue;
                    if (xc == edge_x(other, y)) {
                        int off = (y == P.ymax) ? -1 : 1;
                        float ax = edge_x(cur, y + off), ao = edge_x(other, y + off);
                        float nv;
                        if (y == cur.ymax) nv = cur.dx > 0.0f ? __fadd_rn(fmaxf(ax, ao), 1.0f) : __fsub_rn(fminf(ax, ao), 1.0f);
                        else               nv = cur.dx > 0.0f ? fminf(ax, ao) : __fadd_rn(fmaxf(ax, ao), 1.0f);
                        if (k < j) xx[k] = nv;
                        break;
                    }
                }
            }
        }
    }
    for (int a = 1; a < j; ++a) {                   // qsort(xx): insertion sort, j <= 16
        float v = xx[a];
        int b = a - 1;
        while (b >= 0 && xx[b] > v) { xx[b + 1] = xx[b]; --b; }
        xx[b + 1] = v;
    }
    int x_pos = (j == 0) ? -1 : 0;
    for (int i = 1; i < j; i += 2) {
        int x_end = poly_round_down(xx[i]);
        if (x_end < x_pos) continue;
        int x_start = poly_round_up(xx[i - 1]);
        if (x_pos > x_start) {
            x_start = x_pos;
            if (x_end < x_start) continue;
        }
        poly_hline(row, x_start, x_end, w);
        x_pos = x_end + 1;
    }
}

// ------------------------------------------------------------------ ImageDraw.line (Draw.c ImagingDrawLine / ImagingDrawWideLine)
// Label 3 "long line" (datasets.py:383-388): a poly-line of integer vertices (ImageDraw truncates the floats it is given),
// width 1 (patch mode) or 3.  The line is rasterised once into coverage bit rows over its bounding box (clipped to the
// image); the generator then pastes a flat colour behind that coverage like a polygon patch.
struct LineBox {
    int x0, y0, w, h;            // box of every pixel the poly-line can touch, inside the image (w <= 0: nothing)
};
__device__ __forceinline__ int draw_round_up(double f) {      // Draw.c ROUND_UP on a double
    return (f >= 0.0) ? (int)floor(f + 0.5) : -(int)floor(fabs(f) + 0.5);
}
__device__ __forceinline__ int draw_round_down(double f) {    // Draw.c ROUND_DOWN on a double
    return (f >= 0.0) ? (int)ceil(f - 0.5) : -(int)ceil(fabs(f) - 0.5);
}
__device__ inline LineBox line_box(const int16_t* __restrict__ xy, int n, int width, int W, int H) {
    LineBox B = {0, 0, 0, 0};
    if (n < 2) return B;                                      // fewer than two points: _draw_lines draws nothing
    int xmin = 32767, xmax = -32768, ymin = 32767, ymax = -32768;
    for (int i = 0; i < n; ++i) {
        xmin = min(xmin, (int)xy[2 * i]); xmax = max(xmax, (int)xy[2 * i]);
        ymin = min(ymin, (int)xy[2 * i + 1]); ymax = max(ymax, (int)xy[2 * i + 1]);
    }
    const int m = width > 1 ? width / 2 + 1 : 0;             // the quad of a wide segment reaches ROUND_UP((width-1)/2) further
    const int x0 = max(xmin - m, 0), x1 = min(xmax + m, W - 1), y0 = max(ymin - m, 0), y1 = min(ymax + m, H - 1);
    if (x1 < x0 || y1 < y0) return B;
    B.x0 = x0; B.y0 = y0; B.w = x1 - x0 + 1; B.h = y1 - y0 + 1;
    return B;
}
__device__ __forceinline__ void line_set(uint32_t* cov, int pitch, const LineBox& B, int x, int y) {
    x -= B.x0; y -= B.y0;
    if ((unsigned)x < (unsigned)B.w && (unsigned)y < (unsigned)B.h) atomicOr(&cov[(size_t)y * pitch + (x >> 5)], 1u << (x & 31));
}
// line8 / line32: one width-1 segment, its end point excluded
__device__ inline void line_segment(uint32_t* cov, int pitch, const LineBox& B, int x0, int y0, int x1, int y1) {
    int dx = x1 - x0, dy = y1 - y0, xs = 1, ys = 1;
    if (dx < 0) { dx = -dx; xs = -1; }
    if (dy < 0) { dy = -dy; ys = -1; }
    if (dx == 0) {
        for (int i = 0; i < dy; ++i) { line_set(cov, pitch, B, x0, y0); y0 += ys; }
    } else if (dy == 0) {
        for (int i = 0; i < dx; ++i) { line_set(cov, pitch, B, x0, y0); x0 += xs; }
    } else if (dx > dy) {
        const int n = dx;
        dy += dy;
        int e = dy - dx;
        dx += dx;
        for (int i = 0; i < n; ++i) {
            line_set(cov, pitch, B, x0, y0);
            if (e >= 0) { y0 += ys; e -= dx; }
            e += dy; x0 += xs;
        }
    } else {
        const int n = dy;
        dx += dx;
        int e = dx - dy;
        dy += dy;
        for (int i = 0; i < n; ++i) {
            line_set(cov, pitch, B, x0, y0);
            if (e >= 0) { x0 += xs; e -= dy; }
            e += dx; y0 += ys;
        }
    }
}
// ImagingDrawWideLine: the polygon of one wide segment, in IMAGE coordinates (the filler's float32 edge arithmetic is
// not translation invariant: x0 + (y - y0) dx rounds differently after a shift); false: zero-length segment (a point)
__device__ inline bool wide_line_polygon(Polygon& P, int H, int x0, int y0, int x1, int y1, int width) {
    const int dx = x1 - x0, dy = y1 - y0;
    if (dx == 0 && dy == 0) return false;
    // hypot(dx, dy): dx^2 + dy^2 is an exact integer, its correctly rounded root is what a correctly rounded hypot returns
    // (and no ratio below can sit within an ulp of the .5 the roundings turn on: dy / big = 1/2 needs dx^2 = 3 dy^2)
    const double big = __dsqrt_rn((double)((long long)dx * dx + (long long)dy * dy)), small = (width - 1) / 2.0;
    const double ratio_max = draw_round_up(small) / big, ratio_min = draw_round_down(small) / big;
    const int dxmin = draw_round_down(ratio_min * dy), dxmax = draw_round_down(ratio_max * dy);
    const int dymin = draw_round_up(ratio_min * dx), dymax = draw_round_up(ratio_max * dx);
    int16_t v[4][2] = {{(int16_t)(x0 - dxmin), (int16_t)(y0 + dymax)}, {(int16_t)(x1 - dxmin), (int16_t)(y1 + dymax)},
                       {(int16_t)(x1 + dxmax), (int16_t)(y1 - dymin)}, {(int16_t)(x0 + dxmax), (int16_t)(y0 - dymin)}};
    polygon_setup(P, v, 4, H);
    return true;
}
// The whole poly-line into `cov` (B.h rows of `pitch` words, zeroed by the caller, visible to every thread: a barrier
// lies between the zeroing and this call); `quads`: shared scratch for n - 1 polygons (width > 1).  Every thread of the
// CTA calls it; the caller synchronises before reading `cov`.
__device__ inline void line_rasterize(const int16_t* __restrict__ xy, int n, int width, int W, int H, const LineBox& B, uint32_t* cov,
                                      int pitch, Polygon* quads, int tid, int nt) {
    if (B.w <= 0 || n < 2) return;
    const int nseg = n - 1;
    if (width <= 1) {
        for (int s = tid; s < nseg; s += nt) line_segment(cov, pitch, B, xy[2 * s], xy[2 * s + 1], xy[2 * s + 2], xy[2 * s + 3]);
        if (tid == 0) line_set(cov, pitch, B, xy[2 * nseg], xy[2 * nseg + 1]);       // _draw_lines: "draw last point"
        return;
    }
    for (int s = tid; s < nseg; s += nt) {
        if (!wide_line_polygon(quads[s], H, xy[2 * s], xy[2 * s + 1], xy[2 * s + 2], xy[2 * s + 3], width)) {
            quads[s].n = -1;
            line_set(cov, pitch, B, xy[2 * s], xy[2 * s + 1]);
        }
    }
    __syncthreads();
    const int iw = (W + 31) >> 5;                                                    // words of an image row (W <= 1024)
    for (int item = tid; item < nseg * B.h; item += nt) {
        const int s = item / B.h, v = item - s * B.h, y = B.y0 + v;
        const Polygon& P = quads[s];
        if (P.n < 0) continue;
        bool touches = y >= P.ymin && y <= P.ymax;
        for (int i = 0; i < P.n && !touches; ++i) touches = P.e[i].ymin == P.e[i].ymax && P.e[i].ymin == y;
        if (!touches) continue;
        uint32_t row[33];                                                            // the image row, then into the box
        for (int j = 0; j <= iw; ++j) row[j] = 0u;
        polygon_fill_row(P, y, W, row);
        const int w0 = B.x0 >> 5, sh = B.x0 & 31;
        for (int j = 0; j < pitch; ++j) {
            uint32_t bits = __funnelshift_r(row[min(w0 + j, iw)], row[min(w0 + j + 1, iw)], sh);
            const int left = B.w - 32 * j;                                           // columns of the box in this word
            if (left < 32) bits &= (1u << left) - 1u;
            if (bits) atomicOr(&cov[(size_t)v * pitch + j], bits);
        }
    }
}

// ------------------------------------------------------------------ block reductions
template <int NT>
__device__ __forceinline__ uint32_t block_sum_u32(uint32_t v, uint32_t* scratch /* >= NT/32 words */) {
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xFFFFFFFFu, v, o);
    __syncthreads();                                  // scratch reuse guard
    if ((threadIdx.x & 31) == 0) scratch[threadIdx.x >> 5] = v;
    __syncthreads();
    uint32_t t = 0;
    #pragma unroll
    for (int i = 0; i < NT / 32; ++i) t += scratch[i];
    return t;
}

// cosine of the mean colours as check_color_similarity + sklearn compute it (float64):
// mean = (sum / n) / 255 per channel; rows L2-normalised; dot.
__device__ __forceinline__ double cosine_of_means(const uint32_t* s1, double n1, const uint32_t* s2, double n2) {
    double a[3], b[3];
    for (int c = 0; c < 3; ++c) {
        a[c] = __ddiv_rn(__ddiv_rn((double)s1[c], n1), 255.0);
        b[c] = __ddiv_rn(__ddiv_rn((double)s2[c], n2), 255.0);
    }
    double na = __dsqrt_rn(__dadd_rn(__dadd_rn(__dmul_rn(a[0], a[0]), __dmul_rn(a[1], a[1])), __dmul_rn(a[2], a[2])));
    double nb = __dsqrt_rn(__dadd_rn(__dadd_rn(__dmul_rn(b[0], b[0]), __dmul_rn(b[1], b[1])), __dmul_rn(b[2], b[2])));
    if (na == 0.0) na = 1.0;
    if (nb == 0.0) nb = 1.0;
    double d = 0.0;
    for (int c = 0; c < 3; ++c) d = __dadd_rn(d, __dmul_rn(__ddiv_rn(a[c], na), __ddiv_rn(b[c], nb)));
    return d;
}

}  // namespace ssad
