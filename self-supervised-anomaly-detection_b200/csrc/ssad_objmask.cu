// obj_mask (reference dataset_generator.py:27-39) for a batch, on device:
//
//   convert('L') -> skimage.feature.canny(sigma=1.5, low=5, high=15) -> binary_dilation(3x3) -> binary_closing(3x3)
//   -> binary_fill_holes(3x3) -> binary_erosion(4x4) -> largest 8-connected component (degenerate: no foreground ->
//   everything), bit-packed.
//
// Two kernels.
//   canny_nms_kernel   float64, one CTA per 64x32 tile: L -> separable Gaussian (scipy correlate1d order: centre tap,
//                      then the symmetric pairs from the outermost inwards, every product and sum rounded on its own)
//                      / bleed-over normalisation -> Sobel (reflect) -> magnitude -> bilinear non-maximum suppression
//                      -> two bit planes (>= low, >= high) written with warp ballots.  FP64-pipe bound: ~125
//                      non-fusable double operations per pixel (no FMA: the CPU libraries round every op).
//   mask_planes_kernel one warp per image, one lane per 32-pixel word of a row (W <= 1024): every later step is a
//                      reachability or a shift on bit planes.  Connectivity (hysteresis, hole filling, component
//                      extraction) = alternating down / up sweeps; one row step floods a whole row at once with
//                      the carry trick  reach = (M & ~(M + S)) | S  inside a word and a ballot + add carry chain
//                      across the 32 words, until a sweep pair changes nothing.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "../../include/ssad.h"

int ssad_set_cuda_error(cudaError_t e);

namespace ssad {

struct CannyK {
    double w[13];          // scipy _gaussian_kernel1d(1.5, 0, 6)
    double low, high, eps;
};

// ------------------------------------------------------------------ canny: tiles
constexpr int CT_W = 64, CT_H = 32, CT_NT = 256;
constexpr int G_H = CT_H + 16, G_W = CT_W + 16, G_P = G_W + 1;   // L / 255 with an 8-pixel halo
constexpr int V_H = CT_H + 4, V_P = G_P;                        // after the axis-0 pass (rows: halo 2)
constexpr int S_H = CT_H + 4, S_W = CT_W + 4, S_P = S_W + 1;     // smoothed (halo 2)
constexpr int M_H = CT_H + 2, M_W = CT_W + 2, M_P = M_W + 1;     // magnitude (halo 1)
constexpr int IJ_P = CT_W + 1;
constexpr int TBL_C = 13;

// correlate1d, symmetric 13-tap (scipy ni_filters.c NI_Correlate1D): tmp = c * w[6]; tmp += (l + r) * w[j], j outermost first
__device__ __forceinline__ double sym13(const double* v, const double* w) {
    double t = __dmul_rn(v[6], w[6]);
    #pragma unroll
    for (int j = 0; j < 6; ++j) t = __dadd_rn(t, __dmul_rn(__dadd_rn(v[j], v[12 - j]), w[j]));
    return t;
}

// a / c correctly rounded, with rc = RN(1 / c) at hand (smoothed / (bleed_over + eps): one of a few divisors per tile).
// q0 = RN(a * rc) is within 1.5 ulp; one residual step makes it faithful, the second one correctly rounded (Markstein's
// theorem: y = RN(1/b), q faithful => RN(q + y * (a - b*q)) = RN(a/b); no overflow / underflow here: a in [0, 1.1],
// c in [0.2, 1.1]).  Five FMA-pipe instructions instead of the ~30 of a full division.
#ifdef SSAD_CANNY_FULL_DIV
__device__ __forceinline__ double div_by_table(double a, double c, double) { return __ddiv_rn(a, c); }
#else
__device__ __forceinline__ double div_by_table(double a, double c, double rc) {
    const double q0 = __dmul_rn(a, rc);
    const double q1 = __fma_rn(__fma_rn(-q0, c, a), rc, q0);
    return __fma_rn(__fma_rn(-q1, c, a), rc, q1);
}
#endif

__device__ __forceinline__ int xclass(int x, int n) { return x < 6 ? x : (x >= n - 6 ? 12 - (n - 1 - x) : 6); }
__device__ __forceinline__ int xrep(int c, int n) { return c < 6 ? c : (c == 6 ? 6 : n - 1 - (12 - c)); }

struct CannySmem {
    double V[V_H * V_P];        // Gaussian along axis 0; later: M, the gradient magnitude (halo 1)
    double S[S_H * S_P];        // smoothed (halo 2)
    double tbl[S_H * TBL_C];    // bleed-over + eps per (row, column class)
    double rtbl[S_H * TBL_C];   // its correctly rounded reciprocal
    double lut[256];            // v / 255.0
    uint8_t G8[G_H * G_W];      // L with an 8-pixel halo (0 outside the image = 0.0 after the table: mode='constant')
};
static_assert(M_H * M_P <= V_H * V_P, "the magnitude tile reuses the axis-0 buffer");

// ndi.sobel (mode='reflect': neighbours clamp to the image) of the smoothed tile at image pixel (y, x)
__device__ __forceinline__ void sobel_at(const double* S, int y, int x, int y0, int x0, int H, int W, double& is, double& js) {
    const int ru = max(y - 1, 0) - (y0 - 2), rc = y - (y0 - 2), rd = min(y + 1, H - 1) - (y0 - 2);
    const int cl = max(x - 1, 0) - (x0 - 2), cc = x - (x0 - 2), cr = min(x + 1, W - 1) - (x0 - 2);
    const double s_ul = S[ru * S_P + cl], s_uc = S[ru * S_P + cc], s_ur = S[ru * S_P + cr];
    const double s_cl = S[rc * S_P + cl], s_cr = S[rc * S_P + cr];
    const double s_dl = S[rd * S_P + cl], s_dc = S[rd * S_P + cc], s_dr = S[rd * S_P + cr];
    // isobel: [-1,0,1] along axis 0, then [1,2,1] along axis 1
    const double dl = __dsub_rn(s_dl, s_ul), dc = __dsub_rn(s_dc, s_uc), dr = __dsub_rn(s_dr, s_ur);
    is = __dadd_rn(__dmul_rn(dc, 2.0), __dadd_rn(dl, dr));
    // jsobel: [-1,0,1] along axis 1, then [1,2,1] along axis 0
    const double eu = __dsub_rn(s_ur, s_ul), ec = __dsub_rn(s_cr, s_cl), ed = __dsub_rn(s_dr, s_dl);
    js = __dadd_rn(__dmul_rn(ec, 2.0), __dadd_rn(eu, ed));
}

// resident CTAs per SM the kernel is compiled for: 55 KB of shared memory per CTA and 63 registers allow four
// (measured: 2 CTAs 27.1 ms, 3 CTAs 24.9 ms, 4 CTAs 23.5 ms for 1024 images of 1024 x 1024, planes kernel included)
#ifndef SSAD_CANNY_MINB
#define SSAD_CANNY_MINB 4
#endif

__global__ void __launch_bounds__(CT_NT, SSAD_CANNY_MINB)
canny_nms_kernel(const __grid_constant__ CannyK K, int H, int W, const uint8_t* __restrict__ src, long long src_pitch,
                 long long src_stride, uint32_t* __restrict__ low_plane, uint32_t* __restrict__ high_plane, int nw,
                 long long plane_stride /* words per image between its planes' starts */) {
    extern __shared__ __align__(16) uint8_t smem_raw[];
    CannySmem& sm = *reinterpret_cast<CannySmem*>(smem_raw);
    const int tid = threadIdx.x;
    const int x0 = blockIdx.x * CT_W, y0 = blockIdx.y * CT_H, b = blockIdx.z;
    const uint8_t* img = src + (size_t)b * src_stride;

    // ---- tables: v / 255 (img_as_float) and the bleed-over divisor of every smoothed row / column class
    sm.lut[tid] = __ddiv_rn((double)tid, 255.0);
    {
        const bool edge_x = (x0 - 2 < 6) || (x0 + CT_W + 2 > W - 6);
        for (int idx = tid; idx < S_H * TBL_C; idx += CT_NT) {
            const int r = idx / TBL_C, c = idx - r * TBL_C;
            const int y = y0 - 2 + r;
            if (y < 0 || y >= H || (c != 6 && !edge_x)) continue;
            double v[13];
            #pragma unroll
            for (int k = 0; k < 13; ++k) v[k] = (y + k - 6 >= 0 && y + k - 6 < H) ? 1.0 : 0.0;
            const double a = sym13(v, K.w);                  // gaussian_filter(ones): axis 0 ...
            const int x = xrep(c, W);
            #pragma unroll
            for (int k = 0; k < 13; ++k) v[k] = (x + k - 6 >= 0 && x + k - 6 < W) ? a : 0.0;
            const double dv = __dadd_rn(sym13(v, K.w), K.eps);   // ... then axis 1, + eps
            sm.tbl[idx] = dv;
            sm.rtbl[idx] = __drcp_rn(dv);
        }
    }
    __syncthreads();

    // ---- L = convert('L') (Convert.c); 0 outside the image (mode='constant'); / 255 happens at the first read
    // (five pixels per thread and round, their byte loads issued together: at two CTAs per SM nothing else hides the
    // memory latency of this phase)
    constexpr int LB = 5;
    static_assert((G_H * G_W) % (CT_NT * LB) == 0, "tile pixels are dealt in whole rounds");
    for (int i0 = tid; i0 < G_H * G_W; i0 += CT_NT * LB) {
        uint32_t pr[LB], pg[LB], pb[LB];
        int at[LB];
        #pragma unroll
        for (int k = 0; k < LB; ++k) {
            const int i = i0 + k * CT_NT;
            const int r = i / G_W, c = i - r * G_W;
            const int y = y0 - 8 + r, x = x0 - 8 + c;
            at[k] = i;
            pr[k] = pg[k] = pb[k] = 0;
            if (y >= 0 && y < H && x >= 0 && x < W) {
                const uint8_t* q = img + (size_t)y * src_pitch + (size_t)x * 3;
                pr[k] = q[0]; pg[k] = q[1]; pb[k] = q[2];
            }
        }
        #pragma unroll
        for (int k = 0; k < LB; ++k) {
            sm.G8[at[k]] = (uint8_t)((19595u * pr[k] + 38470u * pg[k] + 7471u * pb[k] + 0x8000u) >> 16);
        }
    }
    __syncthreads();

    // ---- Gaussian, axis 0: V[r][c] over G rows r .. r+12; a thread slides down 12 rows of one column
    for (int it = tid; it < G_W * 3; it += CT_NT) {
        const int c = it % G_W, seg = it / G_W;
        double v[24];
        #pragma unroll
        for (int k = 0; k < 24; ++k) v[k] = sm.lut[sm.G8[(seg * 12 + k) * G_W + c]];
        #pragma unroll
        for (int o = 0; o < 12; ++o) sm.V[(seg * 12 + o) * V_P + c] = sym13(v + o, K.w);
    }
    __syncthreads();

    // ---- Gaussian, axis 1, then / (bleed_over + eps): S[r][c] over V cols c .. c+12; 10 outputs per thread
    for (int it = tid; it < S_H * 7; it += CT_NT) {
        const int r = it % S_H, seg = it / S_H;
        const int y = y0 - 2 + r;
        if (y < 0 || y >= H) continue;
        const int c0 = seg * 10;
        double v[22];
        #pragma unroll
        for (int k = 0; k < 22; ++k) v[k] = (c0 + k < G_W) ? sm.V[r * V_P + c0 + k] : 0.0;
        #pragma unroll
        for (int o = 0; o < 10; ++o) {
            const int c = c0 + o, x = x0 - 2 + c;
            if (c < S_W && x >= 0 && x < W)
                sm.S[r * S_P + c] = div_by_table(sym13(v + o, K.w), sm.tbl[r * TBL_C + xclass(x, W)], sm.rtbl[r * TBL_C + xclass(x, W)]);
        }
    }
    __syncthreads();

    // ---- ndi.sobel and the magnitude, halo 1, into the axis-0 buffer (free since the barrier above)
    double* const M = sm.V;
    for (int i = tid; i < M_H * M_W; i += CT_NT) {
        const int r = i / M_W, c = i - r * M_W;
        const int y = y0 - 1 + r, x = x0 - 1 + c;
        double mag = 0.0;
        if (y >= 0 && y < H && x >= 0 && x < W) {
            double is, js;
            sobel_at(sm.S, y, x, y0, x0, H, W, is, js);
            mag = __dsqrt_rn(__dadd_rn(__dmul_rn(is, is), __dmul_rn(js, js)));
        }
        M[r * M_P + c] = mag;
    }
    __syncthreads();

    // ---- _nonmaximum_suppression_bilinear + the two thresholds; a warp owns 32 pixels of a row
    const int lane = tid & 31, warp = tid >> 5;
    for (int task = warp; task < CT_H * (CT_W / 32); task += CT_NT / 32) {
        const int r = task >> 1, half = task & 1;
        const int c = half * 32 + lane;
        const int y = y0 + r, x = x0 + c;
        bool lo = false, hi = false;
        if (y >= 1 && y <= H - 2 && x >= 1 && x <= W - 2) {        // eroded_mask: the 1-pixel border never qualifies
            const double m = M[(r + 1) * M_P + (c + 1)];
            if (m >= K.low) {
                double is, js;                   // only the few pixels above the low threshold need the gradient again
                sobel_at(sm.S, y, x, y0, x0, H, W, is, js);
                const bool is_down = is <= 0.0, is_up = is >= 0.0, is_left = js <= 0.0, is_right = js >= 0.0;
                const bool cond1 = (is_up && is_right) || (is_down && is_left);
                const bool cond2 = (is_down && is_right) || (is_up && is_left);
                const double ai = fabs(is), aj = fabs(js);
                auto nb = [&](int dr, int dc) { return M[(r + 1 + dr) * M_P + (c + 1 + dc)]; };
                bool have = false;
                double w = 0.0, n11 = 0.0, n12 = 0.0, n21 = 0.0, n22 = 0.0;
                if (cond1) {
                    have = true;
                    if (ai > aj) { w = __ddiv_rn(aj, ai); n11 = nb(1, 0); n12 = nb(1, 1); n21 = nb(-1, 0); n22 = nb(-1, -1); }
                    else         { w = __ddiv_rn(ai, aj); n11 = nb(0, 1); n12 = nb(1, 1); n21 = nb(0, -1); n22 = nb(-1, -1); }
                } else if (cond2) {
                    have = true;
                    if (ai < aj) { w = __ddiv_rn(ai, aj); n11 = nb(0, 1); n12 = nb(-1, 1); n21 = nb(0, -1); n22 = nb(1, -1); }
                    else         { w = __ddiv_rn(aj, ai); n11 = nb(-1, 0); n12 = nb(-1, 1); n21 = nb(1, 0); n22 = nb(1, -1); }
                }
                if (have) {
                    const double omw = __dsub_rn(1.0, w);
                    const bool c_plus = __dadd_rn(__dmul_rn(n12, w), __dmul_rn(n11, omw)) <= m;
                    const bool c_minus = __dadd_rn(__dmul_rn(n22, w), __dmul_rn(n21, omw)) <= m;
                    if (c_plus && c_minus) { lo = m > 0.0; hi = lo && (m >= K.high); }
                }
            }
        }
        const uint32_t wl = __ballot_sync(0xFFFFFFFFu, lo), wh = __ballot_sync(0xFFFFFFFFu, hi);
        const int xw = (x0 >> 5) + half;
        if (lane == 0 && y < H && xw < nw) {
            low_plane[(size_t)b * plane_stride + (size_t)y * nw + xw] = wl;
            high_plane[(size_t)b * plane_stride + (size_t)y * nw + xw] = wh;
        }
    }
}

// ------------------------------------------------------------------ bit planes: one warp per image
constexpr uint32_t FULL = 0xFFFFFFFFu;

// bit x <- bit x-1 / bit x+1 of the row (zero shifted in at the row ends; words beyond the row are zero)
__device__ __forceinline__ uint32_t shl1(uint32_t v, int lane) {
    uint32_t c = __shfl_up_sync(FULL, v, 1);
    if (lane == 0) c = 0;
    return (v << 1) | (c >> 31);
}
__device__ __forceinline__ uint32_t shr1(uint32_t v, int lane) {
    uint32_t c = __shfl_down_sync(FULL, v, 1);
    if (lane == 31) c = 0;
    return (v >> 1) | (c << 31);
}
__device__ __forceinline__ uint32_t spread3(uint32_t v, int lane) { return v | shl1(v, lane) | shr1(v, lane); }

// All pixels of the row that are connected, inside the row, through set pixels of M to a seed (S subset of M).
__device__ __forceinline__ uint32_t flood_row(uint32_t M, uint32_t S, int lane) {
    // towards higher x: adding the seeds to M ripples a carry up every run that holds one
    uint32_t up = (M & ~(M + S)) | S;
    const uint32_t P = __ballot_sync(FULL, M == FULL);                 // words a carry passes straight through
    {
        const uint32_t G = __ballot_sync(FULL, up >> 31) << 1;          // word j+1 receives a carry from word j
        const uint32_t cin = G | ((P + (G & P)) ^ P);
        if ((cin >> lane) & 1u) {
            const uint32_t S2 = S | (M & 1u);
            up = (M & ~(M + S2)) | S2;
        }
    }
    // towards lower x: the same on the bit-reversed row
    const uint32_t Mr = __brev(M), Sr = __brev(S), Pr = __brev(P);
    uint32_t dn = (Mr & ~(Mr + Sr)) | Sr;
    {
        const uint32_t G = __brev(__ballot_sync(FULL, dn >> 31)) << 1;
        const uint32_t cin = G | ((Pr + (G & Pr)) ^ Pr);
        if ((cin >> (31 - lane)) & 1u) {
            const uint32_t S2 = Sr | (Mr & 1u);
            dn = (Mr & ~(Mr + S2)) | S2;
        }
    }
    return up | __brev(dn);
}

struct Plane {
    uint32_t* p;
    int nw, lane;
    bool on;
    __device__ __forceinline__ uint32_t ld(int y) const { return on ? p[(size_t)y * nw + lane] : 0u; }
    __device__ __forceinline__ void st(int y, uint32_t v) const { if (on) p[(size_t)y * nw + lane] = v; }
};

// Grow R (subset of M) to everything 8-connected to it through M, rows [ylo, yhi].  COMPLEMENT: M = ~plane & rowmask.
template <bool COMPLEMENT>
__device__ void reach(const Plane& Mp, const Plane& Rp, int ylo, int yhi, uint32_t rowmask, int lane) {
    bool first = true;
    for (;;) {
        bool changed = false;
        #pragma unroll 1
        for (int dir = 0; dir < 2; ++dir) {
            uint32_t prev = 0;
            int y = dir == 0 ? ylo : yhi;
            const int step = dir == 0 ? 1 : -1, n = yhi - ylo + 1;
            // (six rows of loads in flight instead of one were measured: no gain, the row step itself is the chain)
            uint32_t Mn = Mp.ld(y), Rn = Rp.ld(y);
            #pragma unroll 1
            for (int i = 0; i < n; ++i, y += step) {
                uint32_t M = COMPLEMENT ? (~Mn & rowmask) : Mn;
                const uint32_t R = Rn;
                if (i + 1 < n) { Mn = Mp.ld(y + step); Rn = Rp.ld(y + step); }      // next row is in flight during the step
                const uint32_t S = R | (M & spread3(prev, lane));
                uint32_t Rnew = R;
                if (first || __any_sync(FULL, S != R)) {
                    Rnew = flood_row(M, S, lane);
                    if (Rnew != R) { Rp.st(y, Rnew); changed = true; }
                }
                prev = Rnew;
            }
            first = false;
        }
        if (!__any_sync(FULL, changed)) break;
    }
}

__global__ void __launch_bounds__(128)
mask_planes_kernel(int batch, int H, int W, int nw, uint32_t* __restrict__ ws, long long plane_words,
                   uint32_t* __restrict__ mask_bits, int mask_pitch_words, long long mask_stride_words) {
    const int lane = threadIdx.x & 31;
    const int b = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (b >= batch) return;
    const bool on = lane < nw;
    const uint32_t rowmask = !on ? 0u : ((lane == nw - 1 && (W & 31)) ? (FULL >> (32 - (W & 31))) : FULL);
    Plane P0{ws + (size_t)b * 3 * plane_words, nw, lane, on};
    Plane P1{P0.p + plane_words, nw, lane, on};
    Plane P2{P1.p + plane_words, nw, lane, on};
    Plane OUT{mask_bits + (size_t)b * mask_stride_words, mask_pitch_words, lane, on};

    // 1. hysteresis: edges = pixels of the low plane 8-connected to a pixel of the high plane (P0 = low, P1 = high -> edges)
    reach<false>(P0, P1, 0, H - 1, rowmask, lane);

    // 2. binary_dilation(3x3): P2 = dilate(P1)
    auto dilate = [&](const Plane& A, const Plane& D) {
        uint32_t up = 0, cur = A.ld(0);
        for (int y = 0; y < H; ++y) {
            const uint32_t dn = (y + 1 < H) ? A.ld(y + 1) : 0u;
            D.st(y, spread3(up | cur | dn, lane) & rowmask);
            up = cur; cur = dn;
        }
    };
    dilate(P1, P2);
    // 3. binary_closing(3x3) = erosion(dilation): P0 = dilate(P2); P1 = erode3(P0), outside the image counts as 0
    dilate(P2, P0);
    {
        uint32_t up = 0, cur = P0.ld(0);
        for (int y = 0; y < H; ++y) {
            const uint32_t dn = (y + 1 < H) ? P0.ld(y + 1) : 0u;
            const uint32_t t = up & cur & dn;
            P1.st(y, t & shl1(t, lane) & shr1(t, lane));
            up = cur; cur = dn;
        }
    }
    // 4. binary_fill_holes: background reachable from outside the image through ~P1 (8-connected); filled = the rest
    for (int y = 0; y < H; ++y) {
        const uint32_t bg = ~P1.ld(y) & rowmask;
        uint32_t seed = bg;
        if (y != 0 && y != H - 1) {
            uint32_t e = 0;
            if (lane == 0) e |= 1u;
            if (lane == nw - 1) e |= 1u << ((W - 1) & 31);
            seed &= e;
        }
        P2.st(y, seed);
    }
    reach<true>(P1, P2, 0, H - 1, rowmask, lane);
    // 5. binary_erosion(4x4), origin 0: offsets -2..+1 in both axes; P0 = erode4(~P2)
    {
        uint32_t f2 = 0, f1 = 0, f0 = ~P2.ld(0) & rowmask;           // rows y-2, y-1, y
        for (int y = 0; y < H; ++y) {
            const uint32_t fn = (y + 1 < H) ? (~P2.ld(y + 1) & rowmask) : 0u;
            const uint32_t t = f2 & f1 & f0 & fn;
            const uint32_t l1 = shl1(t, lane), l2 = shl1(l1, lane);
            P0.st(y, l2 & l1 & t & shr1(t, lane));
            f2 = f1; f1 = f0; f0 = fn;
        }
    }
    // 6. largest 8-connected component (skimage label + argmax(bincount)): peel components in raster order of their
    //    first pixel; strictly-greater keeps the first of equal sizes (= the lowest label).  P0 = remaining, P2 = scratch.
    for (int y = 0; y < H; ++y) { P2.st(y, 0u); OUT.st(y, 0u); }
    int best = 0, bmin = 0, bmax = -1, ys = 0;
    for (;;) {
        uint32_t w = 0;
        for (; ys < H; ++ys) {
            w = P0.ld(ys);
            if (__any_sync(FULL, w != 0u)) break;
        }
        if (ys >= H) break;
        const int sl = __ffs(__ballot_sync(FULL, w != 0u)) - 1;
        const uint32_t seed = (lane == sl) ? (w & (0u - w)) : 0u;
        P2.st(ys, flood_row(w, seed, lane));
        int ymax = ys;
        for (;;) {                                             // sweeps restricted to the rows the component spans
            bool changed = false;
            uint32_t prev = 0;
            for (int y = ys; y < H; ++y) {
                const uint32_t M = P0.ld(y), R = P2.ld(y);
                const uint32_t S = R | (M & spread3(prev, lane));
                uint32_t Rn = R;
                if (__any_sync(FULL, S != R)) {
                    Rn = flood_row(M, S, lane);
                    if (Rn != R) { P2.st(y, Rn); changed = true; }
                }
                prev = Rn;
                if (y > ymax) {
                    if (!__any_sync(FULL, Rn != 0u)) break;
                    ymax = y;
                }
            }
            prev = 0;
            for (int y = ymax; y >= ys; --y) {
                const uint32_t M = P0.ld(y), R = P2.ld(y);
                const uint32_t S = R | (M & spread3(prev, lane));
                uint32_t Rn = R;
                if (__any_sync(FULL, S != R)) {
                    Rn = flood_row(M, S, lane);
                    if (Rn != R) { P2.st(y, Rn); changed = true; }
                }
                prev = Rn;
            }
            if (!__any_sync(FULL, changed)) break;
        }
        int cnt = 0;
        for (int y = ys; y <= ymax; ++y) {
            const uint32_t R = P2.ld(y);
            cnt += __popc(R);
            P0.st(y, P0.ld(y) & ~R);
        }
        for (int o = 16; o > 0; o >>= 1) cnt += __shfl_xor_sync(FULL, cnt, o);
        if (cnt > best) {
            for (int y = bmin; y <= bmax; ++y) OUT.st(y, 0u);
            for (int y = ys; y <= ymax; ++y) OUT.st(y, P2.ld(y));
            best = cnt; bmin = ys; bmax = ymax;
        }
        for (int y = ys; y <= ymax; ++y) P2.st(y, 0u);
    }
    if (best == 0)                                             // no foreground: label 0 wins the argmax -> all True
        for (int y = 0; y < H; ++y) OUT.st(y, rowmask);
}

}  // namespace ssad

// ------------------------------------------------------------------ C ABI
extern "C" size_t ssad_obj_mask_workspace(int32_t batch, int32_t height, int32_t width) {
    if (batch <= 0 || height <= 0 || width <= 0) return 0;
    const size_t nw = (size_t)(width + 31) / 32;
    return (size_t)batch * 3 * (size_t)height * nw * sizeof(uint32_t);
}

extern "C" int ssad_obj_mask(int32_t batch, int32_t height, int32_t width, const uint8_t* src, int64_t src_pitch,
                             int64_t src_stride, const SsadCannyConsts* consts, uint32_t* mask_bits, int32_t mask_pitch_words,
                             int64_t mask_stride_words, void* workspace, size_t workspace_bytes, void* stream) {
    if (batch < 0 || height <= 0 || width <= 0 || !src || !mask_bits || !consts) return SSAD_EINVAL;
    if (src_pitch < 3ll * width || mask_pitch_words < (width + 31) / 32) return SSAD_EINVAL;
    if (batch == 0) return SSAD_OK;
    if (width > 1024 || width < 13 || height < 13) return SSAD_EUNSUPPORTED;   // one lane per row word; 6-pixel border classes
    if (!workspace || workspace_bytes < ssad_obj_mask_workspace(batch, height, width)) return SSAD_EWORKSPACE;
    if (reinterpret_cast<uintptr_t>(workspace) & 15) return SSAD_EALIGN;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    ssad::CannyK K;
    for (int i = 0; i < 13; ++i) K.w[i] = consts->gauss_w[i];
    K.low = consts->low; K.high = consts->high; K.eps = consts->eps;
    const int nw = (width + 31) / 32;
    const long long plane_words = (long long)height * nw;
    uint32_t* ws = static_cast<uint32_t*>(workspace);
    cudaError_t e = cudaFuncSetAttribute(ssad::canny_nms_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(ssad::CannySmem));
    if (e != cudaSuccess) return ssad_set_cuda_error(e);
    dim3 grid((width + ssad::CT_W - 1) / ssad::CT_W, (height + ssad::CT_H - 1) / ssad::CT_H, batch);
    if (grid.z > 65535u) return SSAD_EUNSUPPORTED;
    ssad::canny_nms_kernel<<<grid, ssad::CT_NT, sizeof(ssad::CannySmem), st>>>(K, height, width, src, src_pitch, src_stride, ws,
                                                                             ws + plane_words, nw, 3 * plane_words);
    e = cudaGetLastError();
    if (e != cudaSuccess) return ssad_set_cuda_error(e);
    // (cutting the batch into parts so that this warp-per-image kernel overlaps the next part's canny was measured: no
    // gain, the two kernels compete for the same issue slots)
    ssad::mask_planes_kernel<<<(batch + 3) / 4, 128, 0, st>>>(batch, height, width, nw, ws, plane_words, mask_bits, mask_pitch_words,
                                                            mask_stride_words);
    e = cudaGetLastError();
    if (e != cudaSuccess) return ssad_set_cuda_error(e);
    return SSAD_OK;
}
