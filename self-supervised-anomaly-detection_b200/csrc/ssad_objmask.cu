// obj_mask (reference dataset_generator.py:27-39) for a batch, on device:
//
//   convert('L') -> skimage.feature.canny(sigma=1.5, low=5, high=15) -> binary_dilation(3x3) -> binary_closing(3x3)
//   -> binary_fill_holes(3x3) -> binary_erosion(4x4) -> largest 8-connected component (degenerate: no foreground ->
//   everything), bit-packed.
//
// Four kernels.
//   canny_screen_kernel float32, one CTA per 128x64 tile: the same Gaussian / Sobel chain in fast arithmetic with a proven
//                      error bound; per 64x32 tile it records whether any pixel could reach the low / the high threshold.
//   canny_tiles_kernel one warp per image: the tiles that can hold an edge = tiles with a possible low pixel that are
//                      8-connected, through such tiles, to a tile with a possible high pixel (hysteresis keeps a low
//                      pixel only when a chain of low pixels ties it to a high one) -> a work list.  On the bench
//                      images that is 12 % of the tiles; everything else is exactly zero in both planes.
//   canny_nms_kernel   float64, the listed 64x32 tiles only (persistent CTAs walking the list): L -> separable Gaussian (scipy correlate1d order: centre tap,
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
// warps (bands of rows) per image in the bit-plane kernel
#ifndef SSAD_PLANES_WARPS
#define SSAD_PLANES_WARPS 8
#endif

__global__ void __launch_bounds__(CT_NT, SSAD_CANNY_MINB)
canny_nms_kernel(const __grid_constant__ CannyK K, int H, int W, const uint8_t* __restrict__ src, long long src_pitch,
                 long long src_stride, uint32_t* __restrict__ low_plane, uint32_t* __restrict__ high_plane, int nw,
                 long long plane_stride /* words per image between its planes' starts */,
                 const uint32_t* __restrict__ tile_list, const uint32_t* __restrict__ tile_count, int ntx, int nty) {
    extern __shared__ __align__(16) uint8_t smem_raw[];
    CannySmem& sm = *reinterpret_cast<CannySmem*>(smem_raw);
    const int tid = threadIdx.x;
    sm.lut[tid] = __ddiv_rn((double)tid, 255.0);         // v / 255 (img_as_float)
    // persistent CTAs walk the list of tiles the float32 screen could not rule out (canny_tiles_kernel)
    const uint32_t ntiles = *tile_count;
    for (uint32_t item = blockIdx.x; item < ntiles; item += gridDim.x) {
    const uint32_t code = tile_list[item];
    const int tx = (int)(code % (uint32_t)ntx), ty = (int)((code / (uint32_t)ntx) % (uint32_t)nty), b = (int)(code / (uint32_t)(ntx * nty));
    const int x0 = tx * CT_W, y0 = ty * CT_H;
    const uint8_t* img = src + (size_t)b * src_stride;
    __syncthreads();                                     // the previous tile's shared memory is free (and lut is written)

    // ---- the bleed-over divisor of every smoothed row / column class
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
                // w = num / den (the smaller over the larger gradient component) and the two neighbour pairs
                bool have = false;
                double num = 0.0, den = 1.0, n11 = 0.0, n12 = 0.0, n21 = 0.0, n22 = 0.0;
                if (cond1) {
                    have = true;
                    if (ai > aj) { num = aj; den = ai; n11 = nb(1, 0); n12 = nb(1, 1); n21 = nb(-1, 0); n22 = nb(-1, -1); }
                    else         { num = ai; den = aj; n11 = nb(0, 1); n12 = nb(1, 1); n21 = nb(0, -1); n22 = nb(-1, -1); }
                } else if (cond2) {
                    have = true;
                    if (ai < aj) { num = ai; den = aj; n11 = nb(0, 1); n12 = nb(-1, 1); n21 = nb(0, -1); n22 = nb(1, -1); }
                    else         { num = aj; den = ai; n11 = nb(-1, 0); n12 = nb(-1, 1); n21 = nb(1, 0); n22 = nb(1, -1); }
                }
                if (have) {
                    // The reference decides  RN(RN(n12 w) + RN(n11 RN(1 - w))) <= m  with w = RN(num / den): five roundings,
                    // at most 3.4e-16 (n11 + n12) away from v = (n12 num + n11 (den - num)) / den.  First compare
                    // den v with den m directly (no division; its own rounding error is below 3.4e-16 den (n11 + n12 + m)):
                    // when they differ by more than 1e-13 den (n11 + n12 + m) the reference's answer is known (300-fold
                    // margin); only the near ties run the reference's sequence.
                    const double dmn = __dsub_rn(den, num), rhs = __dmul_rn(m, den);
                    const double dp = __dsub_rn(rhs, __fma_rn(n12, num, __dmul_rn(n11, dmn)));
                    const double dm = __dsub_rn(rhs, __fma_rn(n22, num, __dmul_rn(n21, dmn)));
                    const double tp = __dmul_rn(__dmul_rn(1e-13, den), __dadd_rn(__dadd_rn(n11, n12), m));
                    const double tm = __dmul_rn(__dmul_rn(1e-13, den), __dadd_rn(__dadd_rn(n21, n22), m));
                    bool c_plus = dp > 0.0, c_minus = dm > 0.0;
                    if (!(fabs(dp) > tp) || !(fabs(dm) > tm)) {
                        const double w = __ddiv_rn(num, den), omw = __dsub_rn(1.0, w);
                        c_plus = __dadd_rn(__dmul_rn(n12, w), __dmul_rn(n11, omw)) <= m;
                        c_minus = __dadd_rn(__dmul_rn(n22, w), __dmul_rn(n21, omw)) <= m;
                    }
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
    }   // tile list
}

// ------------------------------------------------------------------ canny: float32 screen
// The float64 chain costs ~125 unfusable double operations per pixel, yet hysteresis (skimage canny: ndi.label of the
// low mask, keep the components holding a pixel >= high) can only keep a pixel whose magnitude m reaches the low
// threshold AND that is 8-connected, through such pixels, to one reaching the high threshold.  This kernel evaluates
// the same real-number chain (L -> 13-tap Gaussian along axis 0, along axis 1, / bleed-over -> Sobel (reflect) ->
// |gradient|) in float32 with fused multiply-adds, in grey-level units (x 255), and flags per 64x32 tile whether any
// pixel has m32 >= low - d resp. m32 >= high - d.
// Error bound (grey levels; every float32 operation has relative error <= 2^-24 = 6e-8, partial sums <= 255, the taps
// sum to 1): axis-0 pass <= 13 * 255 * 6e-8 + tap rounding 255 * 6e-8 = 2.2e-4; axis-1 pass doubles it to 4.4e-4;
// the bleed-over factor is >= 0.63^2 = 0.4 at a corner (x 2.5 -> 1.1e-3), 1 in the interior; Sobel sums six values
// with weights (1, 2, 1) of both signs: 8 x 1.1e-3 + its own rounding 6 * 1020 * 6e-8 = 9.2e-3 per component, so
// |m32 - m| <= 1.9e-2 plus the square / add rounding (2e-7 relative of m <= 1443: 3e-4).  The float64 chain itself
// differs from the real-number value by ~1e-12.  d = 0.1 grey level leaves a five-fold margin.
constexpr int SC_W = 128, SC_H = 64, SC_NT = 448;
constexpr int SG_W = SC_W + 16, SG_H = SC_H + 14;       // L: rows y0-7 .. y0+SC_H+6, columns x0-8 .. x0+SC_W+7
constexpr int SV_H = SC_H + 2;                           // V (axis-0 pass): rows y0-1 .. y0+SC_H, the columns of L
constexpr int SS_W = SC_W + 4;                           // S (smoothed): columns x0-1 .. x0+SC_W (+2: rows stay 16-byte multiples)
constexpr float kScreenMargin = 0.1f;
static_assert(SG_W % 4 == 0 && SS_W % 4 == 0, "vector rows");
static_assert(SV_H % 11 == 0 && (SV_H & 1) == 0, "axis-0 pass: 11-row segments; axis-1 pass: row pairs");

struct ScreenK {
    float w[7];              // taps 0..6 (centre last) of the symmetric 13-tap kernel
    float lo2, hi2;          // (255 low - d)^2, (255 high - d)^2
};
struct ScreenSmem {
    float G[SG_H * SG_W];    // L (grey levels); reused for S once the axis-0 pass is done
    float V[SV_H * SG_W];
    float ra[SV_H], rb[SS_W];   // 1 / bleed-over along the rows / the columns (border tiles)
    uint32_t flags[8];       // [sub-tile 0..3][low, high]
};
static_assert(SV_H * SS_W <= SG_H * SG_W, "S reuses the L buffer");

__device__ __forceinline__ uint32_t om_dp2a_lo(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t d;
    asm("dp2a.lo.u32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}
__device__ __forceinline__ uint32_t om_dp2a_hi(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t d;
    asm("dp2a.hi.u32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}
// Convert.c L of the four pixels held by three words (12 bytes)
__device__ __forceinline__ float4 luma4(uint32_t w0, uint32_t w1, uint32_t w2) {
    constexpr uint32_t kRG = (38470u << 16) | 19595u, kB0 = 7471u;          // pixel at bytes 0,1,2 of a word
    constexpr uint32_t k0R = (19595u << 16), kGB = (7471u << 16) | 38470u;  // pixel at bytes 1,2,3 of a word
    const uint32_t p1 = __byte_perm(w0, w1, 0x0543), p2 = __byte_perm(w1, w2, 0x0432);
    float4 L;
    L.x = (float)(om_dp2a_hi(kB0, w0, om_dp2a_lo(kRG, w0, 0x8000u)) >> 16);
    L.y = (float)(om_dp2a_hi(kB0, p1, om_dp2a_lo(kRG, p1, 0x8000u)) >> 16);
    L.z = (float)(om_dp2a_hi(kB0, p2, om_dp2a_lo(kRG, p2, 0x8000u)) >> 16);
    L.w = (float)(om_dp2a_hi(kGB, w2, om_dp2a_lo(k0R, w2, 0x8000u)) >> 16);
    return L;
}
typedef unsigned long long om_f32x2;
__device__ __forceinline__ om_f32x2 om_pack2(float lo, float hi) {
    om_f32x2 r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void om_unpack2(om_f32x2 v, float& lo, float& hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ om_f32x2 om_fma2(om_f32x2 a, om_f32x2 b, om_f32x2 c) {
    om_f32x2 r;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
    return r;
}
__device__ __forceinline__ om_f32x2 om_mul2(om_f32x2 a, om_f32x2 b) {
    om_f32x2 r;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
__device__ __forceinline__ om_f32x2 om_add2(om_f32x2 a, om_f32x2 b) {
    om_f32x2 r;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
// 13 symmetric taps on packed pairs: w[6] v[6] + sum_j w[j] (v[j] + v[12-j])
__device__ __forceinline__ om_f32x2 om_sym13(const om_f32x2* v, const om_f32x2 (&w)[7]) {
    om_f32x2 t = om_mul2(v[6], w[6]);
    #pragma unroll
    for (int j = 0; j < 6; ++j) t = om_fma2(om_add2(v[j], v[12 - j]), w[j], t);
    return t;
}

__global__ void __launch_bounds__(SC_NT, 2)
canny_screen_kernel(const __grid_constant__ ScreenK K, int H, int W, const uint8_t* __restrict__ src, long long src_pitch,
                    long long src_stride, uint8_t* __restrict__ tile_flags, int ntx, int nty,
                    uint32_t* __restrict__ low_plane, uint32_t* __restrict__ high_plane, int nw, long long plane_stride) {
    extern __shared__ __align__(16) uint8_t smem_raw[];
    ScreenSmem& sm = *reinterpret_cast<ScreenSmem*>(smem_raw);
    const int tid = threadIdx.x;
    const int x0 = blockIdx.x * SC_W, y0 = blockIdx.y * SC_H, b = blockIdx.z;
    const uint8_t* img = src + (size_t)b * src_stride;
    const bool border = (x0 - 8 < 6) || (y0 - 7 < 6) || (x0 + SC_W + 8 > W - 6) || (y0 + SC_H + 7 > H - 6);

    if (tid < 8) sm.flags[tid] = 0u;
    // both planes start at zero: the float64 kernel writes the listed tiles only
    for (int i = tid; i < SC_H * (SC_W / 32); i += SC_NT) {
        const int r = i / (SC_W / 32), xw = (x0 >> 5) + (i - r * (SC_W / 32)), y = y0 + r;
        if (y < H && xw < nw) {
            low_plane[(size_t)b * plane_stride + (size_t)y * nw + xw] = 0u;
            high_plane[(size_t)b * plane_stride + (size_t)y * nw + xw] = 0u;
        }
    }
    if (border) {
        // gaussian_filter(ones): the taps that fall inside the image, per row and per column of the tile
        for (int i = tid; i < SV_H + SS_W; i += SC_NT) {
            const bool is_row = i < SV_H;
            const int p = is_row ? y0 - 1 + i : x0 - 1 + (i - SV_H), n = is_row ? H : W;
            float a = 0.0f;
            #pragma unroll
            for (int k = 0; k < 13; ++k) if (p + k - 6 >= 0 && p + k - 6 < n) a += K.w[k < 7 ? k : 12 - k];
            const float r = (p >= 0 && p < n) ? 1.0f / a : 0.0f;
            if (is_row) sm.ra[i] = r; else sm.rb[i - SV_H] = r;
        }
    }

    // ---- L (Convert.c), four pixels = three words per task; 0 outside the image (mode='constant')
    {
        const bool vec_ok = ((src_pitch & 3) == 0) && ((reinterpret_cast<uintptr_t>(img) & 3) == 0);
        constexpr int NG = SG_W / 4, NTASK = SG_H * NG, NR = (NTASK + SC_NT - 1) / SC_NT;
        uint32_t w[NR][3];
        const bool inside = vec_ok && x0 >= 8 && y0 >= 7 && x0 + SC_W + 8 <= W && y0 + SC_H + 7 <= H;   // (uniform)
        if (inside) {
            // every pixel of the tile and its halo is in the image: no bounds, one pointer per task
            const uint8_t* const base = img + (size_t)(y0 - 7) * src_pitch + (size_t)(x0 - 8) * 3;
            #pragma unroll
            for (int k = 0; k < NR; ++k) {
                const int t = min(tid + k * SC_NT, NTASK - 1);
                const int r = t / NG, g = t - r * NG;
                const uint32_t* q = reinterpret_cast<const uint32_t*>(base + (size_t)r * src_pitch + g * 12);
                w[k][0] = __ldg(q); w[k][1] = __ldg(q + 1); w[k][2] = __ldg(q + 2);
            }
        } else {
        #pragma unroll
        for (int k = 0; k < NR; ++k) {
            const int t = tid + k * SC_NT;
            const int r = t / NG, g = t - r * NG;
            const int y = y0 - 7 + r, x = x0 - 8 + 4 * g;
            w[k][0] = w[k][1] = w[k][2] = 0u;
            if (t < NTASK && y >= 0 && y < H && x + 3 >= 0 && x < W) {
                if (vec_ok && x >= 0 && x + 3 < W) {
                    const uint32_t* q = reinterpret_cast<const uint32_t*>(img + (size_t)y * src_pitch + (size_t)x * 3);
                    w[k][0] = __ldg(q); w[k][1] = __ldg(q + 1); w[k][2] = __ldg(q + 2);
                } else {
                    uint8_t bytes[12];
                    #pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const int xx = x + j;
                        const bool in = xx >= 0 && xx < W;
                        const uint8_t* q = img + (size_t)y * src_pitch + (size_t)(in ? xx : 0) * 3;
                        bytes[3 * j] = in ? q[0] : 0; bytes[3 * j + 1] = in ? q[1] : 0; bytes[3 * j + 2] = in ? q[2] : 0;
                    }
                    #pragma unroll
                    for (int j = 0; j < 3; ++j)
                        w[k][j] = bytes[4 * j] | (bytes[4 * j + 1] << 8) | (bytes[4 * j + 2] << 16) | ((uint32_t)bytes[4 * j + 3] << 24);
                }
            }
        }
        }
        #pragma unroll
        for (int k = 0; k < NR; ++k) {
            const int t = tid + k * SC_NT;
            if (t < NTASK) *reinterpret_cast<float4*>(&sm.G[4 * t]) = luma4(w[k][0], w[k][1], w[k][2]);     // (an all-zero group gives L = 0)
        }
    }
    __syncthreads();

    om_f32x2 wp[7];
    #pragma unroll
    for (int j = 0; j < 7; ++j) wp[j] = om_pack2(K.w[j], K.w[j]);

    // ---- Gaussian along axis 0: a thread slides down 11 rows of a column pair
    {
        constexpr int NP = SG_W / 2, NSEG = SV_H / 11;
        for (int t = tid; t < NP * NSEG; t += SC_NT) {
            const int seg = t / NP, pr = t - seg * NP;
            const float* g = &sm.G[(seg * 11) * SG_W + 2 * pr];
            om_f32x2 v[23];
            #pragma unroll
            for (int k = 0; k < 23; ++k) v[k] = *reinterpret_cast<const om_f32x2*>(g + k * SG_W);
            #pragma unroll
            for (int o = 0; o < 11; ++o)
                *reinterpret_cast<om_f32x2*>(&sm.V[(seg * 11 + o) * SG_W + 2 * pr]) = om_sym13(v + o, wp);
        }
    }
    __syncthreads();

    // ---- Gaussian along axis 1 (/ bleed-over in border tiles): a thread takes four columns of a row; lanes walk along
    // the row, so every 128-bit shared-memory access of a warp is a run of consecutive chunks (no bank conflicts)
    float* const S = sm.G;
    {
        constexpr int NG = SS_W / 4;                     // 33 groups: S columns 0 .. 131 (130, 131 are padding)
        for (int t = tid; t < SV_H * NG; t += SC_NT) {
            const int r = t / NG, g = t - r * NG;
            // S[r][c] = sum_k w[k] V[r][c + 1 + k]: columns 4g+1 .. 4g+16 of V, loaded as the chunks 4g .. 4g+19
            const float4* vp = reinterpret_cast<const float4*>(&sm.V[r * SG_W + 4 * g]);
            float v[20];
            #pragma unroll
            for (int k = 0; k < 5; ++k) {
                const float4 q = vp[k];
                v[4 * k] = q.x; v[4 * k + 1] = q.y; v[4 * k + 2] = q.z; v[4 * k + 3] = q.w;
            }
            float o[4];
            #pragma unroll
            for (int j = 0; j < 4; ++j) {
                float a = v[7 + j] * K.w[6];
                #pragma unroll
                for (int i = 0; i < 6; ++i) a = fmaf(v[1 + j + i] + v[13 + j - i], K.w[i], a);
                o[j] = a;
            }
            if (border) {
                const float ra = sm.ra[r];
                #pragma unroll
                for (int j = 0; j < 4; ++j) o[j] *= ra * sm.rb[4 * g + j];
            }
            *reinterpret_cast<float4*>(&S[r * SS_W + 4 * g]) = make_float4(o[0], o[1], o[2], o[3]);
        }
    }
    __syncthreads();
    if (border) {
        // ndi.sobel mode='reflect': a neighbour outside the image is the pixel itself clamped to the image
        for (int i = tid; i < SS_W; i += SC_NT) {
            if (y0 == 0) S[i] = S[SS_W + i];
            const int rl = H - (y0 - 1);                       // row of image row H
            if (rl >= 1 && rl < SV_H) S[rl * SS_W + i] = S[(rl - 1) * SS_W + i];
        }
        __syncthreads();
        for (int i = tid; i < SV_H; i += SC_NT) {
            if (x0 == 0) S[i * SS_W] = S[i * SS_W + 1];
            const int cl = W - (x0 - 1);
            if (cl >= 1 && cl < SS_W) S[i * SS_W + cl] = S[i * SS_W + cl - 1];
        }
        __syncthreads();
    }

    // ---- Sobel and |gradient|^2 against the two thresholds: a thread takes four pixels of a row
    {
        const int lane = tid & 31;
        static_assert((SC_H * (SC_W / 4)) % 32 == 0 && SC_NT % 32 == 0, "whole warps stay in the loop (ballots)");
        for (int t = tid; t < SC_H * (SC_W / 4); t += SC_NT) {
            const int i = t / (SC_W / 4), g = t - i * (SC_W / 4);
            bool lo = false, hi = false;
            if (i < SC_H && y0 + i < H) {
                float s[3][6];
                #pragma unroll
                for (int r = 0; r < 3; ++r) {
                    const float4 a = *reinterpret_cast<const float4*>(&S[(i + r) * SS_W + 4 * g]);
                    const float2 c = *reinterpret_cast<const float2*>(&S[(i + r) * SS_W + 4 * g + 4]);
                    s[r][0] = a.x; s[r][1] = a.y; s[r][2] = a.z; s[r][3] = a.w; s[r][4] = c.x; s[r][5] = c.y;
                }
                float dc[6], e[3][4];
                #pragma unroll
                for (int c = 0; c < 6; ++c) dc[c] = s[2][c] - s[0][c];
                #pragma unroll
                for (int r = 0; r < 3; ++r)
                    #pragma unroll
                    for (int j = 0; j < 4; ++j) e[r][j] = s[r][j + 2] - s[r][j];
                float mx = 0.0f;
                #pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const float is = fmaf(2.0f, dc[j + 1], dc[j]) + dc[j + 2];
                    const float js = fmaf(2.0f, e[1][j], e[0][j]) + e[2][j];
                    float m2 = fmaf(is, is, js * js);
                    if (border && x0 + 4 * g + j >= W) m2 = 0.0f;
                    mx = fmaxf(mx, m2);
                }
                lo = mx >= K.lo2; hi = mx >= K.hi2;
            }
            // a warp = one row, lanes 0-15 / 16-31 = the two 64-pixel halves
            const uint32_t bl = __ballot_sync(0xFFFFFFFFu, lo), bh = __ballot_sync(0xFFFFFFFFu, hi);
            if (lane == 0 && i < SC_H) {
                const int sy = i >> 5;
                if (bl & 0x0000FFFFu) atomicOr(&sm.flags[(2 * sy) * 2], 1u);
                if (bl & 0xFFFF0000u) atomicOr(&sm.flags[(2 * sy + 1) * 2], 1u);
                if (bh & 0x0000FFFFu) atomicOr(&sm.flags[(2 * sy) * 2 + 1], 1u);
                if (bh & 0xFFFF0000u) atomicOr(&sm.flags[(2 * sy + 1) * 2 + 1], 1u);
            }
        }
    }
    __syncthreads();
    if (tid < 4) {
        const int tx = 2 * blockIdx.x + (tid & 1), ty = 2 * blockIdx.y + (tid >> 1);
        if (tx < ntx && ty < nty)
            tile_flags[((size_t)b * nty + ty) * ntx + tx] = (uint8_t)((sm.flags[2 * tid] ? 1u : 0u) | (sm.flags[2 * tid + 1] ? 2u : 0u));
    }
}

// The tiles that can hold an edge pixel: a possible low pixel in the tile, and a chain of such tiles (8-neighbours) to a
// tile with a possible high pixel (connected pixels lie in the same or in neighbouring tiles).  One warp per image;
// appends the tiles to the work list of the float64 kernel.
__global__ void __launch_bounds__(32)
canny_tiles_kernel(int ntx, int nty, const uint8_t* __restrict__ tile_flags, uint32_t* __restrict__ tile_list,
                   uint32_t* __restrict__ tile_count) {
    extern __shared__ uint32_t tsm[];          // [nty + 2] low rows, [nty + 2] reached rows (one zero row above and below)
    uint32_t* const lo = tsm + 1;
    uint32_t* const R = tsm + (nty + 2) + 1;
    const int lane = threadIdx.x, b = blockIdx.x;
    const uint8_t* f = tile_flags + (size_t)b * nty * ntx;
    if (lane == 0) { lo[-1] = lo[nty] = 0u; R[-1] = R[nty] = 0u; }
    for (int r = lane; r < nty; r += 32) {
        uint32_t l = 0, h = 0;
        for (int c = 0; c < ntx; ++c) {
            const uint32_t v = f[r * ntx + c];
            l |= (v & 1u) << c; h |= ((v >> 1) & 1u) << c;
        }
        lo[r] = l; R[r] = h & l;
    }
    __syncwarp();
    for (;;) {
        bool changed = false;
        for (int r = lane; r < nty; r += 32) {
            uint32_t n = R[r - 1] | R[r] | R[r + 1];
            n = (n | (n << 1) | (n >> 1)) & lo[r];
            if (n != R[r]) { R[r] = n; changed = true; }      // (monotone: the order of the updates does not matter)
        }
        __syncwarp();
        if (!__any_sync(0xFFFFFFFFu, changed)) break;
    }
    for (int r0 = 0; r0 < nty; r0 += 32) {
        const int r = r0 + lane;
        const uint32_t bits = (r < nty) ? R[r] : 0u;
        const int n = __popc(bits);
        int incl = n;
        for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xFFFFFFFFu, incl, o); if (lane >= o) incl += t; }
        const int total = __shfl_sync(0xFFFFFFFFu, incl, 31);
        uint32_t base = 0;
        if (lane == 0 && total > 0) base = atomicAdd(tile_count, (uint32_t)total);
        base = __shfl_sync(0xFFFFFFFFu, base, 0) + (uint32_t)(incl - n);
        uint32_t rest = bits;
        while (rest) {
            const int c = __ffs(rest) - 1;
            rest &= rest - 1;
            tile_list[base++] = (uint32_t)((b * nty + r) * ntx + c);
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
    // (p already points at this lane's word of row 0; a plane is H * nw < 2^31 words)
    __device__ __forceinline__ uint32_t ld(int y) const { return on ? p[y * nw] : 0u; }
    // a row another warp of the CTA writes (band boundary): straight from the L2
    __device__ __forceinline__ uint32_t ldcg(int y) const { return on ? __ldcg(p + y * nw) : 0u; }
    __device__ __forceinline__ void st(int y, uint32_t v) const { if (on) p[y * nw] = v; }
};

// Grow R (subset of M) to everything 8-connected to it through M.  COMPLEMENT: M = ~plane & rowmask.
// The rows of the image are cut into one band per warp of the CTA; a round = every warp sweeps its band down and up
// (one row step floods a whole row, see flood_row), entering with the neighbouring bands' boundary rows as they are;
// rounds repeat until no warp changed anything.  R only grows, so reading a neighbour's row while it is being
// rewritten costs at most another round.  Every thread of the CTA must call this.
template <bool COMPLEMENT>
__device__ void reach(const Plane& Mp, const Plane& Rp, int H, int ylo, int yhi, uint32_t rowmask, int lane) {
    bool first = true;
    for (;;) {
        bool changed = false;
        if (ylo <= yhi) {
            #pragma unroll 1
            for (int dir = 0; dir < 2; ++dir) {
                const int step = dir == 0 ? 1 : -1, n = yhi - ylo + 1;
                int y = dir == 0 ? ylo : yhi;
                uint32_t prev = 0;
                if (dir == 0 ? (ylo > 0) : (yhi + 1 < H)) prev = Rp.ldcg(y - step);
                uint32_t Mn = Mp.ld(y), Rn = Rp.ld(y);
                #pragma unroll 1
                for (int i = 0; i < n; ++i, y += step) {
                    const uint32_t M = COMPLEMENT ? (~Mn & rowmask) : Mn;
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
        }
        if (!__syncthreads_or(changed ? 1 : 0)) break;
    }
}

// The same for one component growing from a seed (largest-component peel): a warp only walks the rows its part of the
// component spans.  [alo, ahi] = rows of the band that hold pixels of R (alo > ahi: none), kept by the caller.
__device__ void reach_component(const Plane& Mp, const Plane& Rp, int H, int ylo, int yhi, int lane, int& alo, int& ahi) {
    for (;;) {
        bool changed = false;
        if (ylo <= yhi) {
            {   // down
                uint32_t prev = (ylo > 0) ? Rp.ldcg(ylo - 1) : 0u;
                int y = __any_sync(FULL, prev != 0u) ? ylo : alo;
                if (y != ylo) prev = 0u;
                #pragma unroll 1
                for (; y <= yhi; ++y) {
                    const uint32_t M = Mp.ld(y), R = Rp.ld(y);
                    const uint32_t S = R | (M & spread3(prev, lane));
                    uint32_t Rnew = R;
                    if (__any_sync(FULL, S != R)) {
                        Rnew = flood_row(M, S, lane);
                        if (Rnew != R) { Rp.st(y, Rnew); changed = true; }
                    }
                    prev = Rnew;
                    if (__any_sync(FULL, Rnew != 0u)) { alo = min(alo, y); ahi = max(ahi, y); }
                    else if (y > ahi) break;
                }
            }
            {   // up
                uint32_t prev = (yhi + 1 < H) ? Rp.ldcg(yhi + 1) : 0u;
                int y = __any_sync(FULL, prev != 0u) ? yhi : ahi;
                if (y != yhi) prev = 0u;
                #pragma unroll 1
                for (; y >= ylo; --y) {
                    const uint32_t M = Mp.ld(y), R = Rp.ld(y);
                    const uint32_t S = R | (M & spread3(prev, lane));
                    uint32_t Rnew = R;
                    if (__any_sync(FULL, S != R)) {
                        Rnew = flood_row(M, S, lane);
                        if (Rnew != R) { Rp.st(y, Rnew); changed = true; }
                    }
                    prev = Rnew;
                    if (__any_sync(FULL, Rnew != 0u)) { alo = min(alo, y); ahi = max(ahi, y); }
                    else if (y < alo) break;
                }
            }
        }
        if (!__syncthreads_or(changed ? 1 : 0)) break;
    }
}

// One CTA per image, one band of rows per warp, one lane per 32-pixel word of a row.
__global__ void __launch_bounds__(512)
mask_planes_kernel(int batch, int H, int W, int nw, uint32_t* __restrict__ ws, long long plane_words,
                   uint32_t* __restrict__ mask_bits, int mask_pitch_words, long long mask_stride_words) {
    __shared__ int s_first, s_cnt;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
    const int b = blockIdx.x;
    const int per = (H + nwarp - 1) / nwarp;
    const int ylo = warp * per, yhi = min(H, ylo + per) - 1;            // this warp's band (empty: ylo > yhi)
    const bool on = lane < nw;
    const uint32_t rowmask = !on ? 0u : ((lane == nw - 1 && (W & 31)) ? (FULL >> (32 - (W & 31))) : FULL);
    const int lo_ = on ? lane : 0;
    Plane P0{ws + (size_t)b * 3 * plane_words + lo_, nw, lane, on};
    Plane P1{P0.p + plane_words, nw, lane, on};
    Plane P2{P1.p + plane_words, nw, lane, on};
    Plane OUT{mask_bits + (size_t)b * mask_stride_words + lo_, mask_pitch_words, lane, on};

    // 1. hysteresis: edges = pixels of the low plane 8-connected to a pixel of the high plane (P0 = low, P1 = high -> edges)
    reach<false>(P0, P1, H, ylo, yhi, rowmask, lane);

    // 2. binary_dilation(3x3): P2 = dilate(P1)   (rows above / below the band belong to other warps: barrier between steps)
    auto dilate = [&](const Plane& A, const Plane& D) {
        if (ylo <= yhi) {
            uint32_t up = ylo > 0 ? A.ld(ylo - 1) : 0u, cur = A.ld(ylo);
            for (int y = ylo; y <= yhi; ++y) {
                const uint32_t dn = (y + 1 < H) ? A.ld(y + 1) : 0u;
                D.st(y, spread3(up | cur | dn, lane) & rowmask);
                up = cur; cur = dn;
            }
        }
        __syncthreads();
    };
    dilate(P1, P2);
    // 3. binary_closing(3x3) = erosion(dilation): P0 = dilate(P2); P1 = erode3(P0), outside the image counts as 0
    dilate(P2, P0);
    if (ylo <= yhi) {
        uint32_t up = ylo > 0 ? P0.ld(ylo - 1) : 0u, cur = P0.ld(ylo);
        for (int y = ylo; y <= yhi; ++y) {
            const uint32_t dn = (y + 1 < H) ? P0.ld(y + 1) : 0u;
            const uint32_t t = up & cur & dn;
            P1.st(y, t & shl1(t, lane) & shr1(t, lane));
            up = cur; cur = dn;
        }
    }
    __syncthreads();
    // 4. binary_fill_holes: background reachable from outside the image through ~P1 (8-connected); filled = the rest
    for (int y = ylo; y <= yhi; ++y) {
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
    __syncthreads();
    reach<true>(P1, P2, H, ylo, yhi, rowmask, lane);
    // 5. binary_erosion(4x4), origin 0: offsets -2..+1 in both axes; P0 = erode4(~P2)
    if (ylo <= yhi) {
        auto fg = [&](int y) { return (y >= 0 && y < H) ? (~P2.ld(y) & rowmask) : 0u; };
        uint32_t f2 = fg(ylo - 2), f1 = fg(ylo - 1), f0 = fg(ylo);        // rows y-2, y-1, y
        for (int y = ylo; y <= yhi; ++y) {
            const uint32_t fn = fg(y + 1);
            const uint32_t t = f2 & f1 & f0 & fn;
            const uint32_t l1 = shl1(t, lane), l2 = shl1(l1, lane);
            P0.st(y, l2 & l1 & t & shr1(t, lane));
            f2 = f1; f1 = f0; f0 = fn;
        }
    }
    __syncthreads();
    // 6. largest 8-connected component (skimage label + argmax(bincount)): peel components in raster order of their
    //    first pixel; strictly-greater keeps the first of equal sizes (= the lowest label).  P0 = remaining, P1 = scratch.
    //    (P2 is still being read by slower warps' erosion above until the barrier; P1 is free.)
    for (int y = ylo; y <= yhi; ++y) { P1.st(y, 0u); OUT.st(y, 0u); }
    if (threadIdx.x == 0) { s_first = 0x7FFFFFFF; s_cnt = 0; }
    __syncthreads();
    int best = 0, blo = 1, bhi = 0;          // rows of the band the best component so far occupies in OUT
    int ys = ylo;                            // first row of the band that may still hold a pixel
    for (;;) {
        for (; ys <= yhi; ++ys)
            if (__any_sync(FULL, P0.ld(ys) != 0u)) break;
        if (lane == 0 && ys <= yhi) atomicMin(&s_first, ys);
        __syncthreads();
        const int y0 = s_first;
        if (y0 == 0x7FFFFFFF) break;
        int alo = 0x7FFFFFFF, ahi = -1;
        if (y0 >= ylo && y0 <= yhi) {        // the band that owns the row plants the seed: the row's first pixel, flooded
            const uint32_t w = P0.ld(y0);
            const int sl = __ffs(__ballot_sync(FULL, w != 0u)) - 1;
            const uint32_t seed = (lane == sl) ? (w & (0u - w)) : 0u;
            P1.st(y0, flood_row(w, seed, lane));
            alo = ahi = y0;
        }
        __syncthreads();
        if (threadIdx.x == 0) s_first = 0x7FFFFFFF;
        reach_component(P0, P1, H, ylo, yhi, lane, alo, ahi);
        int cnt = 0;
        for (int y = alo; y <= ahi; ++y) {
            const uint32_t R = P1.ld(y);
            cnt += __popc(R);
            P0.st(y, P0.ld(y) & ~R);
        }
        for (int o = 16; o > 0; o >>= 1) cnt += __shfl_xor_sync(FULL, cnt, o);
        if (lane == 0 && cnt) atomicAdd(&s_cnt, cnt);
        __syncthreads();
        const int total = s_cnt;
        if (total > best) {
            for (int y = blo; y <= bhi; ++y) OUT.st(y, 0u);
            for (int y = alo; y <= ahi; ++y) OUT.st(y, P1.ld(y));
            best = total; blo = alo; bhi = ahi;
        }
        for (int y = alo; y <= ahi; ++y) P1.st(y, 0u);
        __syncthreads();
        if (threadIdx.x == 0) s_cnt = 0;
    }
    if (best == 0)                                             // no foreground: label 0 wins the argmax -> all True
        for (int y = ylo; y <= yhi; ++y) OUT.st(y, rowmask);
}

}  // namespace ssad

// ------------------------------------------------------------------ C ABI
namespace {
struct ObjMaskLayout {
    int nw, ntx, nty;
    long long plane_words;
    size_t o_flags, o_list, o_count, total;
};
ObjMaskLayout obj_mask_layout(int batch, int height, int width) {
    ObjMaskLayout L;
    L.nw = (width + 31) / 32;
    L.ntx = (width + ssad::CT_W - 1) / ssad::CT_W;
    L.nty = (height + ssad::CT_H - 1) / ssad::CT_H;
    L.plane_words = (long long)height * L.nw;
    auto up = [](size_t v) { return (v + 255) / 256 * 256; };
    const size_t ntile = (size_t)batch * L.ntx * L.nty;
    L.o_flags = up((size_t)batch * 3 * (size_t)L.plane_words * sizeof(uint32_t));
    L.o_list = L.o_flags + up(ntile);
    L.o_count = L.o_list + up(ntile * sizeof(uint32_t));
    L.total = L.o_count + 256;
    return L;
}
}  // namespace

extern "C" size_t ssad_obj_mask_workspace(int32_t batch, int32_t height, int32_t width) {
    if (batch <= 0 || height <= 0 || width <= 0) return 0;
    return obj_mask_layout(batch, height, width).total;
}

// kernels one ssad_obj_mask call launches (bench.py's gpu_launches): screen, tiles, float64 canny, bit planes
extern "C" int ssad_obj_mask_launches(void) { return 4; }

extern "C" int ssad_obj_mask(int32_t batch, int32_t height, int32_t width, const uint8_t* src, int64_t src_pitch,
                             int64_t src_stride, const SsadCannyConsts* consts, uint32_t* mask_bits, int32_t mask_pitch_words,
                             int64_t mask_stride_words, void* workspace, size_t workspace_bytes, void* stream) {
    if (batch < 0 || height <= 0 || width <= 0 || !src || !mask_bits || !consts) return SSAD_EINVAL;
    if (src_pitch < 3ll * width || mask_pitch_words < (width + 31) / 32) return SSAD_EINVAL;
    if (batch == 0) return SSAD_OK;
    if (width > 1024 || width < 13 || height < 13) return SSAD_EUNSUPPORTED;   // one lane per row word; 6-pixel border classes
    if (batch > 65535) return SSAD_EUNSUPPORTED;
    const ObjMaskLayout L = obj_mask_layout(batch, height, width);
    if ((unsigned long long)batch * L.ntx * L.nty > 0xFFFFFFFFull) return SSAD_EUNSUPPORTED;
    if (!workspace || workspace_bytes < L.total) return SSAD_EWORKSPACE;
    if (reinterpret_cast<uintptr_t>(workspace) & 15) return SSAD_EALIGN;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    ssad::CannyK K;
    for (int i = 0; i < 13; ++i) K.w[i] = consts->gauss_w[i];
    K.low = consts->low; K.high = consts->high; K.eps = consts->eps;
    ssad::ScreenK SK;
    for (int i = 0; i < 7; ++i) SK.w[i] = (float)consts->gauss_w[i];
    {
        const float lo = (float)(consts->low * 255.0) - ssad::kScreenMargin, hi = (float)(consts->high * 255.0) - ssad::kScreenMargin;
        SK.lo2 = lo > 0.0f ? lo * lo : 0.0f;
        SK.hi2 = hi > 0.0f ? hi * hi : 0.0f;
    }
    const int nw = L.nw;
    const long long plane_words = L.plane_words;
    uint8_t* const wsb = static_cast<uint8_t*>(workspace);
    uint32_t* ws = static_cast<uint32_t*>(workspace);
    uint8_t* const flags = wsb + L.o_flags;
    uint32_t* const list = reinterpret_cast<uint32_t*>(wsb + L.o_list);
    uint32_t* const count = reinterpret_cast<uint32_t*>(wsb + L.o_count);
    static int sm_count[64];                       // SMs per device, read once
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return ssad_set_cuda_error(e);
    if (dev < 0 || dev >= 64) return SSAD_EUNSUPPORTED;
    if (sm_count[dev] == 0) {
        int n = 0;
        e = cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
        if (e != cudaSuccess) return ssad_set_cuda_error(e);
        e = cudaFuncSetAttribute(ssad::canny_nms_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(ssad::CannySmem));
        if (e == cudaSuccess)
            e = cudaFuncSetAttribute(ssad::canny_screen_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(ssad::ScreenSmem));
        if (e != cudaSuccess) return ssad_set_cuda_error(e);
        sm_count[dev] = n > 0 ? n : 1;
    }
    e = cudaMemsetAsync(count, 0, sizeof(uint32_t), st);
    if (e != cudaSuccess) return ssad_set_cuda_error(e);
    // 1. float32 screen: which 64x32 tiles can hold a pixel at the low / the high threshold (also zeroes both planes)
    dim3 sgrid((width + ssad::SC_W - 1) / ssad::SC_W, (height + ssad::SC_H - 1) / ssad::SC_H, batch);
    ssad::canny_screen_kernel<<<sgrid, ssad::SC_NT, sizeof(ssad::ScreenSmem), st>>>(SK, height, width, src, src_pitch, src_stride, flags,
                                                                                L.ntx, L.nty, ws, ws + plane_words, nw, 3 * plane_words);
    // 2. tiles connected to a high tile -> work list
    ssad::canny_tiles_kernel<<<batch, 32, (size_t)(2 * (L.nty + 2)) * sizeof(uint32_t), st>>>(L.ntx, L.nty, flags, list, count);
    // 3. float64 canny + non-maximum suppression on the listed tiles
    ssad::canny_nms_kernel<<<sm_count[dev] * SSAD_CANNY_MINB, ssad::CT_NT, sizeof(ssad::CannySmem), st>>>(
        K, height, width, src, src_pitch, src_stride, ws, ws + plane_words, nw, 3 * plane_words, list, count, L.ntx, L.nty);
    e = cudaGetLastError();
    if (e != cudaSuccess) return ssad_set_cuda_error(e);
    // 4. hysteresis, morphology, hole filling, largest component on the bit planes: a CTA per image, a band of rows per warp
    int nwarp = height / 64;
    nwarp = nwarp < 1 ? 1 : (nwarp > SSAD_PLANES_WARPS ? SSAD_PLANES_WARPS : nwarp);
    ssad::mask_planes_kernel<<<batch, 32 * nwarp, 0, st>>>(batch, height, width, nw, ws, plane_words, mask_bits, mask_pitch_words,
                                                         mask_stride_words);
    e = cudaGetLastError();
    if (e != cudaSuccess) return ssad_set_cuda_error(e);
    return SSAD_OK;
}
