// obj_mask (reference dataset_generator.py:27-39) for a batch, on device:
//
//   convert('L') -> skimage.feature.canny(sigma=1.5, low=5, high=15) -> binary_dilation(3x3) -> binary_closing(3x3)
//   -> binary_fill_holes(3x3) -> binary_erosion(4x4) -> largest 8-connected component (degenerate: no foreground ->
//   everything), bit-packed.
//
// Four kernels (+ ssad_obj_mask_static at the end of the file: the same call with its scratch in a static device array).
//   canny32_kernel     float32, one CTA per 128x32 tile, four CTAs per SM: the whole canny chain in fast arithmetic
//                      (L -> Gaussian axis 0 -> axis 1 / bleed-over -> Sobel, |gradient|^2 -> candidates listed ->
//                      thresholds + non-maximum suppression one candidate per thread); a decision is taken only when
//                      it clears the proven error of the float32 values, the few undecided pixels (near ties) are
//                      listed -- or, when a 64x32 tile has too many, the tile.
//   canny_resolve_kernel  float64, one warp per listed pixel: the reference's operation sequence on the pixel's window.
//   canny_nms_kernel   float64, dense, the listed 64x32 tiles only (persistent CTAs walking the list): L -> separable
//                      Gaussian (scipy correlate1d order: centre tap, then the symmetric pairs from the outermost
//                      inwards, every product and sum rounded on its own) / bleed-over normalisation -> Sobel (reflect)
//                      -> magnitude -> bilinear non-maximum suppression -> two bit planes (>= low, >= high) written
//                      with warp ballots.  FP64-pipe bound: ~125 non-fusable double operations per pixel (no FMA: the
//                      CPU libraries round every op).
//   mask_planes_kernel one CTA per image, one band of 64 rows per warp, one lane per 32-pixel word of a row
//                      (W <= 1024): every later step is a reachability or a shift on bit planes.  Connectivity
//                      (hysteresis, hole filling) = down / up sweeps inside a band to the band's fixed point, then
//                      rounds over the band boundaries; one row step floods a whole row at once with the carry trick
//                      reach = (M & ~(M + S)) | S  inside a word and a ballot + add carry chain across the 32 words.
//                      Dilation + closing + hole-fill seeds are one streaming sweep; the largest component is a
//                      union-find over pixel runs in shared memory (flood peel beyond 4096 runs).
#include <cuda_runtime.h>
#include <mutex>
#include <cstdio>
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
#define SSAD_PLANES_WARPS 16
#endif
// ... and the fewest rows a band should have
#ifndef SSAD_PLANES_ROWS
#define SSAD_PLANES_ROWS 64
#endif
// resident CTAs per SM the plane kernel is compiled for (3 x 512 threads = 40 registers per thread)
#ifndef SSAD_PLANES_MINB
#define SSAD_PLANES_MINB 3
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

// ------------------------------------------------------------------ canny: float32 decisions
// The float64 chain costs ~125 unfusable double operations per pixel, and every decision canny makes with it is a
// comparison: |gradient| against the two thresholds, the signs and the order of the two gradient components, and the
// two interpolated neighbours against the pixel (non-maximum suppression).  This kernel evaluates the same real-number
// chain (L -> 13-tap Gaussian along axis 0, along axis 1, / bleed-over -> Sobel (reflect) -> |gradient| -> NMS) in
// float32 with fused multiply-adds, in grey-level units (x 255), and takes a decision only when it clears the proven
// error of the float32 values; the pixels that stay undecided (near ties, ~0.1 % of an image) are listed for
// canny_resolve_kernel, which repeats the reference's float64 operation sequence for exactly those pixels.  A tile
// with too many undecided pixels goes to the dense float64 kernel above instead (canny_nms_kernel, work list).
//
// Error bounds (grey levels; u = 2^-24; L is an exact integer <= 255; the taps are positive and sum to 1):
//   axis-0 pass   7 rounded operations on partial sums <= 255 (the pair sums v[j] + v[12-j] are exact integers) and the
//                 rounding of the taps:                         |V32 - V|  <= 7u 255 + u 255            = 1.22e-4
//   axis-1 pass   8 rounded operations + the propagated error:  |S32 - S|  <= 8u 255 + u 255 + 1.22e-4 = 2.6e-4
//   bleed-over    (tiles within 6 pixels of the image border) the divisor A(y) B(x) >= 0.4, its float32 reciprocal is
//                 within 2.2e-6 relative:                       |S32 - S|  <= 2.5 * 2.6e-4 + 255 * 2.2e-6 = 1.2e-3
//   Sobel         four differences of two S, weights (1, 2, 1), three roundings on sums <= 1020:
//                 |is32 - is| <= 8 |S32 - S| + 1.8e-4           = 2.3e-3 (interior), 9.8e-3 (border)
//   |gradient|    |m32 - m| <= sqrt(2) |is32 - is| + 3u 1443    = 3.5e-3 (interior), 1.41e-2 (border)
// The float64 chain of the reference is within 1e-11 of the real-number values.  The constants below round these up:
//   interior  Ec = 0.003, Em = 0.0045;   border  Ec = 0.0125, Em = 0.018.
// Decisions (dataset_generator.py:30 -> skimage canny): m >= low, m >= high: certain when m32 is Em away from the
// threshold; sign(is), sign(js): certain when |.| > Ec; |is| > |js|: certain when they differ by more than 2 Ec; the
// interpolated neighbour v = (n12 num + n11 (den - num)) / den against m: compare den v with den m; the float32
// difference is within  2 Em den + Ec (n12 + 2 n11 + m)  (+ 1e-3 second order + 1e-6 (lhs + rhs) evaluation rounding)
// of the true one, so it is certain when it exceeds  T = 2 Em den + 2 Ec (n11 + n12 + m) + 1e-3 + 1e-6 (lhs + rhs).
#ifndef SSAD_C32_NT
#define SSAD_C32_NT 256
#endif
#ifndef SSAD_C32_H
#define SSAD_C32_H 32
#endif
#ifndef SSAD_C32_MINB
#define SSAD_C32_MINB 4
#endif
constexpr int SC_W = 128, SC_H = SSAD_C32_H, SC_NT = SSAD_C32_NT;     // SC_H: a multiple of the 32-row dense tiles
static_assert(SC_H % 32 == 0, "whole 64x32 tiles");
constexpr int SG_W = SC_W + 16, SG_H = SC_H + 16;       // L: rows y0-8 .. y0+SC_H+7, columns x0-8 .. x0+SC_W+7
constexpr int SV_H = SC_H + 4;                           // V (axis-0 pass): rows y0-2 .. y0+SC_H+1, the columns of L
constexpr int SS_W = SC_W + 4;                           // S (smoothed): rows of V, columns x0-2 .. x0+SC_W+1
constexpr int SM_H = SC_H + 2, SM_W = SS_W;              // |gradient|^2: rows y0-1 .. y0+SC_H, columns x0-1 .. (130 used)
constexpr int SUB_CAP = 96;                              // undecided pixels a 64x32 tile may list (else: dense float64 tile)
static_assert(SG_W % 4 == 0 && SS_W % 4 == 0, "vector rows");
constexpr int SV_P = 2 * SG_W + 4;                       // floats per row PAIR of the axis-0 result: an odd number of 16-byte
                                                         // chunks, so lanes on consecutive row pairs hit different banks
static_assert(SM_H * SM_W <= (SV_H / 2) * SV_P, "|gradient|^2 reuses the axis-0 buffer");

struct C32K {
    float w[7];              // taps 0..6 (centre last) of the symmetric 13-tap kernel
    float ec[2], em[2];      // error bounds of a gradient component / of |gradient|: [interior, border]
    float lo_no[2], lo_yes[2], hi_no[2], hi_yes[2];   // 255 low -+ em, 255 high -+ em (lower ones clamped at 0)
    float lo_no2[2];         // lo_no^2, a hair below (the squares' own rounding)
    int sub_cap;             // <= SUB_CAP
    unsigned pix_cap;        // entries of the undecided-pixel list
};
struct C32Smem {
    float G[SG_H * SG_W];    // L (grey levels); reused for S once the axis-0 pass is done
    float V[(SV_H / 2) * SV_P];   // axis-0 pass, rows interleaved in pairs; reused for |gradient|^2
    float ra[SV_H], rb[SS_W];   // 1 / bleed-over along the rows / the columns (border tiles)
    uint32_t cand[SC_H * (SC_W / 32)];   // pixels whose |gradient| may reach the low threshold, one word per 32 pixels of a row
    uint32_t ncand;                      // ... how many: they are listed (i << 8 | j) behind S in the G buffer
    uint32_t lo_w[SC_H * (SC_W / 32)], hi_w[SC_H * (SC_W / 32)];   // the tile's words of the two planes
    uint32_t sub_cnt[SC_H / 16];
    uint16_t sub_list[SC_H / 16][SUB_CAP];
};
static_assert((SV_H * SS_W + 2) * 4 + SC_H * SC_W * 2 <= SG_H * SG_W * 4, "S and the candidate list reuse the L buffer");

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

__global__ void __launch_bounds__(SC_NT, SSAD_C32_MINB)
canny32_kernel(const __grid_constant__ C32K K, int H, int W, const uint8_t* __restrict__ src, long long src_pitch,
               long long src_stride, uint32_t* __restrict__ low_plane, uint32_t* __restrict__ high_plane, int nw,
               long long plane_stride, uint2* __restrict__ pix_list, uint32_t* __restrict__ pix_count,
               uint32_t* __restrict__ tile_list, uint32_t* __restrict__ tile_count, int ntx, int nty) {
    extern __shared__ __align__(16) uint8_t smem_raw[];
    C32Smem& sm = *reinterpret_cast<C32Smem*>(smem_raw);
    const int tid = threadIdx.x;
    const int x0 = blockIdx.x * SC_W, y0 = blockIdx.y * SC_H, b = blockIdx.z;
    const uint8_t* img = src + (size_t)b * src_stride;
    const bool border = (x0 - 8 < 6) || (y0 - 8 < 6) || (x0 + SC_W + 8 > W - 6) || (y0 + SC_H + 8 > H - 6);

    if (tid < SC_H / 16) sm.sub_cnt[tid] = 0u;
    if (tid == 32) sm.ncand = 0u;
    if (tid < SC_H * (SC_W / 32)) { sm.cand[tid] = 0u; sm.lo_w[tid] = 0u; sm.hi_w[tid] = 0u; }
    static_assert(SC_H * (SC_W / 32) <= SC_NT, "one thread per candidate word");
    if (border) {
        // gaussian_filter(ones): the taps that fall inside the image, per row and per column of the tile
        for (int i = tid; i < SV_H + SS_W; i += SC_NT) {
            const bool is_row = i < SV_H;
            const int p = is_row ? y0 - 2 + i : x0 - 2 + (i - SV_H), n = is_row ? H : W;
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
        // a thread keeps one group of four columns and takes every RS-th row: one division per thread, pointers and
        // shared-memory offsets advance by constants
        constexpr int NG = SG_W / 4, RS = SC_NT / NG, NK = (SG_H + RS - 1) / RS;
        static_assert(RS >= 1, "a row of groups per pass");
        const int rr = tid / NG, g = tid - rr * NG;
        const bool active = rr < RS;                       // (the last SC_NT - RS * NG threads repeat rows of others, unstored)
        uint32_t w[NK][3];
        const bool inside = vec_ok && x0 >= 8 && y0 >= 8 && x0 + SC_W + 8 <= W && y0 + SC_H + 8 <= H;   // (uniform)
        if (inside) {
            // every pixel of the tile and its halo is in the image: no bounds
            const uint8_t* const base = img + (size_t)(y0 - 8) * src_pitch + (size_t)(x0 - 8) * 3 + g * 12;
            const uint8_t* row = base + (size_t)rr * src_pitch;             // (one 64-bit product, then constant steps)
            const size_t down = (size_t)RS * src_pitch;
            #pragma unroll
            for (int k = 0; k < NK; ++k) {
                // (rows past the halo -- last round only -- re-read the last row; they are not stored)
                const uint8_t* rq = (k * RS + RS - 1 < SG_H || rr + k * RS < SG_H) ? row : base + (size_t)(SG_H - 1) * src_pitch;
                const uint32_t* q = reinterpret_cast<const uint32_t*>(rq);
                w[k][0] = __ldg(q); w[k][1] = __ldg(q + 1); w[k][2] = __ldg(q + 2);
                row += down;
            }
        } else {
            const int x = x0 - 8 + 4 * g;
            const bool xany = x + 3 >= 0 && x < W, xfull = vec_ok && x >= 0 && x + 3 < W;
            #pragma unroll
            for (int k = 0; k < NK; ++k) {
                const int r = rr + k * RS, y = y0 - 8 + r;
                w[k][0] = w[k][1] = w[k][2] = 0u;
                if (active && r < SG_H && y >= 0 && y < H && xany) {
                    if (xfull) {
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
        for (int k = 0; k < NK; ++k) {
            const int r = rr + k * RS;
            if (active && r < SG_H) *reinterpret_cast<float4*>(&sm.G[(r * NG + g) * 4]) = luma4(w[k][0], w[k][1], w[k][2]);     // (an all-zero group gives L = 0)
        }
    }
    __syncthreads();

    // ---- Gaussian along axis 0: a thread slides down 12 rows of a column pair (the last segment has 8).  The result is
    // stored with the rows interleaved in pairs -- V2[r / 2][c] = (V[r][c], V[r + 1][c]) -- so that the axis-1 pass
    // reads ready-made packed pairs (two rows per FFMA2)
    om_f32x2 wp[7];
    #pragma unroll
    for (int j = 0; j < 7; ++j) wp[j] = om_pack2(K.w[j], K.w[j]);
    {
        constexpr int NP = SG_W / 2, SEG = 12, NSEG = (SV_H + SEG - 1) / SEG;
        static_assert(SEG % 2 == 0 && SV_H % 2 == 0, "row pairs");
        for (int t = tid; t < NP * NSEG; t += SC_NT) {
            const int seg = t / NP, pr = t - seg * NP;
            const int r0 = seg * SEG, nrow = min(SEG, SV_H - r0);
            const float* g = &sm.G[r0 * SG_W + 2 * pr];
            om_f32x2 v[SEG + 12];
            #pragma unroll
            for (int k = 0; k < SEG + 12; ++k) v[k] = (k < nrow + 12) ? *reinterpret_cast<const om_f32x2*>(g + k * SG_W) : 0ull;
            #pragma unroll
            for (int o = 0; o < SEG; o += 2) {
                if (o < nrow) {
                    float a0, a1, b0, b1;
                    om_unpack2(om_sym13(v + o, wp), a0, a1);          // row r0 + o:     columns 2 pr, 2 pr + 1
                    om_unpack2(om_sym13(v + o + 1, wp), b0, b1);      // row r0 + o + 1
                    *reinterpret_cast<float4*>(&sm.V[((r0 + o) >> 1) * SV_P + 4 * pr]) = make_float4(a0, b0, a1, b1);
                }
            }
        }
    }
    __syncthreads();

    // ---- Gaussian along axis 1 (/ bleed-over in border tiles): a thread takes twelve columns of a row pair -- 24 packed
    // pairs = 12 LDS.128 for 12 packed results (two columns per thread cost 7 loads for 2).  Lanes walk down the row
    // pairs: SV_P makes the eight 16-byte accesses of a quarter warp fall on different banks.
    float* const S = sm.G;
    {
        constexpr int SEG1 = 12, NS1 = SS_W / SEG1, NRP = SV_H / 2;     // 11 column segments x 18 row pairs = 198 tasks
        static_assert(SS_W % SEG1 == 0 && SEG1 % 4 == 0 && (NS1 - 1) * SEG1 + SEG1 + 12 <= SG_W, "whole segments inside a row of V");
        for (int t = tid; t < NRP * NS1; t += SC_NT) {
            const int sg = t / NRP, rp = t - sg * NRP, c0 = sg * SEG1;
            // S[r][c] = sum_k w[k] V[r][c + k], c = c0 .. c0 + 11: the pairs V2[rp][c0 .. c0 + 23]
            const float4* vp = reinterpret_cast<const float4*>(&sm.V[rp * SV_P + 2 * c0]);
            om_f32x2 v[SEG1 + 12];
            #pragma unroll
            for (int k = 0; k < (SEG1 + 12) / 2; ++k) {
                const float4 f = vp[k];
                v[2 * k] = om_pack2(f.x, f.y); v[2 * k + 1] = om_pack2(f.z, f.w);
            }
            float s0[SEG1], s1[SEG1];                     // rows 2 rp, 2 rp + 1
            #pragma unroll
            for (int j = 0; j < SEG1; ++j) om_unpack2(om_sym13(v + j, wp), s0[j], s1[j]);
            const int r = 2 * rp;
            if (border) {
                const float ra0 = sm.ra[r], ra1 = sm.ra[r + 1];
                #pragma unroll
                for (int j = 0; j < SEG1; ++j) { const float cb = sm.rb[c0 + j]; s0[j] *= ra0 * cb; s1[j] *= ra1 * cb; }
            }
            #pragma unroll
            for (int j = 0; j < SEG1; j += 4) {
                *reinterpret_cast<float4*>(&S[r * SS_W + c0 + j]) = make_float4(s0[j], s0[j + 1], s0[j + 2], s0[j + 3]);
                *reinterpret_cast<float4*>(&S[(r + 1) * SS_W + c0 + j]) = make_float4(s1[j], s1[j + 1], s1[j + 2], s1[j + 3]);
            }
        }
    }
    __syncthreads();
    if (border) {
        // ndi.sobel mode='reflect': a neighbour outside the image is the pixel itself clamped to the image
        // (S row r <-> image row y0 - 2 + r, column c <-> x0 - 2 + c; only the rows / columns -1 and H / W are read)
        for (int i = tid; i < SS_W; i += SC_NT) {
            if (y0 == 0) S[1 * SS_W + i] = S[2 * SS_W + i];
            const int rl = H - (y0 - 2);                       // row of image row H
            if (rl >= 1 && rl < SV_H) S[rl * SS_W + i] = S[(rl - 1) * SS_W + i];
        }
        __syncthreads();
        for (int i = tid; i < SV_H; i += SC_NT) {
            if (x0 == 0) S[i * SS_W + 1] = S[i * SS_W + 2];
            const int cl = W - (x0 - 2);
            if (cl >= 1 && cl < SS_W) S[i * SS_W + cl] = S[i * SS_W + cl - 1];
        }
        __syncthreads();
    }

    // ---- Sobel, |gradient|^2 for the tile and a 1-pixel halo: a thread takes four pixels of a row; the pixels that may
    // reach the low threshold (a few per cent of an image) are marked in the candidate words
    float* const M2 = sm.V;          // M2[r][c] <-> pixel (y0 - 1 + r, x0 - 1 + c) from S rows r .. r+2, columns c .. c+2
    const int bc = border ? 1 : 0;
    const float Ec = K.ec[bc], Em = K.em[bc];
    const float lo_no = K.lo_no[bc], lo_yes = K.lo_yes[bc], hi_no = K.hi_no[bc], hi_yes = K.hi_yes[bc], lo_no2 = K.lo_no2[bc];
    {
        // a thread slides down five rows of a group of four columns: every row of S is loaded once (6 floats) and its
        // horizontal differences computed once for the three result rows that use it
        constexpr int NG = SM_W / 4, SEG5 = 5, NS5 = (SM_H + SEG5 - 1) / SEG5;       // 33 x 7 = 231 tasks
        static_assert(NG * NS5 <= SC_NT, "one round");
        const om_f32x2 NEG1 = om_pack2(-1.0f, -1.0f), TWO = om_pack2(2.0f, 2.0f);
        for (int t = tid; t < NG * NS5; t += SC_NT) {
            const int sg = t / NG, g = t - sg * NG;
            const int r0 = sg * SEG5, nr = min(SEG5, SM_H - r0);
            om_f32x2 spa[3], spb[3], ea[2], eb[2];          // rows r, r + 1 of S: column pairs, right - left of the pixel pairs
            #pragma unroll
            for (int k = 0; k < SEG5 + 2; ++k) {
                if (k < nr + 2) {
                    // (the last group reads two floats past the row: finite values of the next row / the V buffer, for
                    //  the unused columns 130, 131)
                    const float4 a = *reinterpret_cast<const float4*>(&S[(r0 + k) * SS_W + 4 * g]);
                    const float2 c = *reinterpret_cast<const float2*>(&S[(r0 + k) * SS_W + 4 * g + 4]);
                    om_f32x2 sp[3], e[2];
                    sp[0] = om_pack2(a.x, a.y); sp[1] = om_pack2(a.z, a.w); sp[2] = om_pack2(c.x, c.y);
                    e[0] = om_fma2(sp[0], NEG1, sp[1]); e[1] = om_fma2(sp[1], NEG1, sp[2]);
                    if (k >= 2) {
                        const int r = r0 + k - 2;           // result row: S rows r (a), r + 1 (b), r + 2 (this one)
                        float dc[6], o[4];
                        #pragma unroll
                        for (int p2 = 0; p2 < 3; ++p2) om_unpack2(om_fma2(spa[p2], NEG1, sp[p2]), dc[2 * p2], dc[2 * p2 + 1]);   // down - up
                        #pragma unroll
                        for (int p2 = 0; p2 < 2; ++p2) {
                            const om_f32x2 js2 = om_add2(om_fma2(TWO, eb[p2], ea[p2]), e[p2]);
                            const float is0 = fmaf(2.0f, dc[2 * p2 + 1], dc[2 * p2]) + dc[2 * p2 + 2];
                            const float is1 = fmaf(2.0f, dc[2 * p2 + 2], dc[2 * p2 + 1]) + dc[2 * p2 + 3];
                            const om_f32x2 is2 = om_pack2(is0, is1);
                            om_unpack2(om_fma2(is2, is2, om_mul2(js2, js2)), o[2 * p2], o[2 * p2 + 1]);
                        }
                        *reinterpret_cast<float4*>(&M2[r * SM_W + 4 * g]) = make_float4(o[0], o[1], o[2], o[3]);
                        if (fmaxf(fmaxf(o[0], o[1]), fmaxf(o[2], o[3])) >= lo_no2 && r >= 1 && r <= SC_H) {
                            // the four pixels j = 4g - 1 .. 4g + 2 of tile row i as bits of one 64-bit window (they may straddle a word)
                            const int i = r - 1, y = y0 + i, j0 = 4 * g - 1;
                            uint32_t nib = (o[0] >= lo_no2 ? 1u : 0u) | (o[1] >= lo_no2 ? 2u : 0u) | (o[2] >= lo_no2 ? 4u : 0u) | (o[3] >= lo_no2 ? 8u : 0u);
                            if (j0 < 0) nib &= ~1u;                         // column -1 is halo
                            if (j0 + 3 >= SC_W) nib &= (1u << (SC_W - j0)) - 1u;     // columns >= SC_W are halo
                            if (border) {                                   // eroded_mask: the 1-pixel border of the image never qualifies
                                if (y < 1 || y > H - 2) nib = 0u;
                                #pragma unroll
                                for (int q = 0; q < 4; ++q) { const int x = x0 + j0 + q; if (x < 1 || x > W - 2) nib &= ~(1u << q); }
                            }
                            if (nib) {
                                const int jb = max(j0, 0);                   // first valid column of the window
                                const uint32_t bits = (j0 < 0) ? (nib >> 1) : nib;
                                const int w0 = jb >> 5, sh = jb & 31;
                                atomicOr(&sm.cand[i * (SC_W / 32) + w0], bits << sh);
                                if (sh > 28 && (bits >> (32 - sh))) atomicOr(&sm.cand[i * (SC_W / 32) + w0 + 1], bits >> (32 - sh));
                            }
                        }
                    }
                    #pragma unroll
                    for (int p2 = 0; p2 < 3; ++p2) { spa[p2] = spb[p2]; spb[p2] = sp[p2]; }
                    #pragma unroll
                    for (int p2 = 0; p2 < 2; ++p2) { ea[p2] = eb[p2]; eb[p2] = e[p2]; }
                }
            }
        }
    }
    __syncthreads();

    // ---- the candidates, listed (any order): one thread per pixel below instead of one warp per 32-pixel word, of which
    // only the few candidate lanes would work
    uint16_t* const clist = reinterpret_cast<uint16_t*>(sm.G + SV_H * SS_W + 2);     // (S's last row is read 2 floats past its end)
    if (tid < SC_H * (SC_W / 32)) {
        uint32_t cw = sm.cand[tid];
        if (cw) {
            uint32_t pos = atomicAdd(&sm.ncand, (uint32_t)__popc(cw));
            const uint32_t ij = (uint32_t)((tid >> 2) << 8) | (uint32_t)((tid & 3) * 32);
            for (; cw; cw &= cw - 1) clist[pos++] = (uint16_t)(ij + (__ffs(cw) - 1));
        }
    }
    __syncthreads();
    // ---- thresholds and non-maximum suppression of the candidates, decided in float32 where the error bounds allow
    {
        const int ncand = (int)sm.ncand;
        for (int k = tid; k < ncand; k += SC_NT) {
            const int i = clist[k] >> 8, j = clist[k] & 255;
            bool lo = false, hi = false, undecided = false;
            {
                const float m = __fsqrt_rn(M2[(i + 1) * SM_W + (j + 1)]);       // (the thresholds want the correctly rounded root)
                if (m >= lo_no) {
                    undecided = m < lo_yes;
                    // the gradient again (S rows i+1 .. i+3 around row i+2 <-> y, columns j+1 .. j+3 around j+2 <-> x)
                    const float* sp = &S[(i + 1) * SS_W + (j + 1)];
                    const float s_ul = sp[0], s_uc = sp[1], s_ur = sp[2];
                    const float s_cl = sp[SS_W], s_cr = sp[SS_W + 2];
                    const float s_dl = sp[2 * SS_W], s_dc = sp[2 * SS_W + 1], s_dr = sp[2 * SS_W + 2];
                    const float is = fmaf(2.0f, s_dc - s_uc, s_dl - s_ul) + (s_dr - s_ur);
                    const float js = fmaf(2.0f, s_cr - s_cl, s_ur - s_ul) + (s_dr - s_dl);
                    const float ai = fabsf(is), aj = fabsf(js);
                    bool keep = false;
                    if (!(ai > Ec) || !(aj > Ec) || !(fabsf(ai - aj) > 2.0f * Ec)) {
                        undecided = true;
                    } else {
                        // neighbours along the gradient: (dr1, dc1) straight, (dr2, dc2) diagonal, and their mirrors
                        const bool up = is > 0.0f, right = js > 0.0f, steep = ai > aj;
                        const bool c1 = up == right;             // cond1; else cond2
                        const int dr1 = steep ? (c1 ? 1 : -1) : 0, dc1 = steep ? 0 : 1;
                        const int dr2 = c1 ? 1 : -1, dc2 = 1;
                        const float num = steep ? aj : ai, den = steep ? ai : aj;
                        const float* mp = &M2[(i + 1) * SM_W + (j + 1)];
                        const float n11 = __fsqrt_rn(mp[dr1 * SM_W + dc1]), n12 = __fsqrt_rn(mp[dr2 * SM_W + dc2]);
                        const float n21 = __fsqrt_rn(mp[-dr1 * SM_W - dc1]), n22 = __fsqrt_rn(mp[-dr2 * SM_W - dc2]);
                        const float dmn = den - num, rhs = m * den;
                        const float lp = fmaf(n12, num, n11 * dmn), lm = fmaf(n22, num, n21 * dmn);
                        const float dp = rhs - lp, dm = rhs - lm;
                        const float base = 2.0f * Em * den + 1e-3f;
                        const float tp = base + 2.0f * Ec * (n11 + n12 + m) + 1e-6f * (rhs + lp);
                        const float tm = base + 2.0f * Ec * (n21 + n22 + m) + 1e-6f * (rhs + lm);
                        if (!(fabsf(dp) > tp) || !(fabsf(dm) > tm)) undecided = true;
                        else keep = dp > 0.0f && dm > 0.0f;
                    }
                    if (!undecided && keep) {
                        lo = true;
                        if (m >= hi_yes) hi = true;
                        else if (m >= hi_no) { undecided = true; lo = false; }
                    }
                }
            }
            if (undecided) {
                const int sub = (i >> 5) * 2 + (j >> 6);
                const uint32_t pos = atomicAdd(&sm.sub_cnt[sub], 1u);
                if (pos < (uint32_t)K.sub_cap) sm.sub_list[sub][pos] = (uint16_t)((i << 8) | j);
            }
            if (lo) atomicOr(&sm.lo_w[i * (SC_W / 32) + (j >> 5)], 1u << (j & 31));
            if (hi) atomicOr(&sm.hi_w[i * (SC_W / 32) + (j >> 5)], 1u << (j & 31));
        }
    }
    __syncthreads();
    if (tid < SC_H * (SC_W / 32)) {                  // the tile's words of both planes, zero where nothing qualified
        const int xw = (x0 >> 5) + (tid & 3), y = y0 + (tid >> 2);
        if (y < H && xw < nw) {
            low_plane[(size_t)b * plane_stride + (size_t)y * nw + xw] = sm.lo_w[tid];
            high_plane[(size_t)b * plane_stride + (size_t)y * nw + xw] = sm.hi_w[tid];
        }
    }
    // ---- the undecided pixels of every 64x32 tile: listed for canny_resolve_kernel, or the whole tile to the dense
    // float64 kernel when there are too many (or the list is full)
    {
        const int warp = tid >> 5, lane = tid & 31;
        for (int sub = warp; sub < SC_H / 16; sub += SC_NT / 32) {
            const uint32_t n = sm.sub_cnt[sub];
            const int tx = 2 * blockIdx.x + (sub & 1), ty = (SC_H / 32) * blockIdx.y + (sub >> 1);
            if (n > 0 && tx < ntx && ty < nty) {
                bool dense = n > (uint32_t)K.sub_cap;
                uint32_t base = 0;
                if (!dense) {
                    if (lane == 0) base = atomicAdd(pix_count, n);
                    base = __shfl_sync(0xFFFFFFFFu, base, 0);
                    if (base + n > K.pix_cap) {
                        dense = true;                // the reservation does not fit: void the part of it inside the list
                        for (uint32_t k = base + lane; k < min(base + n, K.pix_cap); k += 32) pix_list[k] = make_uint2(0xFFFFFFFFu, 0u);
                    }
                }
                if (dense) {
                    if (lane == 0) tile_list[atomicAdd(tile_count, 1u)] = (uint32_t)((b * nty + ty) * ntx + tx);
                } else {
                    for (uint32_t k = lane; k < n; k += 32) {
                        const uint32_t c = sm.sub_list[sub][k];
                        pix_list[base + k] = make_uint2((uint32_t)b, ((uint32_t)(y0 + (int)(c >> 8)) << 16) | (uint32_t)(x0 + (int)(c & 255u)));
                    }
                }
            }
        }
    }
}

// ------------------------------------------------------------------ canny: float64 for single pixels
// The reference's float64 operation sequence (the same device functions as the dense kernel above) for one listed
// pixel per warp: L on the 17x17 window -> axis-0 pass on 5 rows -> smoothed 5x5 (/ bleed-over) -> |gradient| 3x3 ->
// thresholds and non-maximum suppression of the centre -> its bits are OR-ed into the two planes.
constexpr int RS_NT = 256, RS_WARPS = RS_NT / 32;
struct ResolveWarp {
    double L[17 * 17];       // L / 255, rows y-8 .. y+8, columns x-8 .. x+8; 0 outside the image
    double V[5 * 17];        // rows y-2 .. y+2
    double S[5 * 5];         // rows y-2 .. y+2, columns x-2 .. x+2 (in-image entries)
    double M[3 * 3];
};

__global__ void __launch_bounds__(RS_NT)
canny_resolve_kernel(const __grid_constant__ CannyK K, int H, int W, const uint8_t* __restrict__ src, long long src_pitch,
                     long long src_stride, uint32_t* __restrict__ low_plane, uint32_t* __restrict__ high_plane, int nw,
                     long long plane_stride, const uint2* __restrict__ pix_list, const uint32_t* __restrict__ pix_count,
                     unsigned pix_cap) {
    __shared__ double lut[256];
    __shared__ ResolveWarp wsm[RS_WARPS];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    lut[tid] = __ddiv_rn((double)tid, 255.0);
    __syncthreads();
    ResolveWarp& R = wsm[warp];
    const uint32_t n = min(*pix_count, pix_cap);
    for (uint32_t item = blockIdx.x * RS_WARPS + warp; item < n; item += gridDim.x * RS_WARPS) {
        const uint2 e = pix_list[item];
        if (e.x == 0xFFFFFFFFu) continue;                  // voided reservation (its tile went to the dense kernel)
        const int b = (int)e.x, y = (int)(e.y >> 16), x = (int)(e.y & 0xFFFFu);
        const uint8_t* img = src + (size_t)b * src_stride;
        __syncwarp();
        {
            // the 17 x 17 window: all byte loads of the lane issued before the first is used (their latency is the
            // kernel's top stall)
            constexpr int NR = (17 * 17 + 31) / 32;
            uint32_t px[NR];
            #pragma unroll
            for (int k = 0; k < NR; ++k) {
                const int i = lane + 32 * k, r = i / 17, c = i - r * 17;
                const int yy = y - 8 + r, xx = x - 8 + c;
                px[k] = 0xFFFFFFFFu;
                if (i < 17 * 17 && yy >= 0 && yy < H && xx >= 0 && xx < W) {
                    const uint8_t* q = img + (size_t)yy * src_pitch + (size_t)xx * 3;
                    px[k] = (uint32_t)__ldg(q) | ((uint32_t)__ldg(q + 1) << 8) | ((uint32_t)__ldg(q + 2) << 16);
                }
            }
            #pragma unroll
            for (int k = 0; k < NR; ++k) {
                const int i = lane + 32 * k;
                if (i < 17 * 17)
                    R.L[i] = px[k] == 0xFFFFFFFFu ? 0.0
                           : lut[(19595u * (px[k] & 255u) + 38470u * ((px[k] >> 8) & 255u) + 7471u * (px[k] >> 16) + 0x8000u) >> 16];
            }
        }
        __syncwarp();
        for (int i = lane; i < 5 * 17; i += 32) {
            const int r = i / 17, c = i - r * 17;
            double v[13];
            #pragma unroll
            for (int k = 0; k < 13; ++k) v[k] = R.L[(r + k) * 17 + c];
            R.V[i] = sym13(v, K.w);
        }
        __syncwarp();
        if (lane < 25) {
            const int r = lane / 5, c = lane - r * 5;
            const int yy = y - 2 + r, xx = x - 2 + c;
            double s = 0.0;
            if (yy >= 0 && yy < H && xx >= 0 && xx < W) {
                double v[13];
                #pragma unroll
                for (int k = 0; k < 13; ++k) v[k] = R.V[r * 17 + c + k];
                const double t = sym13(v, K.w);
                #pragma unroll
                for (int k = 0; k < 13; ++k) v[k] = (yy + k - 6 >= 0 && yy + k - 6 < H) ? 1.0 : 0.0;
                const double a = sym13(v, K.w);                  // gaussian_filter(ones): axis 0 ...
                #pragma unroll
                for (int k = 0; k < 13; ++k) v[k] = (xx + k - 6 >= 0 && xx + k - 6 < W) ? a : 0.0;
                s = __ddiv_rn(t, __dadd_rn(sym13(v, K.w), K.eps));   // ... then axis 1, + eps
            }
            R.S[lane] = s;
        }
        __syncwarp();
        // ndi.sobel (reflect) at image pixel (yy, xx) from the 5x5 window
        auto sobel = [&](int yy, int xx, double& is, double& js) {
            const int ru = max(yy - 1, 0) - (y - 2), rc = yy - (y - 2), rd = min(yy + 1, H - 1) - (y - 2);
            const int cl = max(xx - 1, 0) - (x - 2), cc = xx - (x - 2), cr = min(xx + 1, W - 1) - (x - 2);
            const double s_ul = R.S[ru * 5 + cl], s_uc = R.S[ru * 5 + cc], s_ur = R.S[ru * 5 + cr];
            const double s_cl = R.S[rc * 5 + cl], s_cr = R.S[rc * 5 + cr];
            const double s_dl = R.S[rd * 5 + cl], s_dc = R.S[rd * 5 + cc], s_dr = R.S[rd * 5 + cr];
            const double dl = __dsub_rn(s_dl, s_ul), dc = __dsub_rn(s_dc, s_uc), dr = __dsub_rn(s_dr, s_ur);
            is = __dadd_rn(__dmul_rn(dc, 2.0), __dadd_rn(dl, dr));
            const double eu = __dsub_rn(s_ur, s_ul), ec = __dsub_rn(s_cr, s_cl), ed = __dsub_rn(s_dr, s_dl);
            js = __dadd_rn(__dmul_rn(ec, 2.0), __dadd_rn(eu, ed));
        };
        if (lane < 9) {
            const int r = lane / 3, c = lane - r * 3;
            double is, js;
            sobel(y - 1 + r, x - 1 + c, is, js);
            R.M[lane] = __dsqrt_rn(__dadd_rn(__dmul_rn(is, is), __dmul_rn(js, js)));
        }
        __syncwarp();
        if (lane == 0) {
            const double m = R.M[4];
            bool lo = false, hi = false;
            if (m >= K.low) {
                double is, js;
                sobel(y, x, is, js);
                const bool is_down = is <= 0.0, is_up = is >= 0.0, is_left = js <= 0.0, is_right = js >= 0.0;
                const bool cond1 = (is_up && is_right) || (is_down && is_left);
                const bool cond2 = (is_down && is_right) || (is_up && is_left);
                const double ai = fabs(is), aj = fabs(js);
                auto nb = [&](int dr, int dc) { return R.M[(1 + dr) * 3 + (1 + dc)]; };
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
            const size_t wi = (size_t)b * plane_stride + (size_t)y * nw + (x >> 5);
            if (lo) atomicOr(&low_plane[wi], 1u << (x & 31));
            if (hi) atomicOr(&high_plane[wi], 1u << (x & 31));
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
// The rows of the image are cut into one band per warp of the CTA.  A warp brings its band to a fixed point on its own --
// a sweep down and a sweep up (one row step floods a whole row, see flood_row), then sweeps over just the rows the
// previous sweep changed until one changes nothing -- with the neighbouring bands' boundary rows as it read them; a
// round ends at a CTA barrier.  In the following rounds a band that
// is at its fixed point only looks at the two boundary rows again: unchanged -> nothing to do; changed -> changes can
// only run inwards from that edge, so the sweep stops at the first row it leaves as it was, and the band iterates again
// only if something did change.  Rounds repeat until no warp changed anything (then every boundary row read in that
// round was final).  R only grows, so reading a neighbour's row while it is being rewritten costs at most another round.
// Every thread of the CTA must call this.  Returns the number of rounds.
template <bool COMPLEMENT>
__device__ int reach(const Plane& Mp, const Plane& Rp, int H, int ylo, int yhi, uint32_t rowmask, int lane) {
    bool first = true, settled = false;
    uint32_t seen_top = 0u, seen_bot = 0u;            // the neighbours' boundary rows as last read
    // One sweep in direction dir (0 down, 1 up) from row `from` to the band's end, `prev` = the row before `from` in
    // sweep order.  It may stop at the first row it leaves unchanged once it is past row `need`: up to `need` a row's
    // predecessor may have changed since the band was last closed in this direction, beyond it only this sweep's own
    // changes run on.  [clo, chi] = the rows it changed (chi < 0: none).
    auto sweep = [&](int dir, int from, uint32_t prev, int need, int& clo, int& chi) {
        const int step = dir == 0 ? 1 : -1, n = dir == 0 ? yhi - from + 1 : from - ylo + 1;
        clo = 0x7FFFFFFF; chi = -1;
        if (n <= 0) return;
        int y = from;
        uint32_t Mn = Mp.ld(y), Rn = Rp.ld(y);
        #pragma unroll 1
        for (int i = 0; i < n; ++i, y += step) {
            const uint32_t M = COMPLEMENT ? (~Mn & rowmask) : Mn;
            const uint32_t R = Rn;
            if (i + 1 < n) { Mn = Mp.ld(y + step); Rn = Rp.ld(y + step); }      // next row is in flight during the step
            const uint32_t S = R | (M & spread3(prev, lane));
            uint32_t Rnew = R;
            bool rowch = false;
            // (the very first sweep floods every row that holds a seed at all: R has not been spread inside its rows yet)
            if (__any_sync(FULL, first ? S != 0u : S != R)) {
                Rnew = flood_row(M, S, lane);
                if (Rnew != R) Rp.st(y, Rnew);
                rowch = __any_sync(FULL, Rnew != R);
            }
            if (rowch) { clo = min(clo, y); chi = max(chi, y); }
            else if (dir == 0 ? y > need : y < need) break;
            prev = Rnew;
        }
    };
    for (int rounds = 1;; ++rounds) {
        bool changed = false;
        if (ylo <= yhi) {
            const uint32_t top = ylo > 0 ? Rp.ldcg(ylo - 1) : 0u, bot = yhi + 1 < H ? Rp.ldcg(yhi + 1) : 0u;
            const bool same = __all_sync(FULL, top == seen_top && bot == seen_bot);
            seen_top = top; seen_bot = bot;
            bool work = !settled;
            int clo, chi;
            if (settled && !same) {
                // the band is at its fixed point: changes can only run inwards from a boundary row that changed
                sweep(0, ylo, top, ylo - 1, clo, chi);
                work = chi >= 0;
                sweep(1, yhi, bot, yhi + 1, clo, chi);
                work = work || chi >= 0;
                changed = work;
            }
            if (work) {
                // to the band's own fixed point: one full sweep down and one up, then sweeps over the rows the previous
                // sweep changed (and as far beyond as their own changes run) until one changes nothing
                sweep(0, ylo, top, yhi, clo, chi);
                first = false;
                changed = changed || chi >= 0;
                sweep(1, yhi, bot, ylo, clo, chi);
                for (int dir = 0; chi >= 0; dir ^= 1) {
                    changed = true;
                    const int a = clo, b = chi;
                    if (dir == 0) sweep(0, a + 1, Rp.ld(a), b, clo, chi);
                    else sweep(1, b - 1, Rp.ld(b), a, clo, chi);
                }
                settled = true;
            }
        }
        if (!__syncthreads_or(changed ? 1 : 0)) return rounds;
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

// ---- largest component by runs (the usual case: a few runs per row after the hole fill and the 4x4 erosion) ----------
// A run = a maximal stretch of set pixels of one row; runs are numbered in raster order, so the smallest number of a
// component is the run that holds its first pixel (skimage's label order).  Rows are independent here: every warp takes
// rows round-robin, there is no sweep that has to walk the image top to bottom.
constexpr int CCL_MAXH = 2048;     // rows the row-offset table holds
constexpr int CCL_CAP = 4096;      // runs the union-find tables hold; more -> the caller falls back to the flood peel

struct CclSmem {
    int off[CCL_MAXH + 1];         // off[r] = number of runs in rows < r
    uint32_t parent[CCL_CAP];
    uint32_t size[CCL_CAP];
    int wsum[32];
    unsigned long long wbest[32];
};

__device__ __forceinline__ uint32_t le_mask(int q) { return FULL >> (31 - q); }

// exclusive prefix over the lanes of a warp
__device__ __forceinline__ int lane_prefix(int v, int lane) {
    int x = v;
    #pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int y = __shfl_up_sync(FULL, x, o);
        if (lane >= o) x += y;
    }
    return x - v;
}

__device__ __forceinline__ uint32_t ccl_find(volatile uint32_t* parent, uint32_t a) {
    uint32_t pa;
    while ((pa = parent[a]) != a) a = pa;
    return a;
}
// roots only ever get a smaller parent: no cycles, and the root of a component ends as its smallest run number
__device__ __forceinline__ void ccl_unite(uint32_t* parent, uint32_t a, uint32_t b) {
    for (;;) {
        a = ccl_find(parent, a);
        b = ccl_find(parent, b);
        if (a == b) return;
        if (a < b) { const uint32_t t = a; a = b; b = t; }
        if (atomicCAS(parent + a, a, b) == a) return;
    }
}

// OUT = the largest 8-connected component of P0 (ties: the first in raster order), all ones when P0 is empty.
// false (nothing written) when P0 holds more than `cap` runs.  Every thread of the CTA must call this.
__device__ bool largest_component_runs(const Plane& P0, const Plane& OUT, int H, uint32_t rowmask, int cap, CclSmem& S) {
    const int tid = threadIdx.x, NT = blockDim.x, lane = tid & 31, warp = tid >> 5, nwarp = NT >> 5;
    // 1. runs per row.  (The three row passes of this routine take the same 8-row groups per warp and issue the eight row
    //    loads together: a row is a 128-byte line that the first pass brings into this SM's L1.)
    for (int r0 = warp * 8; r0 < H; r0 += nwarp * 8) {
        uint32_t Bv[8];
        #pragma unroll
        for (int k = 0; k < 8; ++k) Bv[k] = (r0 + k < H) ? P0.ld(r0 + k) : 0u;
        #pragma unroll
        for (int k = 0; k < 8; ++k) {
            const int c = __reduce_add_sync(FULL, (unsigned)__popc(Bv[k] & ~shl1(Bv[k], lane)));
            if (lane == 0 && r0 + k < H) S.off[r0 + k + 1] = c;
        }
    }
    if (tid == 0) S.off[0] = 0;
    __syncthreads();
    // 2. off[] <- inclusive prefix (each thread a stretch of rows, a warp scan of the stretch sums, then the warps' sums)
    {
        const int per = (H + NT - 1) / NT, r0 = 1 + tid * per, r1 = min(H + 1, r0 + per);
        int sum = 0;
        for (int r = r0; r < r1; ++r) sum += S.off[r];
        const int before = lane_prefix(sum, lane);
        if (lane == 31) S.wsum[warp] = before + sum;
        __syncthreads();
        int base = before;
        for (int w = 0; w < warp; ++w) base += S.wsum[w];
        for (int r = r0; r < r1; ++r) { base += S.off[r]; S.off[r] = base; }
        __syncthreads();
    }
    const int n = S.off[H];
    if (n > cap) return false;
    if (n == 0) {
        for (int r = warp; r < H; r += nwarp) OUT.st(r, rowmask);
        return true;
    }
    for (int i = tid; i < n; i += NT) { S.parent[i] = i; S.size[i] = 0; }
    __syncthreads();
    // 3. unions with the row above and pixel counts, eight rows at a time per warp (the row above is carried in registers).
    //    Run i of this row and run j of the row above touch (8-connected) iff i meets j widened by one pixel either
    //    side; each such pair has exactly one first common column, which is either the start of i (then the row above has
    //    a pixel there or one column to the left) or the start of j minus one (then this row has a pixel there).  With
    //    SB(x) / SA(x) = run starts at columns <= x, the runs at such a column q are SB(q) - 1 and SA(q) - 1 in both cases.
    for (int r0 = warp * 8; r0 < H; r0 += nwarp * 8) {
        uint32_t A = 0, A1 = 0, sA = 0;
        int PA = 0;
        if (r0 > 0) {
            A = P0.ld(r0 - 1); A1 = shl1(A, lane); sA = A & ~A1;
            PA = S.off[r0 - 1] + lane_prefix(__popc(sA), lane);
        }
        uint32_t Bv[8];
        #pragma unroll
        for (int k = 0; k < 8; ++k) Bv[k] = (r0 + k < H) ? P0.ld(r0 + k) : 0u;
        #pragma unroll
        for (int k = 0; k < 8; ++k) {
            const int r = r0 + k;
            const uint32_t B = Bv[k];
            if (!__any_sync(FULL, B != 0u)) { A = A1 = sA = 0; PA = 0; continue; }
            const uint32_t B1 = shl1(B, lane), sB = B & ~B1;
            const int PB = S.off[r] + lane_prefix(__popc(sB), lane);
            uint32_t m = (sB & (A | A1)) | (sA & B1);
            while (m) {
                const int q = __ffs(m) - 1;
                m &= m - 1;
                const uint32_t le = le_mask(q);
                ccl_unite(S.parent, PB + __popc(sB & le) - 1, PA + __popc(sA & le) - 1);
            }
            m = B;
            while (m) {                                   // the pieces of runs inside this word
                const uint32_t low = m & (0u - m), t = m + low, piece = m & ~t;
                atomicAdd(&S.size[PB + __popc(sB & le_mask(__ffs(low) - 1)) - 1], (uint32_t)__popc(piece));
                m &= t;
            }
            A = B; A1 = B1; sA = sB; PA = PB;
        }
    }
    __syncthreads();
    // 4. every run learns its root; the roots collect the pixel counts
    for (int i = tid; i < n; i += NT) {
        const uint32_t r = ccl_find(S.parent, i);
        if (r != (uint32_t)i) { S.parent[i] = r; atomicAdd(&S.size[r], S.size[i]); }
    }
    __syncthreads();
    // 5. the largest root, the smallest number among equals: key = size << 32 | ~number
    unsigned long long best = 0;
    for (int i = tid; i < n; i += NT)
        if (S.parent[i] == (uint32_t)i) {
            const unsigned long long k = ((unsigned long long)S.size[i] << 32) | (uint32_t)~i;
            best = k > best ? k : best;
        }
    #pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const unsigned long long k = __shfl_xor_sync(FULL, best, o);
        best = k > best ? k : best;
    }
    if (lane == 0) S.wbest[warp] = best;
    __syncthreads();
    for (int w = 0; w < nwarp; ++w) best = S.wbest[w] > best ? S.wbest[w] : best;
    const uint32_t keep = ~(uint32_t)best;
    // 6. the rows of the result: the runs whose root is the winner
    for (int r0 = warp * 8; r0 < H; r0 += nwarp * 8) {
        uint32_t Bv[8];
        #pragma unroll
        for (int k = 0; k < 8; ++k) Bv[k] = (r0 + k < H) ? P0.ld(r0 + k) : 0u;
        #pragma unroll
        for (int k = 0; k < 8; ++k) {
            const int r = r0 + k;
            const uint32_t B = Bv[k];
            uint32_t out = 0;
            if (__any_sync(FULL, B != 0u)) {
                const uint32_t sB = B & ~shl1(B, lane);
                int id = S.off[r] + lane_prefix(__popc(sB), lane);
                uint32_t ks = 0;
                for (uint32_t m = sB; m; m &= m - 1, ++id)
                    if (S.parent[id] == keep) ks |= m & (0u - m);
                if (__any_sync(FULL, ks != 0u)) out = flood_row(B, ks, lane);
            }
            if (r < H) OUT.st(r, out);
        }
    }
    return true;
}

// One CTA per image, one band of rows per warp, one lane per 32-pixel word of a row.
__global__ void __launch_bounds__(32 * SSAD_PLANES_WARPS, SSAD_PLANES_MINB)
mask_planes_kernel(int batch, int H, int W, int nw, uint32_t* __restrict__ ws, long long plane_words,
                   uint32_t* __restrict__ mask_bits, int mask_pitch_words, long long mask_stride_words, int ccl_cap) {
    __shared__ int s_first, s_cnt;
    __shared__ CclSmem s_ccl;
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

#ifdef SSAD_PLANES_TIMING
    long long tk[8]; int ncomp = 0; tk[0] = clock64();
#define SSAD_TICK(i) tk[i] = clock64()
#else
#define SSAD_TICK(i)
#endif
    // 1. hysteresis: edges = pixels of the low plane 8-connected to a pixel of the high plane (P0 = low, P1 = high -> edges)
#ifdef SSAD_PLANES_TIMING
    const int rounds_h =
#endif
    reach<false>(P0, P1, H, ylo, yhi, rowmask, lane);
    SSAD_TICK(1);

    // 2-4a. binary_dilation(3x3), then binary_closing(3x3) = erosion(dilation), then the seeds of the hole fill, in ONE
    //    sweep: row y of the closing needs rows y-3..y+3 of the edges, so the warp streams its band plus three rows
    //    either side and keeps the two previous rows of each stage in registers (D1 = dilate(E), D2 = dilate(D1),
    //    C = erode3(D2); rows outside the image are 0 at every stage, as scipy's border_value=0).  P0 = C,
    //    P2 = seeds = background pixels of C on the image border.
    if (ylo <= yhi) {
        uint32_t e2 = 0, e1 = 0, d1b = 0, d1a = 0, d2b = 0, d2a = 0;       // E[yy-2], E[yy-1], D1[yy-3], D1[yy-2], D2[yy-4], D2[yy-3]
        #pragma unroll 1
        for (int yb = ylo - 3; yb <= yhi + 3; yb += 4) {                    // four row loads in flight
        uint32_t ev[4];
        #pragma unroll
        for (int k = 0; k < 4; ++k) ev[k] = (yb + k >= 0 && yb + k < H && yb + k <= yhi + 3) ? P1.ld(yb + k) : 0u;
        #pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int yy = yb + k;
            const uint32_t e0 = ev[k];
            const uint32_t d1 = (yy - 1 >= 0 && yy - 1 < H) ? (spread3(e2 | e1 | e0, lane) & rowmask) : 0u;      // D1[yy-1]
            const uint32_t d2 = (yy - 2 >= 0 && yy - 2 < H) ? (spread3(d1b | d1a | d1, lane) & rowmask) : 0u;    // D2[yy-2]
            const uint32_t t = d2b & d2a & d2;
            const uint32_t c = t & shl1(t, lane) & shr1(t, lane);                                                  // C[yy-3]
            const int y = yy - 3;
            if (y >= ylo && y <= yhi) {
                P0.st(y, c);
                uint32_t seed = ~c & rowmask;
                if (y != 0 && y != H - 1) {
                    uint32_t e = 0;
                    if (lane == 0) e |= 1u;
                    if (lane == nw - 1) e |= 1u << ((W - 1) & 31);
                    seed &= e;
                }
                P2.st(y, seed);
            }
            e2 = e1; e1 = e0; d1b = d1a; d1a = d1; d2b = d2a; d2a = d2;
        }
        }
    }
    __syncthreads();
    SSAD_TICK(2);
    // 4b. binary_fill_holes: background reachable from outside the image through ~C (8-connected); filled = the rest
#ifdef SSAD_PLANES_TIMING
    const int rounds_f =
#endif
    reach<true>(P0, P2, H, ylo, yhi, rowmask, lane);
    SSAD_TICK(3);
    // 5. binary_erosion(4x4), origin 0: offsets -2..+1 in both axes; P0 = erode4(~P2)
    if (ylo <= yhi) {
        auto fg = [&](int y) { return (y >= 0 && y < H) ? (~P2.ld(y) & rowmask) : 0u; };
        uint32_t f2 = fg(ylo - 2), f1 = fg(ylo - 1), f0 = fg(ylo);        // rows y-2, y-1, y
        #pragma unroll 1
        for (int yb = ylo; yb <= yhi; yb += 4) {                            // four row loads in flight
            uint32_t fv[4];
            #pragma unroll
            for (int k = 0; k < 4; ++k) fv[k] = (yb + k <= yhi) ? fg(yb + k + 1) : 0u;
            #pragma unroll
            for (int k = 0; k < 4; ++k) {
                const int y = yb + k;
                const uint32_t fn = fv[k];
                const uint32_t t = f2 & f1 & f0 & fn;
                const uint32_t l1 = shl1(t, lane), l2 = shl1(l1, lane);
                if (y <= yhi) {
                    P0.st(y, l2 & l1 & t & shr1(t, lane));
                    P1.st(y, 0u); OUT.st(y, 0u);                            // (step 6's scratch and result planes)
                }
                f2 = f1; f1 = f0; f0 = fn;
            }
        }
    }
    if (threadIdx.x == 0) { s_first = 0x7FFFFFFF; s_cnt = 0; }
    __syncthreads();
    SSAD_TICK(4);
    // 6. largest 8-connected component (skimage label + argmax(bincount)): peel components in raster order of their
    //    first pixel; strictly-greater keeps the first of equal sizes (= the lowest label).  P0 = remaining, P1 = scratch
    //    (zeroed by the erosion sweep above: the edges it held were last read before the hole fill).
    //    Usual case: few runs -> union-find over the runs; the flood peel below is the path for planes with more runs.
    if (H <= CCL_MAXH && ccl_cap > 0 && largest_component_runs(P0, OUT, H, rowmask, min(ccl_cap, CCL_CAP), s_ccl)) {
#ifdef SSAD_PLANES_TIMING
        SSAD_TICK(5);
        if (threadIdx.x == 0 && (b % 128) == 0)
            printf("planes b=%d hyst %lld (%d rounds) closing %lld fill %lld (%d rounds) erode %lld runs %lld total %lld\n", b, tk[1] - tk[0],
                   rounds_h, tk[2] - tk[1], tk[3] - tk[2], rounds_f, tk[4] - tk[3], tk[5] - tk[4], tk[5] - tk[0]);
#endif
        return;
    }
    int best = 0, blo = 1, bhi = 0;          // rows of the band the best component so far occupies in OUT
    int ys = ylo;                            // first row of the band that may still hold a pixel
    for (;;) {
        for (; ys <= yhi; ++ys)
            if (__any_sync(FULL, P0.ld(ys) != 0u)) break;
        if (lane == 0 && ys <= yhi) atomicMin(&s_first, ys);
        __syncthreads();
        const int y0 = s_first;
        if (y0 == 0x7FFFFFFF) break;
#ifdef SSAD_PLANES_TIMING
        ++ncomp;
#endif
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
#ifdef SSAD_PLANES_TIMING
    SSAD_TICK(5);
    if (threadIdx.x == 0 && (b % 128) == 0)
        printf("planes b=%d hyst %lld closing %lld fill %lld erode %lld peel %lld (components %d) total %lld\n", b, tk[1] - tk[0],
               tk[2] - tk[1], tk[3] - tk[2], tk[4] - tk[3], tk[5] - tk[4], ncomp, tk[5] - tk[0]);
#endif
}

}  // namespace ssad

// ------------------------------------------------------------------ C ABI
namespace {
struct ObjMaskLayout {
    int nw, ntx, nty;
    long long plane_words;
    unsigned pix_cap;
    size_t o_tiles, o_pix, o_count, total;
};
// test knobs (ssad_obj_mask_test_knobs): scale of the float32 error margins, undecided pixels a tile may list,
// list entries per image
float g_margin_scale = 1.0f;
int g_sub_cap = ssad::SUB_CAP;
int g_pix_per_image = 4096;
int g_ccl_cap = ssad::CCL_CAP;          // runs the union-find of the largest component takes (0: always the flood peel)

ObjMaskLayout obj_mask_layout(int batch, int height, int width) {
    ObjMaskLayout L;
    L.nw = (width + 31) / 32;
    L.ntx = (width + ssad::CT_W - 1) / ssad::CT_W;
    L.nty = (height + ssad::CT_H - 1) / ssad::CT_H;
    L.plane_words = (long long)height * L.nw;
    auto up = [](size_t v) { return (v + 255) / 256 * 256; };
    const size_t ntile = (size_t)batch * L.ntx * L.nty;
    const size_t cap = (size_t)batch * 4096;              // (the workspace is sized for the default knob)
    L.pix_cap = cap > 0x7FFFFFFFull ? 0x7FFFFFFFu : (unsigned)cap;
    L.o_tiles = up((size_t)batch * 3 * (size_t)L.plane_words * sizeof(uint32_t));
    L.o_pix = L.o_tiles + up(ntile * sizeof(uint32_t));
    L.o_count = L.o_pix + up((size_t)L.pix_cap * sizeof(uint2));
    L.total = L.o_count + 256;
    return L;
}
}  // namespace

extern "C" size_t ssad_obj_mask_workspace(int32_t batch, int32_t height, int32_t width) {
    if (batch <= 0 || height <= 0 || width <= 0) return 0;
    return obj_mask_layout(batch, height, width).total;
}

// kernels one ssad_obj_mask call launches (bench.py's gpu_launches): float32 canny, dense float64 tiles, float64 pixels,
// bit planes
extern "C" int ssad_obj_mask_launches(void) { return 4; }

// Test hook (tests/test_gpu_objmask.py): widen the float32 error margins by `margin_scale` (more pixels take the
// float64 paths), cap the undecided pixels a tile may list (0: every tile with one goes to the dense kernel) and the
// list entries per image.  Arguments <= 0 restore the defaults.  Results must not depend on any of them.
extern "C" void ssad_obj_mask_test_knobs(float margin_scale, int32_t sub_cap, int32_t pix_per_image) {
    g_margin_scale = margin_scale > 0.0f ? margin_scale : 1.0f;
    g_sub_cap = (sub_cap >= 0 && sub_cap <= ssad::SUB_CAP) ? sub_cap : ssad::SUB_CAP;
    g_pix_per_image = (pix_per_image >= 0 && pix_per_image <= 4096) ? pix_per_image : 4096;
}

extern "C" void ssad_obj_mask_test_runs(int32_t max_runs) {
    g_ccl_cap = (max_runs >= 0 && max_runs <= ssad::CCL_CAP) ? max_runs : ssad::CCL_CAP;
}

extern "C" int ssad_obj_mask(int32_t batch, int32_t height, int32_t width, const uint8_t* src, int64_t src_pitch,
                             int64_t src_stride, const SsadCannyConsts* consts, uint32_t* mask_bits, int32_t mask_pitch_words,
                             int64_t mask_stride_words, void* workspace, size_t workspace_bytes, void* stream) {
    if (batch < 0 || height <= 0 || width <= 0 || !src || !mask_bits || !consts) return SSAD_EINVAL;
    if (src_pitch < 3ll * width || mask_pitch_words < (width + 31) / 32) return SSAD_EINVAL;
    if (batch == 0) return SSAD_OK;
    if (width > 1024 || width < 13 || height < 13) return SSAD_EUNSUPPORTED;   // one lane per row word; 6-pixel border classes
    if (batch > 65535 || height > 65535) return SSAD_EUNSUPPORTED;
    const ObjMaskLayout L = obj_mask_layout(batch, height, width);
    if ((unsigned long long)batch * L.ntx * L.nty > 0xFFFFFFFFull) return SSAD_EUNSUPPORTED;
    if (!workspace || workspace_bytes < L.total) return SSAD_EWORKSPACE;
    if (reinterpret_cast<uintptr_t>(workspace) & 15) return SSAD_EALIGN;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    ssad::CannyK K;
    for (int i = 0; i < 13; ++i) K.w[i] = consts->gauss_w[i];
    K.low = consts->low; K.high = consts->high; K.eps = consts->eps;
    ssad::C32K FK;
    for (int i = 0; i < 7; ++i) FK.w[i] = (float)consts->gauss_w[i];
    FK.ec[0] = 0.003f * g_margin_scale; FK.em[0] = 0.0045f * g_margin_scale;
    FK.ec[1] = 0.0125f * g_margin_scale; FK.em[1] = 0.018f * g_margin_scale;
    for (int c = 0; c < 2; ++c) {
        const float tl = (float)(consts->low * 255.0), th = (float)(consts->high * 255.0);
        FK.lo_no[c] = tl - FK.em[c] > 0.0f ? tl - FK.em[c] : 0.0f;
        FK.lo_yes[c] = tl + FK.em[c];
        FK.hi_no[c] = th - FK.em[c] > 0.0f ? th - FK.em[c] : 0.0f;
        FK.hi_yes[c] = th + FK.em[c];
        FK.lo_no2[c] = FK.lo_no[c] * FK.lo_no[c] * 0.999999f;
    }
    FK.sub_cap = g_sub_cap;
    {
        const unsigned long long cap = (unsigned long long)batch * (unsigned)g_pix_per_image;
        FK.pix_cap = cap < L.pix_cap ? (unsigned)cap : L.pix_cap;
    }
    const int nw = L.nw;
    const long long plane_words = L.plane_words;
    uint8_t* const wsb = static_cast<uint8_t*>(workspace);
    uint32_t* ws = static_cast<uint32_t*>(workspace);
    uint32_t* const tiles = reinterpret_cast<uint32_t*>(wsb + L.o_tiles);
    uint2* const pix = reinterpret_cast<uint2*>(wsb + L.o_pix);
    uint32_t* const count = reinterpret_cast<uint32_t*>(wsb + L.o_count);     // [0] undecided pixels, [1] dense tiles
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
            e = cudaFuncSetAttribute(ssad::canny32_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(ssad::C32Smem));
        if (e != cudaSuccess) return ssad_set_cuda_error(e);
        sm_count[dev] = n > 0 ? n : 1;
    }
    e = cudaMemsetAsync(count, 0, 2 * sizeof(uint32_t), st);
    if (e != cudaSuccess) return ssad_set_cuda_error(e);
    // 1. float32 canny: both planes for every pixel it can decide, the undecided pixels / tiles listed
    dim3 sgrid((width + ssad::SC_W - 1) / ssad::SC_W, (height + ssad::SC_H - 1) / ssad::SC_H, batch);
    ssad::canny32_kernel<<<sgrid, ssad::SC_NT, sizeof(ssad::C32Smem), st>>>(FK, height, width, src, src_pitch, src_stride, ws,
                                                                          ws + plane_words, nw, 3 * plane_words, pix, count,
                                                                          tiles, count + 1, L.ntx, L.nty);
    // 2. dense float64 canny on the listed tiles (usually none), 3. float64 for the listed pixels
    ssad::canny_nms_kernel<<<sm_count[dev] * SSAD_CANNY_MINB, ssad::CT_NT, sizeof(ssad::CannySmem), st>>>(
        K, height, width, src, src_pitch, src_stride, ws, ws + plane_words, nw, 3 * plane_words, tiles, count + 1, L.ntx, L.nty);
    ssad::canny_resolve_kernel<<<sm_count[dev] * 4, ssad::RS_NT, 0, st>>>(K, height, width, src, src_pitch, src_stride, ws,
                                                                        ws + plane_words, nw, 3 * plane_words, pix, count, FK.pix_cap);
    e = cudaGetLastError();
    if (e != cudaSuccess) return ssad_set_cuda_error(e);
    // 4. hysteresis, morphology, hole filling, largest component on the bit planes: a CTA per image, a band of rows per warp
    int nwarp = height / SSAD_PLANES_ROWS;
    nwarp = nwarp < 1 ? 1 : (nwarp > SSAD_PLANES_WARPS ? SSAD_PLANES_WARPS : nwarp);
    ssad::mask_planes_kernel<<<batch, 32 * nwarp, 0, st>>>(batch, height, width, nw, ws, plane_words, mask_bits, mask_pitch_words,
                                                         mask_stride_words, g_ccl_cap);
    e = cudaGetLastError();
    if (e != cudaSuccess) return ssad_set_cuda_error(e);
    return SSAD_OK;
}

// ------------------------------------------------------------------ ssad_obj_mask_static: no caller workspace, no allocation
// The scratch of ssad_obj_mask in a file-scope device array (include/ssad.h: why, the limit, and the sharing rules).
// In host code the NAME of the array is a host-side stand-in, not a device address: the address the kernels get comes
// from cudaGetSymbolAddress, asked once per device with that device current.
static __device__ __align__(256) unsigned char g_obj_mask_static_scratch[SSAD_OBJ_MASK_STATIC_BYTES];

extern "C" int ssad_obj_mask_static(int32_t batch, int32_t height, int32_t width, const uint8_t* src, int64_t src_pitch,
                                    int64_t src_stride, const SsadCannyConsts* consts, uint32_t* mask_bits,
                                    int32_t mask_pitch_words, int64_t mask_stride_words, void* stream) {
    if (batch < 0 || height <= 0 || width <= 0 || !src || !mask_bits || !consts) return SSAD_EINVAL;
    if (batch == 0) return SSAD_OK;
    if (width > 1024 || batch > 65535 || height > 65535) return SSAD_EUNSUPPORTED;      // (keeps the layout arithmetic in range)
    if (obj_mask_layout(batch, height, width).total > (size_t)SSAD_OBJ_MASK_STATIC_BYTES) return SSAD_ESTATIC;
    static void* scratch_of[64];                   // device address of the array on every device, asked once
    static std::mutex mu;
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return ssad_set_cuda_error(e);
    if (dev < 0 || dev >= 64) return SSAD_EUNSUPPORTED;
    void* scratch = nullptr;
    {
        std::lock_guard<std::mutex> lock(mu);
        if (!scratch_of[dev]) {
            void* p = nullptr;
            e = cudaGetSymbolAddress(&p, g_obj_mask_static_scratch);       // the current device's copy
            if (e != cudaSuccess || !p) return ssad_set_cuda_error(e != cudaSuccess ? e : cudaErrorInvalidSymbol);
            scratch_of[dev] = p;
        }
        scratch = scratch_of[dev];
    }
    // the planes, lists and counters are laid out exactly as in a caller workspace; ssad_obj_mask zeroes the two
    // counters on the stream (cudaMemsetAsync) and writes every other word before reading it
    return ssad_obj_mask(batch, height, width, src, src_pitch, src_stride, consts, mask_bits, mask_pitch_words, mask_stride_words,
                         scratch, (size_t)SSAD_OBJ_MASK_STATIC_BYTES, stream);
}
