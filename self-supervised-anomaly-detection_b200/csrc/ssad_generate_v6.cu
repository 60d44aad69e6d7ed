// ssad_generate, B200 path v6: one cluster of C CTAs per image, rows split over the CTAs, pixels streamed twice
// through registers (HBM, then L2); the ColorJitter chain runs as per-image TABLES around an unclamped Saturation.
//
// PretextTaskDataset.__getitem__ (datasets.py:209-394) has two whole-image dependencies: the mean colour of x
// (check_color_similarity, dataset_generator.py:147-159) and the mean of L that ImageEnhance.Contrast blends towards
// (after the paste and after the ColorJitter ops that precede Contrast).  Per image:
//
//   A  geometry   parameters -> smem; mask-window prefix -> coords_map[randint] centres -> container clamp ->
//                 paste boxes; rect2poly / rotated-scar coverage bit rows; rows-with-a-paste flags; the jitter plan
//                 and the tables of the ops that precede Contrast.  Nothing here reads a pixel.
//   B  pass 1     stream the slab: (RandomAffine gather) ; RGB sums of x ; ops that precede Contrast ; sum of L of
//                 that (unpasted) image ; RGB sums of the would-be cut.  The bytes are kept in the x buffer when they
//                 were expensive (gather, Saturation).
//      exchange 1 partial sums -> every CTA of the cluster (st.shared::cluster) -> barrier.cluster
//   C  decide     cosine > 0.99 -> double-Brightness branch, flat colours
//   D  correct    covered pixels only: sum of L(new colour) - L(old pixel) -> exchange 2 -> Contrast mean -> tables of
//                 Contrast and the ops after it
//   E  pass 2     stream the slab again (L2 hits): paste on the fly -> rest of the chain -> stores
//
// Jitter arithmetic.  Brightness and Contrast map a channel value to a channel value, so any run of them -- in the
// image's own order, with their clips -- is ONE 256-entry table per image, filled with Pillow's float32 sequence
// (ImageEnhance -> Blend.c).  Saturation (ImageEnhance.Color) needs the pixel's L: it runs in
// registers on 8-pixel units as packed f32x2 operations, each product and sum rounded on its own, and leaves its
// result UNCLIPPED as 2^23 + t rounded toward zero; the low bits of that float index the table of whatever follows
// (at least the clip), so no min/max is ever executed per pixel.  Results are bit-identical to Pillow's.
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "ssad_device.cuh"
#include "ssad_sm100.cuh"
#include "ssad_unit8.cuh"

int ssad_set_cuda_error(cudaError_t e);

#include "ssad_v6_common.cuh"

namespace ssad {
namespace v6 {


// ------------------------------------------------------------------ the kernel
#ifndef SSAD_V6_MINB
#define SSAD_V6_MINB 4
#endif
__global__ void __launch_bounds__(NT, SSAD_V6_MINB)
generate_v6_kernel(const __grid_constant__ SsadBatchDesc d, const __grid_constant__ Geom g,
                   const SsadImageParams* __restrict__ params, const uint8_t* __restrict__ src,
                   const uint8_t* __restrict__ donor_bank, const uint32_t* __restrict__ mask_bits,
                   const void* __restrict__ norm_lut, uint8_t* __restrict__ out_u8, void* __restrict__ out_norm,
                   int64_t* __restrict__ y_out, SsadDebug* __restrict__ debug, uint8_t* xbuf, uint32_t* canvas_all,
                   unsigned long long* prof) {
    extern __shared__ __align__(16) uint8_t smem[];
    __shared__ SsadImageParams P;
    __shared__ Plan D;
    __shared__ Polygon G;
    __shared__ Tables TP, TF;                    // tables of the `pre` and `fin` programs
    __shared__ uint32_t red[NWARP * 8 + 8];
    __shared__ uint32_t xchg[MAXC][8];
    __shared__ uint32_t stage[8];
    uint32_t* const rowpre = reinterpret_cast<uint32_t*>(smem + g.o_rowpre);
    uint32_t* const cov = reinterpret_cast<uint32_t*>(smem + g.o_cov);
    uint8_t* const flags = smem + g.o_flags;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int C = g.C;
    const uint32_t rank = (C > 1) ? cluster_ctarank() : 0u;
    const int b = (int)(blockIdx.x / (unsigned)C);
    const int H = g.H, W = g.W;
    const int npix = H * W;
    if (C > 1) cluster_arrive();                 // #0: "this CTA runs" (waited for before the first remote store)
    long long tprev = prof ? clock64() : 0;
    #define SSAD_MARK(i) if (prof && tid == 0) { const long long tn = clock64(); atomicAdd(prof + (i), (unsigned long long)(tn - tprev)); tprev = tn; }

    // ================================================================ A: geometry (no pixel is read)
    if (tid < (int)(sizeof(SsadImageParams) / 4))
        reinterpret_cast<uint32_t*>(&P)[tid] = reinterpret_cast<const uint32_t*>(params + b)[tid];
    if (out_norm != nullptr) {
        const int nw = (d.out_norm_dtype == SSAD_NORM_F32) ? 768 : 384;
        for (int i = tid; i < nw; i += NT)
            reinterpret_cast<uint32_t*>(smem + g.o_lut)[i] = reinterpret_cast<const uint32_t*>(norm_lut)[i];
    }
    __syncthreads();

    const int r0 = (int)rank * g.rows_cta, r1 = min(H, r0 + g.rows_cta);
    const int cx0 = P.crop_off[0], cy0 = P.crop_off[1];
    const uint32_t* mrow0 = nullptr;
    if (d.mask_mode != SSAD_MASK_ALL)
        mrow0 = mask_bits + (d.mask_mode == SSAD_MASK_PER_IMAGE ? (size_t)b * d.mask_stride_words : 0);
    const bool defect = P.label != 0;
    const int pw = P.rect[2], ph = P.rect[3];
    if (defect && d.mask_mode != SSAD_MASK_ALL) {
        // mask window: per-row popcounts -> prefix (coords_map is row-major, datasets.py:263-264)
        for (int r = tid; r < H; r += NT)
            rowpre[r + 1] = row_window_popc(mrow0 + (size_t)(cy0 + r) * d.mask_pitch_words, cx0, W);
        __syncthreads();
        if (tid < 32) {
            uint32_t carry = 0;
            for (int base = 0; base < H; base += 32) {
                int r = base + tid;
                uint32_t v = (r < H) ? rowpre[r + 1] : 0, s = v;
                for (int o = 1; o < 32; o <<= 1) {
                    uint32_t t = __shfl_up_sync(0xFFFFFFFFu, s, o);
                    if (tid >= o) s += t;
                }
                if (r < H) rowpre[r + 1] = carry + s;
                carry += __shfl_sync(0xFFFFFFFFu, s, 31);
            }
            if (tid == 0) rowpre[0] = 0;
        }
        __syncthreads();
    }
    const int n_coords = (d.mask_mode == SSAD_MASK_ALL) ? npix : (defect ? (int)rowpre[H] : 0);
    int y = P.label;
    if (defect && d.coverage_min3 > 0 && 3 * n_coords < d.coverage_min3) y = 0;   // datasets.py:258-259
    const int npaste = (y == 0) ? 0 : (y == 1 ? 1 : P.scar_k);
    const int bw = (y == 1) ? pw : P.scar_size[0], bh = (y == 1) ? ph : P.scar_size[1];
    if (tid < npaste) {
        // centres + container clamp (dataset_generator.py:104-144, 259-265): one thread per paste
        int cx, cy;
        if (n_coords == 0) {
            cx = P.fallback_c[tid][0]; cy = P.fallback_c[tid][1];
        } else {
            uint32_t idx = (n_coords < 2) ? 0u : (uint32_t)(((uint64_t)P.centre_r[tid] * (uint64_t)n_coords) >> 32);
            if (d.mask_mode == SSAD_MASK_ALL) {
                cy = idx / W; cx = idx - cy * W;
            } else {
                int lo = 0, hi = H;
                while (hi - lo > 1) { int mid = (lo + hi) >> 1; if (rowpre[mid] <= idx) lo = mid; else hi = mid; }
                cy = lo;
                cx = row_window_select(mrow0 + (size_t)(cy0 + cy) * d.mask_pitch_words, cx0, W, idx - rowpre[lo]);
            }
        }
        int pl, pt;
        clamp_to_container(y == 1 ? d.box_poly : d.box_scar, bw, bh, cx, cy, pl, pt);
        D.paste[tid][0] = pl; D.paste[tid][1] = pt;
    }
    if (y == 1 && tid == 32) polygon_setup(G, P.poly_xy, P.poly_n, ph);
    if (tid == 64) {
        D.y = y; D.npaste = npaste; D.n_coords = n_coords; D.bw = bw; D.bh = bh;
        D.sim = 0; D.flat = 0; D.cosine = 0.0; D.cmean = -1;
        D.flat_rgb[0] = D.flat_rgb[1] = D.flat_rgb[2] = 0;
        D.avg_rgb[0] = D.avg_rgb[1] = D.avg_rgb[2] = 0;
        D.bf1 = P.bright_f[0]; D.bf2 = P.bright_f[1];
        // ColorJitter plan (datasets.py:391): ops in fn_idx order, positions of Contrast and Saturation
        int nops = 0, cpos = -1, spos = -1;
        D.ops[0] = D.ops[1] = D.ops[2] = 0; D.fac[0] = D.fac[1] = D.fac[2] = 1.0f;
        for (int st = 0; st < 4; ++st) {
            int fn = P.jit_order[st];
            if (fn < 3 && ((P.jit_en >> fn) & 1)) {
                if (fn == 1) cpos = nops;
                if (fn == 2) spos = nops;
                D.ops[nops] = fn; D.fac[nops] = P.jit_f[fn]; ++nops;
            }
        }
        D.nops = nops; D.cpos = cpos;
        D.s_in_pre = (cpos >= 0 && spos >= 0 && spos < cpos) ? 1 : 0;
    }
    __syncthreads();
    const int cpos = D.cpos, nops = D.nops;
    const bool s_in_pre = D.s_in_pre != 0;
    build_prog(D.pre, TP, D.ops, D.fac, 0, max(cpos, 0), 0, tid);      // ops ahead of Contrast (never Contrast itself)
    const int pitchw = (int)g.cov_pitch;
    if (y == 1) {
        // rect2poly bit rows (dataset_generator.py:63-100) of the canvas rows that meet this CTA's slab
        const int pt = D.paste[0][1];
        const int v_lo = max(0, r0 - pt), v_hi = min(ph, r1 - pt);
        for (int v = v_lo + tid; v < v_hi; v += NT) {
            uint32_t* row = cov + (size_t)v * pitchw;
            for (int j = 0; j < pitchw; ++j) row[j] = 0;
            polygon_fill_row(G, v, pw, row);
        }
    } else if (y == 2) {
        // alpha of scar.rotate(angle, expand=True) (datasets.py:344): canvas pixel (u, v) is opaque iff its inverse
        // map (affine_fixed, 16.16) lands inside the pw x ph scar -- per canvas row one span of u, in closed form
        for (int v = tid; v < bh; v += NT) {
            uint32_t* row = cov + (size_t)v * pitchw;
            for (int j = 0; j < pitchw; ++j) row[j] = 0;
            int lo = 0, hi = bw - 1;
            affine_span((long long)P.scar_fix[2] + (long long)v * P.scar_fix[1], P.scar_fix[0], (long long)pw << 16, lo, hi);
            affine_span((long long)P.scar_fix[5] + (long long)v * P.scar_fix[4], P.scar_fix[3], (long long)ph << 16, lo, hi);
            if (lo <= hi) row_set_bits(row, lo, hi);
        }
    }
    for (int r = r0 + tid; r < r1; r += NT) {
        uint32_t f = 0;
        for (int k = 0; k < npaste; ++k) {
            const int v = r - D.paste[k][1];
            if (v >= 0 && v < bh) f |= 1u << k;
        }
        flags[r - r0] = (uint8_t)f;
    }
    __syncthreads();

    SSAD_MARK(0)
    // ================================================================ B: pass 1
    const bool warped = P.warp_en != 0;
    const bool stored = warped || s_in_pre;      // pass 1 leaves its bytes in the x buffer
    const int nunits = (r1 - r0) * g.upr;
    const uint8_t* const sbase = src + (size_t)b * d.src_stride + (size_t)cy0 * d.src_pitch + (size_t)cx0 * 3;
    const int spitch = (int)d.src_pitch;
    uint8_t* const xb = xbuf + (size_t)b * npix * 3;
    const int xpitch = W * 3;
    const Donor dn = donor_of(d, src, donor_bank, b, P);

    uint32_t* const canvas = canvas_all + (size_t)b * g.canvas_stride;
    Boxes BX;
    BX.cov = cov; BX.canvas = canvas; BX.pitch = pitchw; BX.cpitch = (int)g.canvas_pitch; BX.bw = bw; BX.bh = bh; BX.npaste = npaste;

    uint32_t acc[7] = {0, 0, 0, 0, 0, 0, 0};          // 0-2 RGB of x, 3 sum of L, 4-6 RGB of the cut
    const Slab SL = make_slab(r0, nunits, g.upr, tid);
    if (warped) {
        // RandomAffine(3, scale=(1.05, 1.1)) NEAREST, fill 0 (datasets.py:220-222): x gathered once into the x buffer
        const int32_t a0 = P.warp_fix[0], a1 = P.warp_fix[1], a2 = P.warp_fix[2];
        const int32_t a3 = P.warp_fix[3], a4 = P.warp_fix[4], a5 = P.warp_fix[5];
        SSAD_SLAB_LOOP(SL) {
            const int yy = r0 + rr, x0 = cu * 8;
            const int xr = a2 + yy * a1 + x0 * a0, yr = a5 + yy * a4 + x0 * a3;
            uint32_t w[6] = {0, 0, 0, 0, 0, 0};
            #define SSAD_GATHER(PIX)                                                                      \
                { const int xin = (xr + PIX * a0) >> 16, yin = (yr + PIX * a3) >> 16;                     \
                  if ((unsigned)xin < (unsigned)W && (unsigned)yin < (unsigned)H) {                       \
                      const uint8_t* q = sbase + yin * spitch + xin * 3;                                  \
                      put_pixel<PIX>(w, __ldg(q), __ldg(q + 1), __ldg(q + 2)); } }
            SSAD_GATHER(0) SSAD_GATHER(1) SSAD_GATHER(2) SSAD_GATHER(3) SSAD_GATHER(4) SSAD_GATHER(5) SSAD_GATHER(6) SSAD_GATHER(7)
            #undef SSAD_GATHER
            uint2* const o = reinterpret_cast<uint2*>(xb + (size_t)r0 * xpitch + (size_t)lin * 24);
            o[0] = make_uint2(w[0], w[1]); o[1] = make_uint2(w[2], w[3]); o[2] = make_uint2(w[4], w[5]);
        }
    }
    // x as the passes read it (each thread re-reads only what it wrote itself)
    const uint8_t* const xin_base = warped ? xb : sbase;
    const int xin_pitch = warped ? xpitch : spitch;
    if (y != 0 || cpos >= 0) {
        ProgRegs R;
        R.fs = D.pre.fs; R.ta = smem_u32(TP.a); R.tab = smem_u32(TP.ab); R.tbc = smem_u32(TP.b) + (uint32_t)TB_OFF - kMagicBits;
        const bool need_rgb = y != 0, need_luma = cpos >= 0;
        switch (loop_kind(D.pre)) {
            case LK_COPY:    pass1_loop<LK_COPY>(SL, xin_base, xin_pitch, !warped, xb, xpitch, false, need_rgb, need_luma, R, D, BX, flags, acc); break;
            case LK_BYTELUT: pass1_loop<LK_BYTELUT>(SL, xin_base, xin_pitch, !warped, xb, xpitch, false, need_rgb, need_luma, R, D, BX, flags, acc); break;
            case LK_SAT_UC:  pass1_loop<LK_SAT_UC>(SL, xin_base, xin_pitch, !warped, xb, xpitch, true, need_rgb, need_luma, R, D, BX, flags, acc); break;
            case LK_SAT_UL:  pass1_loop<LK_SAT_UL>(SL, xin_base, xin_pitch, !warped, xb, xpitch, true, need_rgb, need_luma, R, D, BX, flags, acc); break;
            case LK_SAT_LC:  pass1_loop<LK_SAT_LC>(SL, xin_base, xin_pitch, !warped, xb, xpitch, true, need_rgb, need_luma, R, D, BX, flags, acc); break;
            default:         pass1_loop<LK_SAT_LL>(SL, xin_base, xin_pitch, !warped, xb, xpitch, true, need_rgb, need_luma, R, D, BX, flags, acc); break;
        }
        // RGB sums of the would-be cut (dataset_generator.py:202-207 / 147-159): rows dealt over the cluster's warps
        if (y != 0 && P.src_variant != 2) {
            for (int v = (int)rank * NWARP + warp; v < dn.v_hi; v += C * NWARP) {
                const uint8_t* q = dn.base + (size_t)v * dn.pitch;
                for (int u = lane; u < dn.u_hi; u += 32) {
                    acc[4] += __ldg(q + 3 * u); acc[5] += __ldg(q + 3 * u + 1); acc[6] += __ldg(q + 3 * u + 2);
                }
            }
        }
    }

    SSAD_MARK(1)
    // ---------------------------------------------------------------- exchange 1: whole-image sums
    // every CTA leaves its partials in slot [rank] of every CTA of the cluster, then all add the C slots
    if (C > 1) cluster_wait();                   // #0: every CTA of the cluster has started
    uint32_t tot[7];
    {
        block_sum_vec<7>(acc, red);
        #pragma unroll
        for (int i = 0; i < 7; ++i) tot[i] = acc[i];
        if (C > 1) {
            #pragma unroll
            for (int i = 0; i < 7; ++i) if (tid == i) stage[i] = tot[i];
            __syncthreads();
            if (tid < 7 * C) {
                const int target = tid / 7, i = tid - target * 7;
                st_cluster_u32(&xchg[rank][i], (uint32_t)target, stage[i]);
            }
            cluster_arrive(); cluster_wait();
            #pragma unroll
            for (int i = 0; i < 7; ++i) {
                uint32_t t = 0;
                for (int r = 0; r < C; ++r) t += xchg[r][i];
                tot[i] = t;
            }
        }
    }

    SSAD_MARK(2)
    // ================================================================ C: check_color_similarity + variant colours
    if (y != 0) {
        if (tid == 0) {
            const uint32_t np = (uint32_t)(pw * ph);
            uint32_t xs[3] = {tot[0], tot[1], tot[2]};
            uint32_t psum[3] = {tot[4], tot[5], tot[6]};
            int flat = 0;
            uint32_t frgb[3] = {0, 0, 0};
            if (P.src_variant == 1) {           // colour = int(mean) (dataset_generator.py:202-207)
                flat = 1;
                for (int k = 0; k < 3; ++k) { frgb[k] = psum[k] / np; D.avg_rgb[k] = (int)frgb[k]; psum[k] = frgb[k] * np; }
            } else if (P.src_variant == 2) {    // random colour (dataset_generator.py:194-199)
                flat = 1;
                for (int k = 0; k < 3; ++k) { frgb[k] = P.rnd_rgb[k]; psum[k] = frgb[k] * np; }
            }
            double cs;
            const bool sim = similarity_branch(xs, (double)npix, psum, (double)np, cs);     // datasets.py:295 / 335
            if (flat && sim)                    // brightness of a flat patch is a flat patch
                for (int k = 0; k < 3; ++k) frgb[k] = brightness_u8(brightness_u8(frgb[k], D.bf1), D.bf2);
            D.flat = flat; D.sim = sim ? 1 : 0; D.cosine = cs;
            for (int k = 0; k < 3; ++k) D.flat_rgb[k] = frgb[k];
        }
        __syncthreads();
    }
    Colour K;
    K.flat = D.flat; K.sim = D.sim; K.bf1 = D.bf1; K.bf2 = D.bf2;
    K.frgb[0] = D.flat_rgb[0]; K.frgb[1] = D.flat_rgb[1]; K.frgb[2] = D.flat_rgb[2];

    // ================================================================ D: patch canvas, Contrast mean
    // canvas pixel (u, v) -> packed (r, g, b, L): the colour pass 2 pastes (after the `pre` program when pass 1 stored
    // pre(x)) and the L of that colour after the ops ahead of Contrast.  Each CTA fills the canvas rows its slab meets,
    // one warp per row.  The double Brightness of the similarity branch (datasets.py:298-299) is a per-channel map: a
    // 256-entry table (TF.ab is free until the final tables are built).
    uint32_t lnew = 0;                           // sum of L over the pixels the pastes leave visible
    if (y != 0) {
        const bool cut_sim = !K.flat && K.sim;
        if (cut_sim) {
            for (int v = tid; v < 256; v += NT) TF.ab[v] = (uint8_t)brightness_u8(brightness_u8((uint32_t)v, K.bf1), K.bf2);
            __syncthreads();
        }
        int cv_lo = bh, cv_hi = 0;
        for (int k = 0; k < npaste; ++k) {
            const int pt = D.paste[k][1];
            const int v_lo = max(0, r0 - pt), v_hi = min(bh, r1 - pt);
            if (v_lo < v_hi) { cv_lo = min(cv_lo, v_lo); cv_hi = max(cv_hi, v_hi); }
        }
        const int cpitch = (int)g.canvas_pitch;
        const int pl0 = D.paste[0][0];
        const bool want_l = cpos >= 0;
        for (int v = cv_lo + warp; v < cv_hi; v += NWARP) {
            const uint32_t* crow = cov + (size_t)v * pitchw;
            uint32_t* out = canvas + (size_t)v * cpitch;
            const int xr = P.scar_fix[2] + v * P.scar_fix[1], yr = P.scar_fix[5] + v * P.scar_fix[4];
            for (int u = lane; u < bw; u += 32) {
                if (!((crow[u >> 5] >> (u & 31)) & 1u)) continue;
                uint32_t r, gg, bl;
                if (K.flat) { r = K.frgb[0]; gg = K.frgb[1]; bl = K.frgb[2]; }
                else {
                    int su = u, sv = v;          // source pixel of the cut: itself (polygon) | inverse rotation (scar)
                    if (y == 2) { su = (xr + u * P.scar_fix[0]) >> 16; sv = (yr + u * P.scar_fix[3]) >> 16; }
                    r = gg = bl = 0;
                    if (su < dn.u_hi && sv < dn.v_hi) {
                        const uint8_t* q = dn.base + (size_t)sv * dn.pitch + 3 * su;
                        r = __ldg(q); gg = __ldg(q + 1); bl = __ldg(q + 2);
                    }
                    if (cut_sim) { r = TF.ab[r]; gg = TF.ab[gg]; bl = TF.ab[bl]; }
                }
                uint32_t L = 0;
                if (want_l) {
                    uint32_t pr = r, pg = gg, pb = bl;
                    pixel_prog(D.pre, TP, pr, pg, pb);
                    L = luma_u8(pr, pg, pb);
                    if (s_in_pre) { r = pr; gg = pg; bl = pb; }
                    // a polygon is pasted once: visible wherever it lands inside the image (rows: this slab by construction)
                    if (y == 1 && (unsigned)(pl0 + u) < (unsigned)W) lnew += L;
                }
                out[u] = r | (gg << 8) | (bl << 16) | (L << 24);
            }
        }
        __syncthreads();
    }
    SSAD_MARK(3)
    int cmean = -1;
    if (cpos >= 0) {
        if (y != 0) {
            if (y == 2) {
                // scars: k pastes of the same canvas, later ones overwrite earlier ones
                const int cpitch = (int)g.canvas_pitch;
                for (int k = 0; k < npaste; ++k) {
                    const int pl = D.paste[k][0], pt = D.paste[k][1];
                    const int v_lo = max(0, r0 - pt), v_hi = min(bh, r1 - pt);
                    for (int v = v_lo + warp; v < v_hi; v += NWARP) {
                        const int Y = pt + v;
                        for (int u = lane; u < bw; u += 32) {
                            const int X = pl + u;
                            if (X < 0 || X >= W) continue;
                            if (!((cov[(size_t)v * pitchw + (u >> 5)] >> (u & 31)) & 1u)) continue;
                            bool later = false;          // a later paste overwrites this pixel: it is counted there
                            for (int j = k + 1; j < npaste; ++j) {
                                const int uj = X - D.paste[j][0], vj = Y - D.paste[j][1];
                                if (uj >= 0 && uj < bw && vj >= 0 && vj < bh && ((cov[(size_t)vj * pitchw + (uj >> 5)] >> (uj & 31)) & 1u)) later = true;
                            }
                            if (later) continue;
                            lnew += canvas[(size_t)v * cpitch + u] >> 24;
                        }
                    }
                }
            }
            uint32_t v1[1] = {lnew};
            block_sum_vec<1>(v1, red);
            lnew = v1[0];
            if (C > 1) {
                if (tid < C) st_cluster_u32(&xchg[rank][7], (uint32_t)tid, lnew);
                cluster_arrive(); cluster_wait();
                lnew = 0;
                for (int r = 0; r < C; ++r) lnew += xchg[r][7];
            }
        }
        // int(mean + 0.5) with mean = sum / count in float64 (ImageEnhance.py Contrast) == integer rounding
        const uint32_t lsum = tot[3] + lnew;
        cmean = (int)((2ull * lsum + (uint32_t)npix) / (2ull * (uint32_t)npix));
    }
    // what pass 2 still has to do: Contrast and the ops after it on pre(x), or the whole chain on x
    build_prog(D.fin, TF, D.ops, D.fac, s_in_pre ? cpos : 0, nops, cmean, tid);
    __syncthreads();

    SSAD_MARK(4)
    // ================================================================ E: pass 2 -- paste, rest of the chain, stores
    {
        ProgRegs R;
        R.fs = D.fin.fs; R.ta = smem_u32(TF.a); R.tab = smem_u32(TF.ab); R.tbc = smem_u32(TF.b) + (uint32_t)TB_OFF - kMagicBits;
        const uint8_t* const in2 = stored ? xb : sbase;
        const int in2_pitch = stored ? xpitch : spitch;
        uint8_t* const obase = out_u8 ? out_u8 + (size_t)b * d.out_u8_stride : nullptr;
        void* onorm = nullptr;
        if (out_norm) onorm = (d.out_norm_dtype == SSAD_NORM_F32) ? (void*)(reinterpret_cast<float*>(out_norm) + (size_t)b * 3 * npix)
                                                                   : (void*)(reinterpret_cast<uint16_t*>(out_norm) + (size_t)b * 3 * npix);
        const uint8_t* lutbase = smem + g.o_lut;
        switch (loop_kind(D.fin)) {
            case LK_COPY:    pass2_loop<LK_COPY>(SL, in2, in2_pitch, !stored, R, D, BX, flags, W, npix, obase, onorm, d.out_norm_dtype, lutbase); break;
            case LK_BYTELUT: pass2_loop<LK_BYTELUT>(SL, in2, in2_pitch, !stored, R, D, BX, flags, W, npix, obase, onorm, d.out_norm_dtype, lutbase); break;
            case LK_SAT_UC:  pass2_loop<LK_SAT_UC>(SL, in2, in2_pitch, !stored, R, D, BX, flags, W, npix, obase, onorm, d.out_norm_dtype, lutbase); break;
            case LK_SAT_UL:  pass2_loop<LK_SAT_UL>(SL, in2, in2_pitch, !stored, R, D, BX, flags, W, npix, obase, onorm, d.out_norm_dtype, lutbase); break;
            case LK_SAT_LC:  pass2_loop<LK_SAT_LC>(SL, in2, in2_pitch, !stored, R, D, BX, flags, W, npix, obase, onorm, d.out_norm_dtype, lutbase); break;
            default:         pass2_loop<LK_SAT_LL>(SL, in2, in2_pitch, !stored, R, D, BX, flags, W, npix, obase, onorm, d.out_norm_dtype, lutbase); break;
        }
    }

    SSAD_MARK(5)
    #undef SSAD_MARK
    if (rank == 0 && tid == 0) {
        if (y_out) y_out[b] = y;
        if (debug) {
            SsadDebug o;
            memset(&o, 0, sizeof(o));
            o.n_coords = n_coords;
            o.label_out = y;
            o.contrast_mean = cmean;
            if (y != 0) {
                o.sim = D.sim;
                o.cosine = D.cosine;
                for (int k = 0; k < npaste; ++k) { o.paste[k][0] = D.paste[k][0]; o.paste[k][1] = D.paste[k][1]; }
                for (int k = 0; k < 3; ++k) o.avg_rgb[k] = D.avg_rgb[k];
            }
            debug[b] = o;
        }
    }
}

static int ilog2_exact(int v) {
    if (v <= 0 || (v & (v - 1))) return -1;
    int s = 0;
    while ((1 << s) < v) ++s;
    return s;
}

static bool make_geom(const SsadBatchDesc& d, Geom& g) {
    g.H = d.height; g.W = d.width;
    if (g.W % 8 != 0 || g.W < 8) return false;
    if (!(d.sat_factor_max <= kSatMax)) return false;       // the table after Saturation covers factors up to kSatMax
    g.upr = g.W / 8;
    g.upr_shift = ilog2_exact(g.upr);
    const long long units = (long long)g.upr * g.H;
    // cluster size: about 2048 units (8 per thread) per CTA, at most 8 CTAs (portable cluster size)
    int C = 1;
    while (C < MAXC && units / (2 * C) >= 2048 && g.H / (2 * C) >= 8) C *= 2;
    // ... but no more CTAs per image than it takes to give every SM several waves of work: the geometry phase is
    // repeated by every CTA of a cluster.  (148 SMs x 4 resident CTAs; a wave and a half.)
    while (C > 2 && (long long)d.batch * (C / 2) >= 3 * 148 * 2) C /= 2;
    if (const char* f = getenv("SSAD_V6_C")) { int v = atoi(f); if (v == 1 || v == 2 || v == 4 || v == 8) C = v; }
    g.C = C;
    g.rows_cta = (g.H + C - 1) / C;
    // canvas: polygon (pw x ph) or rotated scar (nw x nh); max_patch_* bound both
    g.cov_pitch = (uint32_t)(d.max_patch_w + 31) / 32u;
    g.cov_rows = (uint32_t)d.max_patch_h;
    uint32_t o = 0;
    g.o_rowpre = o; o += (((uint32_t)g.H + 1u) * 4u + 15u) & ~15u;
    g.o_cov = o;    o += (g.cov_rows * g.cov_pitch * 4u + 15u) & ~15u;
    g.o_flags = o;  o += ((uint32_t)g.rows_cta + 15u) & ~15u;
    g.o_lut = o;    o += (d.out_norm_dtype == SSAD_NORM_F32) ? 3072u : (d.out_norm_dtype == SSAD_NORM_BF16 ? 1536u : 0u);
    g.total = o;
    g.canvas_pitch = (uint32_t)d.max_patch_w;
    g.canvas_stride = (size_t)d.max_patch_w * (size_t)d.max_patch_h;
    g.o_canvas = ((size_t)d.batch * d.height * d.width * 3 + 255) / 256 * 256;       // after the x buffer
    g.ws_total = g.o_canvas + ((size_t)d.batch * g.canvas_stride * 4 + 255) / 256 * 256;
    return true;
}

}  // namespace v6
}  // namespace ssad

size_t ssad_v6_workspace(const SsadBatchDesc& d) {
    ssad::v6::Geom g;
    if (!ssad::v6::make_geom(d, g)) return 0;
    return g.ws_total;       // the x buffer + the patch canvases
}

// returns SSAD_OK when launched, SSAD_EUNSUPPORTED when the caller should use another kernel
int ssad_launch_v6(const SsadBatchDesc& d, const SsadImageParams* params, const uint8_t* src, const uint8_t* donor_bank,
                   const uint32_t* mask_bits, const void* norm_lut, uint8_t* out_u8, void* out_norm, int64_t* y_out,
                   SsadDebug* debug, void* workspace, size_t workspace_bytes, cudaStream_t stream) {
    using namespace ssad::v6;
    Geom g;
    if (!make_geom(d, g)) return SSAD_EUNSUPPORTED;
    if ((reinterpret_cast<uintptr_t>(src) & 15) || (d.src_pitch & 7) || (d.src_stride & 7)) return SSAD_EUNSUPPORTED;
    if (g.total > 96u * 1024u) return SSAD_EUNSUPPORTED;
    const size_t need = ssad_v6_workspace(d);
    if (!workspace || workspace_bytes < need) return SSAD_EWORKSPACE;
    if (reinterpret_cast<uintptr_t>(workspace) & 255) return SSAD_EALIGN;
    uint8_t* xbuf = static_cast<uint8_t*>(workspace);
    uint32_t* canvas = reinterpret_cast<uint32_t*>(xbuf + g.o_canvas);
    static int attr_smem = -1;
    cudaError_t e;
    if ((int)g.total > attr_smem) {
        e = cudaFuncSetAttribute(generate_v6_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)g.total);
        if (e != cudaSuccess) return ssad_set_cuda_error(e);
        attr_smem = (int)g.total;
    }
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = dim3((unsigned)((long long)d.batch * g.C), 1, 1);
    cfg.blockDim = dim3(NT, 1, 1);
    cfg.dynamicSmemBytes = g.total;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = (unsigned)g.C; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = (g.C > 1) ? 1 : 0;
    // SSAD_V6_PROF=1 (debugging aid): per-phase CTA clock sums, printed after a synchronising copy
    static unsigned long long* prof = nullptr;
    static const bool want_prof = getenv("SSAD_V6_PROF") != nullptr;
    if (want_prof && !prof) { cudaMalloc(&prof, 8 * sizeof(unsigned long long)); }
    if (want_prof) cudaMemsetAsync(prof, 0, 8 * sizeof(unsigned long long), stream);
    e = cudaLaunchKernelEx(&cfg, generate_v6_kernel, d, g, params, src, donor_bank, mask_bits, norm_lut, out_u8, out_norm,
                           y_out, debug, xbuf, canvas, want_prof ? prof : nullptr);
    if (e != cudaSuccess) return ssad_set_cuda_error(e);
    if (want_prof) {
        unsigned long long h[8];
        cudaMemcpyAsync(h, prof, sizeof(h), cudaMemcpyDeviceToHost, stream);
        cudaStreamSynchronize(stream);
        const double n = (double)d.batch * g.C;
        fprintf(stderr, "v6 phases (cycles per CTA): geometry %.0f  pass1 %.0f  exchange1 %.0f  decide+canvas %.0f  mean+tables %.0f  pass2 %.0f\n",
                h[0] / n, h[1] / n, h[2] / n, h[3] / n, h[4] / n, h[5] / n);
    }
    return SSAD_OK;
}
