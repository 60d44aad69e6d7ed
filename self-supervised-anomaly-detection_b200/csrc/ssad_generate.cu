// ssad_generate: PretextTaskDataset.__getitem__ (reference datasets.py:209-394) for a whole batch.
//
// One CTA per image; the whole working image (H*W*3 bytes, 192 KB at 256x256) stays resident in
// shared memory from the load to the final store, so HBM sees one read and one write per image.
// Stages (reference line numbers in the code):
//   load / RandomAffine gather -> mask-window prefix + centre selection -> RGB sums + cosine branch
//   -> polygon table + select-paste | rotated scar x k -> ColorJitter chain (contrast mean = block
//   reduction at its position in the permutation) -> uint8 HWC and/or normalised NCHW store.
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "ssad_device.cuh"

namespace ssad {

constexpr int NT = 512;          // threads per CTA
constexpr int kUnitPix = 16;     // pixels per thread-iteration: 48 B = 3 x 128-bit, channel aligned

struct Decisions {
    int y;
    int n_coords;
    int sim;
    int npaste;
    int paste[5][2];
    int flat;            // 1: patch is a flat colour (variants 1, 2)
    uint32_t flat_rgb[3];
    int cmean;
    float bf1, bf2;
    double cosine;
    uint32_t avg_rgb[3];
};

struct SmemLayout {
    uint32_t img, poly, rowpre, lut, red, par, dec, total;
    uint32_t poly_pitch;
};

__host__ __device__ inline uint32_t align16(uint32_t v) { return (v + 15u) & ~15u; }

__host__ __device__ inline SmemLayout smem_layout(const SsadBatchDesc& d) {
    SmemLayout L;
    uint32_t o = 0;
    L.img = o;    o += align16((uint32_t)d.height * d.width * 3u);
    L.poly_pitch = (uint32_t)(d.max_patch_w + 31) / 32u;
    L.poly = o;   o += align16((uint32_t)d.max_patch_h * L.poly_pitch * 4u);
    L.rowpre = o; o += align16(((uint32_t)d.height + 1u) * 4u);
    L.lut = o;    o += (d.out_norm_dtype == SSAD_NORM_F32) ? 3072u : (d.out_norm_dtype == SSAD_NORM_BF16 ? 1536u : 0u);
    L.red = o;    o += 64u * 4u;
    L.par = o;    o += align16((uint32_t)sizeof(SsadImageParams));
    L.dec = o;    o += align16((uint32_t)sizeof(Decisions));
    L.total = o;
    return L;
}

// ---------------------------------------------------------------- 16-pixel unit in magic form
struct Unit {
    float m[48];   // m[3 * p + c]
};

__device__ __forceinline__ void unit_unpack(Unit& u, const uint4& a, const uint4& b, const uint4& c) {
    const uint32_t w[12] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w, c.x, c.y, c.z, c.w};
    #pragma unroll
    for (int i = 0; i < 12; ++i) {
        #pragma unroll
        for (int k = 0; k < 4; ++k) u.m[4 * i + k] = magic_from_byte(w[i], k);
    }
}
__device__ __forceinline__ void unit_pack(const Unit& u, uint4& a, uint4& b, uint4& c) {
    uint32_t w[12];
    #pragma unroll
    for (int i = 0; i < 12; ++i) w[i] = magic_pack4(u.m[4 * i], u.m[4 * i + 1], u.m[4 * i + 2], u.m[4 * i + 3]);
    a = make_uint4(w[0], w[1], w[2], w[3]);
    b = make_uint4(w[4], w[5], w[6], w[7]);
    c = make_uint4(w[8], w[9], w[10], w[11]);
}

template <bool CLAMP>
__device__ __forceinline__ void unit_brightness(Unit& u, float f) {
    #pragma unroll
    for (int i = 0; i < 48; ++i) u.m[i] = magic_brightness<CLAMP>(u.m[i], f);
}
template <bool CLAMP>
__device__ __forceinline__ void unit_contrast(Unit& u, float f, int mean) {
    const float mA = magic_from_u8((uint32_t)mean), aF = (float)mean;
    #pragma unroll
    for (int i = 0; i < 48; ++i) u.m[i] = magic_blend<CLAMP>(u.m[i], mA, aF, f);
}
template <bool CLAMP>
__device__ __forceinline__ void unit_saturation(Unit& u, float f) {
    #pragma unroll
    for (int p = 0; p < 16; ++p) {
        uint32_t acc = luma_acc(u.m[3 * p], u.m[3 * p + 1], u.m[3 * p + 2]);
        float mL = __uint_as_float(__byte_perm(acc, kMagicBits, 0x7542));
        float lF = __fsub_rn(mL, kMagic);
        #pragma unroll
        for (int c = 0; c < 3; ++c) u.m[3 * p + c] = magic_blend<CLAMP>(u.m[3 * p + c], mL, lF, f);
    }
}
__device__ __forceinline__ uint32_t unit_luma_sum(const Unit& u) {
    uint32_t s = 0;
    #pragma unroll
    for (int p = 0; p < 16; ++p) s += luma_acc(u.m[3 * p], u.m[3 * p + 1], u.m[3 * p + 2]) >> 16;
    return s;
}
__device__ __forceinline__ void unit_apply(Unit& u, int op, float f) {
    // op: 0 brightness, 2 saturation (contrast is handled by the caller: it needs the mean)
    if (op == 0) {
        if (f > 1.0f) unit_brightness<true>(u, f); else unit_brightness<false>(u, f);
    } else {
        if (blend_needs_clip(f)) unit_saturation<true>(u, f); else unit_saturation<false>(u, f);
    }
}

// ---------------------------------------------------------------- the fused kernel
__global__ void __launch_bounds__(NT, 1)
generate_fused_kernel(const __grid_constant__ SsadBatchDesc d, const SsadImageParams* __restrict__ params,
                      const uint8_t* __restrict__ src, const uint8_t* __restrict__ donor_bank,
                      const uint32_t* __restrict__ mask_bits, const void* __restrict__ norm_lut,
                      uint8_t* __restrict__ out_u8, void* __restrict__ out_norm, int64_t* __restrict__ y_out,
                      SsadDebug* __restrict__ debug) {
    extern __shared__ __align__(16) uint8_t smem[];
    const SmemLayout L = smem_layout(d);
    uint8_t* img = smem + L.img;
    uint32_t* poly = reinterpret_cast<uint32_t*>(smem + L.poly);
    uint32_t* rowpre = reinterpret_cast<uint32_t*>(smem + L.rowpre);
    uint32_t* red = reinterpret_cast<uint32_t*>(smem + L.red);
    SsadImageParams& P = *reinterpret_cast<SsadImageParams*>(smem + L.par);
    Decisions& D = *reinterpret_cast<Decisions*>(smem + L.dec);

    const int tid = threadIdx.x;
    const int b = blockIdx.x;
    const int H = d.height, W = d.width;
    const int npix = H * W;
    const int nbytes = npix * 3;

    // ---- parameters -> smem
    if (tid < (int)(sizeof(SsadImageParams) / 4))
        reinterpret_cast<uint32_t*>(&P)[tid] = reinterpret_cast<const uint32_t*>(params + b)[tid];
    if (d.out_norm_dtype != SSAD_NORM_NONE && out_norm != nullptr) {
        const int nw = (d.out_norm_dtype == SSAD_NORM_F32) ? 768 : 384;
        for (int i = tid; i < nw; i += NT)
            reinterpret_cast<uint32_t*>(smem + L.lut)[i] = reinterpret_cast<const uint32_t*>(norm_lut)[i];
    }
    __syncthreads();

    const uint8_t* simg = src + (size_t)b * d.src_stride;
    const int cx0 = P.crop_off[0], cy0 = P.crop_off[1];

    // ---- load x: datasets.py:218-222 (copy / RandomAffine NEAREST, fill 0) and :242-251 (window)
    if (P.warp_en) {
        const int32_t a0 = P.warp_fix[0], a1 = P.warp_fix[1], a2 = P.warp_fix[2];
        const int32_t a3 = P.warp_fix[3], a4 = P.warp_fix[4], a5 = P.warp_fix[5];
        for (int p = tid; p < npix; p += NT) {
            int yo = p / W, xo = p - yo * W;
            int xin = (a2 + yo * a1 + xo * a0) >> 16;
            int yin = (a5 + yo * a4 + xo * a3) >> 16;
            uint32_t r = 0, g = 0, bl = 0;
            if (xin >= 0 && xin < W && yin >= 0 && yin < H) {
                const uint8_t* q = simg + (size_t)(cy0 + yin) * d.src_pitch + (size_t)(cx0 + xin) * 3;
                r = __ldg(q); g = __ldg(q + 1); bl = __ldg(q + 2);
            }
            img[3 * p] = (uint8_t)r; img[3 * p + 1] = (uint8_t)g; img[3 * p + 2] = (uint8_t)bl;
        }
    } else {
        const int rowb = W * 3;
        const uint8_t* base = simg + (size_t)cy0 * d.src_pitch + (size_t)cx0 * 3;
        const bool vec = ((rowb & 15) == 0) && ((d.src_pitch & 15) == 0) && ((reinterpret_cast<uintptr_t>(base) & 15) == 0);
        if (vec) {
            const int vpr = rowb >> 4;
            for (int i = tid; i < H * vpr; i += NT) {
                int r = i / vpr, c = i - r * vpr;
                reinterpret_cast<uint4*>(img)[i] = __ldg(reinterpret_cast<const uint4*>(base + (size_t)r * d.src_pitch) + c);
            }
        } else {
            for (int i = tid; i < nbytes; i += NT) {
                int r = i / rowb, c = i - r * rowb;
                img[i] = __ldg(base + (size_t)r * d.src_pitch + c);
            }
        }
    }

    // ---- mask window: per-row popcounts (coords_map is row-major, datasets.py:263-264)
    const uint32_t* mrow0 = nullptr;
    if (d.mask_mode != SSAD_MASK_ALL)
        mrow0 = mask_bits + (d.mask_mode == SSAD_MASK_PER_IMAGE ? (size_t)b * d.mask_stride_words : 0);
    const bool defect = P.label != 0;
    if (defect && d.mask_mode != SSAD_MASK_ALL) {
        for (int r = tid; r < H; r += NT)
            rowpre[r + 1] = row_window_popc(mrow0 + (size_t)(cy0 + r) * d.mask_pitch_words, cx0, W);
    }
    __syncthreads();
    if (defect && d.mask_mode != SSAD_MASK_ALL && tid < 32) {      // exclusive prefix by warp 0
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

    // ---- label after the coverage rule (datasets.py:258-259)
    int n_coords = (d.mask_mode == SSAD_MASK_ALL) ? npix : (defect ? (int)rowpre[H] : 0);
    int y = P.label;
    if (defect && d.coverage_min3 > 0 && 3 * n_coords < d.coverage_min3) y = 0;

    // donor image (datasets.py:225-228, 253)
    const uint8_t* dimg = simg;
    int64_t dpitch = d.src_pitch;
    int dH = d.src_height, dW = d.src_width;
    if (P.donor_idx >= 0) {
        dimg = donor_bank + (size_t)P.donor_idx * d.donor_stride;
        dpitch = d.donor_pitch; dH = d.donor_height; dW = d.donor_width;
    }
    const int rl = P.rect[0], rt = P.rect[1], pw = P.rect[2], ph = P.rect[3];
    const int dox = P.donor_off[0], doy = P.donor_off[1];
    auto donor_px = [&](int u, int v, uint32_t& r, uint32_t& g, uint32_t& bl) {
        // patch pixel (u, v) = donor window pixel (rl+u, rt+v); outside the window PIL's crop pads with 0
        int wx = rl + u, wy = rt + v;
        r = g = bl = 0;
        if (wx >= 0 && wy >= 0 && wx < d.donor_win_w && wy < d.donor_win_h) {
            int gx = dox + wx, gy = doy + wy;
            if (gx < dW && gy < dH) {
                const uint8_t* q = dimg + (size_t)gy * dpitch + (size_t)gx * 3;
                r = __ldg(q); g = __ldg(q + 1); bl = __ldg(q + 2);
            }
        }
    };

    if (y != 0) {
        // ---- check_color_similarity (dataset_generator.py:147-159): exact integer channel sums
        uint32_t sx[3] = {0, 0, 0};
        {
            const uint32_t* w = reinterpret_cast<const uint32_t*>(img);
            const int ngrp = nbytes / 12;                       // 4 pixels = 3 words
            for (int i = tid; i < ngrp; i += NT) {
                uint32_t w0 = w[3 * i], w1 = w[3 * i + 1], w2 = w[3 * i + 2];
                sx[0] += __dp4a(w0, 0x01000001u, 0u) + __dp4a(w1, 0x00010000u, 0u) + __dp4a(w2, 0x00000100u, 0u);
                sx[1] += __dp4a(w0, 0x00000100u, 0u) + __dp4a(w1, 0x01000001u, 0u) + __dp4a(w2, 0x00010000u, 0u);
                sx[2] += __dp4a(w0, 0x00010000u, 0u) + __dp4a(w1, 0x00000100u, 0u) + __dp4a(w2, 0x01000001u, 0u);
            }
            for (int p = ngrp * 4 + tid; p < npix; p += NT) {   // tail pixels (npix % 4)
                sx[0] += img[3 * p]; sx[1] += img[3 * p + 1]; sx[2] += img[3 * p + 2];
            }
        }
        uint32_t sp[3] = {0, 0, 0};
        if (P.src_variant != 2) {
            for (int i = tid; i < pw * ph; i += NT) {
                int v = i / pw, u = i - v * pw;
                uint32_t r, g, bl;
                donor_px(u, v, r, g, bl);
                sp[0] += r; sp[1] += g; sp[2] += bl;
            }
        }
        uint32_t tot[6];
        #pragma unroll
        for (int c = 0; c < 3; ++c) tot[c] = block_sum_u32<NT>(sx[c], red);
        #pragma unroll
        for (int c = 0; c < 3; ++c) tot[3 + c] = block_sum_u32<NT>(sp[c], red);

        if (tid == 0) {
            D.n_coords = n_coords;
            const uint32_t np = (uint32_t)(pw * ph);
            uint32_t psum[3] = {tot[3], tot[4], tot[5]};
            D.flat = 0;
            D.avg_rgb[0] = D.avg_rgb[1] = D.avg_rgb[2] = 0;
            if (P.src_variant == 1) {           // colour = int(mean) (dataset_generator.py:202-207)
                D.flat = 1;
                for (int c = 0; c < 3; ++c) { D.flat_rgb[c] = psum[c] / np; D.avg_rgb[c] = D.flat_rgb[c]; psum[c] = D.flat_rgb[c] * np; }
            } else if (P.src_variant == 2) {    // random colour (dataset_generator.py:194-199)
                D.flat = 1;
                for (int c = 0; c < 3; ++c) { D.flat_rgb[c] = P.rnd_rgb[c]; psum[c] = D.flat_rgb[c] * np; }
            }
            double cs = cosine_of_means(tot, (double)npix, psum, (double)np);
            D.cosine = cs;
            D.sim = cs > 0.99;                  // datasets.py:295 / 335
            D.bf1 = P.bright_f[0]; D.bf2 = P.bright_f[1];
            if (D.flat && D.sim) {              // brightness of a flat patch is a flat patch
                for (int c = 0; c < 3; ++c) D.flat_rgb[c] = brightness_u8(brightness_u8(D.flat_rgb[c], D.bf1), D.bf2);
            }
            D.npaste = (y == 1) ? 1 : P.scar_k;
        }
        __syncthreads();
        // ---- centres + container clamp (dataset_generator.py:104-144, 259-265), one thread per paste
        if (tid < D.npaste) {
            int cx, cy;
            if (n_coords == 0) {
                cx = P.fallback_c[tid][0]; cy = P.fallback_c[tid][1];
            } else {
                uint32_t idx = (n_coords < 2) ? 0u : (uint32_t)(((uint64_t)P.centre_r[tid] * (uint64_t)n_coords) >> 32);
                if (d.mask_mode == SSAD_MASK_ALL) {
                    cy = idx / W; cx = idx - cy * W;
                } else {
                    int lo = 0, hi = H;          // largest r with rowpre[r] <= idx
                    while (hi - lo > 1) { int mid = (lo + hi) >> 1; if (rowpre[mid] <= idx) lo = mid; else hi = mid; }
                    cy = lo;
                    cx = row_window_select(mrow0 + (size_t)(cy0 + cy) * d.mask_pitch_words, cx0, W, idx - rowpre[lo]);
                }
            }
            int sw = (y == 1) ? pw : P.scar_size[0], sh = (y == 1) ? ph : P.scar_size[1];
            int pl, pt;
            clamp_to_container(y == 1 ? d.box_poly : d.box_scar, sw, sh, cx, cy, pl, pt);
            D.paste[tid][0] = pl; D.paste[tid][1] = pt;
        }
        // ---- rect2poly table (dataset_generator.py:63-100): one thread per canvas row
        if (y == 1) {
            Polygon G;
            polygon_setup(G, P.poly_xy, P.poly_n, ph);
            for (int r = tid; r < ph; r += NT) {
                uint32_t* row = poly + (size_t)r * L.poly_pitch;
                for (int j = 0; j < (int)L.poly_pitch; ++j) row[j] = 0;
                polygon_fill_row(G, r, pw, row);
            }
        }
        __syncthreads();

        const bool sim = D.sim != 0, flat = D.flat != 0;
        const float bf1 = D.bf1, bf2 = D.bf2;
        auto patch_colour = [&](int u, int v, uint32_t& r, uint32_t& g, uint32_t& bl) {
            if (flat) { r = D.flat_rgb[0]; g = D.flat_rgb[1]; bl = D.flat_rgb[2]; return; }
            donor_px(u, v, r, g, bl);
            if (sim) {      // two successive ImageEnhance.Brightness, each re-quantised (datasets.py:298-299)
                r = brightness_u8(brightness_u8(r, bf1), bf2);
                g = brightness_u8(brightness_u8(g, bf1), bf2);
                bl = brightness_u8(brightness_u8(bl, bf1), bf2);
            }
        };
        if (y == 1) {
            // ---- paste_patch(x, patch, coords, mask) (dataset_generator.py:268-275)
            const int pl = D.paste[0][0], pt = D.paste[0][1];
            for (int i = tid; i < pw * ph; i += NT) {
                int v = i / pw, u = i - v * pw;
                int X = pl + u, Y = pt + v;
                if (X < 0 || Y < 0 || X >= W || Y >= H) continue;
                if (!((poly[(size_t)v * L.poly_pitch + (u >> 5)] >> (u & 31)) & 1u)) continue;
                uint32_t r, g, bl;
                patch_colour(u, v, r, g, bl);
                uint8_t* q = img + 3 * (Y * W + X);
                q[0] = (uint8_t)r; q[1] = (uint8_t)g; q[2] = (uint8_t)bl;
            }
        } else {
            // ---- scar: rotate(angle, expand=True) NEAREST evaluated at paste time, k pastes in order
            //      (datasets.py:340-355); later pastes overwrite earlier ones
            const int nw = P.scar_size[0], nh = P.scar_size[1];
            for (int k = 0; k < D.npaste; ++k) {
                const int pl = D.paste[k][0], pt = D.paste[k][1];
                for (int i = tid; i < nw * nh; i += NT) {
                    int v = i / nw, u = i - v * nw;
                    int X = pl + u, Y = pt + v;
                    if (X < 0 || Y < 0 || X >= W || Y >= H) continue;
                    int xin, yin;
                    affine_fixed_map(P.scar_fix, u, v, xin, yin);
                    if (xin < 0 || yin < 0 || xin >= pw || yin >= ph) continue;     // alpha 0 outside the source
                    uint32_t r, g, bl;
                    patch_colour(xin, yin, r, g, bl);
                    uint8_t* q = img + 3 * (Y * W + X);
                    q[0] = (uint8_t)r; q[1] = (uint8_t)g; q[2] = (uint8_t)bl;
                }
                __syncthreads();
            }
        }
    }
    __syncthreads();

    // ---- CPP.jitter_transforms (datasets.py:391): ColorJitter ops in fn_idx order
    int ops[3];
    float fac[3];
    int nops = 0, cpos = -1;
    #pragma unroll
    for (int s = 0; s < 4; ++s) {
        int fn = P.jit_order[s];
        if (fn < 3 && ((P.jit_en >> fn) & 1)) {
            if (fn == 1) cpos = nops;
            ops[nops] = fn; fac[nops] = P.jit_f[fn]; ++nops;
        }
    }
    const int nunits = npix / kUnitPix;
    uint4* img4 = reinterpret_cast<uint4*>(img);
    int cmean = -1;
    if (cpos >= 0) {
        // pass A: ops before contrast, in place, + sum of L over the image they produce
        uint32_t lsum = 0;
        for (int uidx = tid; uidx < nunits; uidx += NT) {
            Unit u;
            unit_unpack(u, img4[3 * uidx], img4[3 * uidx + 1], img4[3 * uidx + 2]);
            if (cpos > 0) {
                unit_apply(u, ops[0], fac[0]);
                if (cpos > 1) unit_apply(u, ops[1], fac[1]);
                unit_pack(u, img4[3 * uidx], img4[3 * uidx + 1], img4[3 * uidx + 2]);
            }
            lsum += unit_luma_sum(u);
        }
        lsum = block_sum_u32<NT>(lsum, red);
        // int(mean + 0.5) with mean = sum / count in float64 (ImageEnhance.py Contrast) == integer rounding
        cmean = (int)((2ull * lsum + (uint32_t)npix) / (2ull * (uint32_t)npix));
    }
    // pass B: contrast + the remaining ops, then the stores
    const float cf = (cpos >= 0) ? fac[cpos] : 1.0f;
    const bool cclip = blend_needs_clip(cf);
    const int first_post = (cpos >= 0) ? cpos + 1 : 0;
    uint8_t* obase = out_u8 ? out_u8 + (size_t)b * d.out_u8_stride : nullptr;
    for (int uidx = tid; uidx < nunits; uidx += NT) {
        Unit u;
        unit_unpack(u, img4[3 * uidx], img4[3 * uidx + 1], img4[3 * uidx + 2]);
        if (cpos >= 0) {
            if (cclip) unit_contrast<true>(u, cf, cmean); else unit_contrast<false>(u, cf, cmean);
        }
        if (first_post < nops) {
            unit_apply(u, ops[first_post], fac[first_post]);
            if (first_post + 1 < nops) {
                unit_apply(u, ops[first_post + 1], fac[first_post + 1]);
                if (first_post + 2 < nops) unit_apply(u, ops[first_post + 2], fac[first_post + 2]);
            }
        }
        if (obase) {
            uint4 a, bq, c;
            unit_pack(u, a, bq, c);
            uint4* o = reinterpret_cast<uint4*>(obase) + 3 * (size_t)uidx;
            o[0] = a; o[1] = bq; o[2] = c;
        }
        if (out_norm) {
            // ToTensor + Normalize (datasets.py:392-393): exact table per channel, NCHW planes
            const size_t p0 = (size_t)uidx * kUnitPix;
            if (d.out_norm_dtype == SSAD_NORM_F32) {
                const float* lut = reinterpret_cast<const float*>(smem + L.lut);
                float* ob = reinterpret_cast<float*>(out_norm) + (size_t)b * 3 * npix;
                #pragma unroll
                for (int c = 0; c < 3; ++c) {
                    float4* o = reinterpret_cast<float4*>(ob + (size_t)c * npix + p0);
                    #pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        float4 v;
                        v.x = lut[c * 256 + magic_to_u8(u.m[3 * (4 * q + 0) + c])];
                        v.y = lut[c * 256 + magic_to_u8(u.m[3 * (4 * q + 1) + c])];
                        v.z = lut[c * 256 + magic_to_u8(u.m[3 * (4 * q + 2) + c])];
                        v.w = lut[c * 256 + magic_to_u8(u.m[3 * (4 * q + 3) + c])];
                        o[q] = v;
                    }
                }
            } else {
                const uint16_t* lut = reinterpret_cast<const uint16_t*>(smem + L.lut);
                uint16_t* ob = reinterpret_cast<uint16_t*>(out_norm) + (size_t)b * 3 * npix;
                #pragma unroll
                for (int c = 0; c < 3; ++c) {
                    uint4* o = reinterpret_cast<uint4*>(ob + (size_t)c * npix + p0);
                    uint32_t w[8];
                    #pragma unroll
                    for (int q = 0; q < 8; ++q) {
                        uint32_t lo = lut[c * 256 + magic_to_u8(u.m[3 * (2 * q) + c])];
                        uint32_t hi = lut[c * 256 + magic_to_u8(u.m[3 * (2 * q + 1) + c])];
                        w[q] = lo | (hi << 16);
                    }
                    o[0] = make_uint4(w[0], w[1], w[2], w[3]);
                    o[1] = make_uint4(w[4], w[5], w[6], w[7]);
                }
            }
        }
    }

    if (tid == 0) {
        if (y_out) y_out[b] = y;
        if (debug) {
            SsadDebug g;
            memset(&g, 0, sizeof(g));
            g.n_coords = n_coords;
            g.label_out = y;
            g.contrast_mean = cmean;
            if (y != 0) {
                g.sim = D.sim;
                g.cosine = D.cosine;
                for (int k = 0; k < D.npaste; ++k) { g.paste[k][0] = D.paste[k][0]; g.paste[k][1] = D.paste[k][1]; }
                for (int c = 0; c < 3; ++c) g.avg_rgb[c] = (int)D.avg_rgb[c];
            }
            debug[b] = g;
        }
    }
}

}  // namespace ssad

// ------------------------------------------------------------------ host side of the C ABI
static thread_local char g_cuda_err[256] = "";

extern "C" int ssad_version(void) { return SSAD_VERSION; }
extern "C" size_t ssad_sizeof_params(void) { return sizeof(SsadImageParams); }
extern "C" size_t ssad_sizeof_desc(void) { return sizeof(SsadBatchDesc); }
extern "C" size_t ssad_sizeof_debug(void) { return sizeof(SsadDebug); }
extern "C" const char* ssad_last_cuda_error(void) { return g_cuda_err; }

extern "C" const char* ssad_strerror(int code) {
    switch (code) {
        case SSAD_OK: return "ok";
        case SSAD_EINVAL: return "invalid argument";
        case SSAD_EALIGN: return "pointer or pitch not 16-byte aligned";
        case SSAD_EUNSUPPORTED: return "unsupported configuration for this build";
        case SSAD_ELAUNCH: return "CUDA launch failed";
        case SSAD_EWORKSPACE: return "workspace too small";
        default: return "unknown error";
    }
}

int ssad_launch_cluster(const SsadBatchDesc& d, const SsadImageParams* params, const uint8_t* src, const uint8_t* donor_bank,
                        const uint32_t* mask_bits, const void* norm_lut, uint8_t* out_u8, void* out_norm, int64_t* y_out,
                        SsadDebug* debug, cudaStream_t stream);

int ssad_launch_v3(const SsadBatchDesc& d, const SsadImageParams* params, const uint8_t* src, const uint8_t* donor_bank,
                   const uint32_t* mask_bits, const void* norm_lut, uint8_t* out_u8, void* out_norm, int64_t* y_out,
                   SsadDebug* debug, cudaStream_t stream);

int ssad_launch_v4(const SsadBatchDesc& d, const SsadImageParams* params, const uint8_t* src, const uint8_t* donor_bank,
                   const uint32_t* mask_bits, const void* norm_lut, uint8_t* out_u8, void* out_norm, int64_t* y_out,
                   SsadDebug* debug, void* workspace, size_t workspace_bytes, cudaStream_t stream);
size_t ssad_v4_workspace(const SsadBatchDesc& d);

int ssad_launch_v5(const SsadBatchDesc& d, const SsadImageParams* params, const uint8_t* src, const uint8_t* donor_bank,
                   const uint32_t* mask_bits, const void* norm_lut, uint8_t* out_u8, void* out_norm, int64_t* y_out,
                   SsadDebug* debug, void* workspace, size_t workspace_bytes, cudaStream_t stream);
size_t ssad_v5_workspace(const SsadBatchDesc& d);

int ssad_launch_v6(const SsadBatchDesc& d, const SsadImageParams* params, const uint8_t* src, const uint8_t* donor_bank,
                   const uint32_t* mask_bits, const void* norm_lut, uint8_t* out_u8, void* out_norm, int64_t* y_out,
                   SsadDebug* debug, void* workspace, size_t workspace_bytes, cudaStream_t stream);
size_t ssad_v6_workspace(const SsadBatchDesc& d);
// v7 (pipeline of five kernels: plan / pass 1 / canvas / finish / pass 2, ssad_generate_v7.cu)
int ssad_launch_v7(const SsadBatchDesc& d, const SsadImageParams* params, const uint8_t* src, const uint8_t* donor_bank,
                   const uint32_t* mask_bits, const void* norm_lut, uint8_t* out_u8, void* out_norm, int64_t* y_out,
                   SsadDebug* debug, void* workspace, size_t workspace_bytes, cudaStream_t stream);
size_t ssad_v7_workspace(const SsadBatchDesc& d);
int ssad_v7_launches(const SsadBatchDesc& d, bool capturing);

int ssad_set_cuda_error(cudaError_t e) {
    strncpy(g_cuda_err, cudaGetErrorString(e), sizeof(g_cuda_err) - 1);
    return SSAD_ELAUNCH;
}

extern "C" size_t ssad_generate_workspace(const SsadBatchDesc* desc) {
    if (!desc || desc->batch <= 0) return 0;
    size_t a = ssad_v5_workspace(*desc), b = ssad_v4_workspace(*desc), c = ssad_v6_workspace(*desc);
    if (b > a) a = b;
    if (c > a) a = c;
    c = ssad_v7_workspace(*desc);
    return c > a ? c : a;               // 0 for shapes only the on-chip kernels serve (they need no workspace)
}

extern "C" int ssad_generate_launches(const SsadBatchDesc* desc, int captured) {
    if (!desc || desc->batch <= 0) return 0;
    const char* force = getenv("SSAD_KERNEL");
    if ((!force || strcmp(force, "v7") == 0) && ssad_v7_workspace(*desc) > 0) return ssad_v7_launches(*desc, captured != 0);
    if (force && strcmp(force, "v4") == 0) return 7;
    return 1;
}

extern "C" int ssad_generate(const SsadBatchDesc* desc, const SsadImageParams* params, const uint8_t* src,
                             const uint8_t* donor_bank, const uint32_t* mask_bits, const void* norm_lut,
                             uint8_t* out_u8, void* out_norm, int64_t* y_out, SsadDebug* debug,
                             void* workspace, size_t workspace_bytes, void* stream) {
    if (!desc || !params || !src) return SSAD_EINVAL;
    const SsadBatchDesc& d = *desc;
    if (d.batch < 0 || d.height <= 0 || d.width <= 0 || d.src_pitch < 3ll * d.width) return SSAD_EINVAL;
    if (d.batch == 0) return SSAD_OK;
    if (d.mask_mode < SSAD_MASK_ALL || d.mask_mode > SSAD_MASK_PER_IMAGE) return SSAD_EINVAL;
    if (d.mask_mode != SSAD_MASK_ALL && !mask_bits) return SSAD_EINVAL;
    if (d.out_norm_dtype < SSAD_NORM_NONE || d.out_norm_dtype > SSAD_NORM_BF16) return SSAD_EINVAL;
    if (out_norm && (d.out_norm_dtype == SSAD_NORM_NONE || !norm_lut)) return SSAD_EINVAL;
    if (d.max_patch_w <= 0 || d.max_patch_h <= 0) return SSAD_EINVAL;
    if ((reinterpret_cast<uintptr_t>(params) & 15) || (reinterpret_cast<uintptr_t>(out_u8) & 15) ||
        (reinterpret_cast<uintptr_t>(out_norm) & 15))
        return SSAD_EALIGN;
    if (out_u8 && (d.out_u8_pitch != 3ll * d.width || d.out_u8_stride != 3ll * d.width * d.height)) return SSAD_EUNSUPPORTED;
    if (((long long)d.height * d.width) % ssad::kUnitPix != 0) return SSAD_EUNSUPPORTED;
    // B200 paths, in order of preference: v4 (pipeline of barrier-free streaming kernels, ssad_generate_v4.cu), v3
    // (clusters of 1-8 CTAs, rows resident or streamed, ssad_generate_v3.cu), the 2-CTA cluster kernel
    // (ssad_generate_cluster.cu) and the generic one-CTA-per-image kernel below for shapes the others cannot express.
    // SSAD_KERNEL=v4|v3|cluster|generic pins one (tests, profiling).
    const char* force = getenv("SSAD_KERNEL");
    const bool only_v4 = force && strcmp(force, "v4") == 0, only_v5 = force && strcmp(force, "v5") == 0;
    const bool only_v6 = force && strcmp(force, "v6") == 0, only_v7 = force && strcmp(force, "v7") == 0;
    if (!force || only_v7) {
        int rc = ssad_launch_v7(d, params, src, donor_bank, mask_bits, norm_lut, out_u8, out_norm, y_out, debug, workspace,
                                workspace_bytes, static_cast<cudaStream_t>(stream));
        if (rc != SSAD_EUNSUPPORTED || only_v7) return rc;
    }
    if (!force || only_v6) {
        int rc = ssad_launch_v6(d, params, src, donor_bank, mask_bits, norm_lut, out_u8, out_norm, y_out, debug, workspace,
                                workspace_bytes, static_cast<cudaStream_t>(stream));
        if (rc != SSAD_EUNSUPPORTED || only_v6) return rc;
    }
    if (!force || only_v5) {
        int rc = ssad_launch_v5(d, params, src, donor_bank, mask_bits, norm_lut, out_u8, out_norm, y_out, debug, workspace,
                                workspace_bytes, static_cast<cudaStream_t>(stream));
        if (rc != SSAD_EUNSUPPORTED || only_v5) return rc;
    }
    const bool only_v3 = force && strcmp(force, "v3") == 0, only_cluster = force && strcmp(force, "cluster") == 0;
    const bool only_generic = force && strcmp(force, "generic") == 0;
    if (d.batch > 65535) return SSAD_EUNSUPPORTED;
    if (!only_v3 && !only_cluster && !only_generic) {
        int rc = ssad_launch_v4(d, params, src, donor_bank, mask_bits, norm_lut, out_u8, out_norm, y_out, debug, workspace,
                                workspace_bytes, static_cast<cudaStream_t>(stream));
        if (rc != SSAD_EUNSUPPORTED || only_v4) return rc;
    }
    if (!only_cluster && !only_generic) {
        int rc = ssad_launch_v3(d, params, src, donor_bank, mask_bits, norm_lut, out_u8, out_norm, y_out, debug,
                                static_cast<cudaStream_t>(stream));
        if (rc != SSAD_EUNSUPPORTED || only_v3) return rc;
    }
    if (!only_generic) {
        int rc = ssad_launch_cluster(d, params, src, donor_bank, mask_bits, norm_lut, out_u8, out_norm, y_out, debug,
                                     static_cast<cudaStream_t>(stream));
        if (rc != SSAD_EUNSUPPORTED || only_cluster) return rc;
    }
    const ssad::SmemLayout L = ssad::smem_layout(d);
    int dev = 0, max_smem = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return ssad_set_cuda_error(e);
    e = cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
    if (e != cudaSuccess) return ssad_set_cuda_error(e);
    if ((long long)L.total > max_smem) return SSAD_EUNSUPPORTED;   // image does not fit one SM (tiled path: next row)
    e = cudaFuncSetAttribute(ssad::generate_fused_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)L.total);
    if (e != cudaSuccess) return ssad_set_cuda_error(e);
    ssad::generate_fused_kernel<<<d.batch, ssad::NT, L.total, static_cast<cudaStream_t>(stream)>>>(
        d, params, src, donor_bank, mask_bits, norm_lut, out_u8, out_norm, y_out, debug);
    e = cudaGetLastError();
    if (e != cudaSuccess) return ssad_set_cuda_error(e);
    return SSAD_OK;
}
