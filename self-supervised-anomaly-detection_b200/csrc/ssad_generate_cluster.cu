// ssad_generate, B200 path: persistent 2-CTA clusters, one image per cluster at a time.
//
//   * each CTA owns half the rows of the working image, fetched by TMA (cp.async.bulk.tensor over a
//     3-D u32 tensor map [batch][rows][row words], one box per half image) while the CTA takes the
//     pixel-independent decisions (mask prefix, centres, container clamp, polygon table).
//     TMA box starts must be 16-byte aligned, a crop window starts at any byte: the box covers the
//     aligned superset (row bytes + 16) and a register-staged pass shifts the rows down to a dense
//     row-major layout (skipped when every window of the batch is aligned);
//   * whole-image reductions (RGB sums for check_color_similarity, sum of L for the contrast mean)
//     are exchanged between the two CTAs through distributed shared memory + barrier.cluster;
//   * the ColorJitter chain runs on 16-pixel register units in packed f32x2 arithmetic
//     (FADD2 / FMUL2 / FADD2.RZ), bit-identical to Pillow's separate float32 multiply and add;
//   * 148 clusters x 2 CTAs of 256 threads, TWO CTAs resident per SM (2 x ~110 KB shared memory): while one CTA
//     waits for its TMA fetch or runs a serial decision phase, its co-resident runs the jitter arithmetic;
//     images are dealt round-robin, so every SM pair gets 7 (or 6) of 512 images: 1.2 % tail.
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <string.h>

#include "ssad_device.cuh"
#include "ssad_sm100.cuh"

int ssad_set_cuda_error(cudaError_t e);

namespace ssad {

constexpr int NT2 = 256;

struct ClusterGeom {
    int H, W, Hh;              // image rows / cols, rows per CTA
    int row_bytes;             // 3 * W
    int raw_pitch;             // bytes per row as landed by TMA: row_bytes (aligned batch) or row_bytes + 16
    int cpr;                   // 16-byte chunks per row
    int units_half;            // Hh * W / 16
    uint32_t buf_bytes;        // Hh * row_bytes (dense)
    uint32_t raw_bytes;        // Hh * raw_pitch (TMA transaction size)
    uint32_t poly_pitch, poly_rows;
    // smem offsets
    uint32_t o_buf, o_poly, o_rowpre, o_lut, o_red, o_par, o_dec, o_xchg, o_bar, o_pgn, total;
};

struct Decisions2 {
    int sim, npaste, flat, cmean;
    int paste[5][2];
    uint32_t flat_rgb[3];
    uint32_t avg_rgb[3];
    float bf1, bf2;
    double cosine;
};

static inline uint32_t up(uint32_t v, uint32_t a) { return (v + a - 1) / a * a; }

static bool make_geom(const SsadBatchDesc& d, bool aligned, ClusterGeom& g) {
    g.H = d.height; g.W = d.width;
    if (g.W % 16 != 0 || g.H % 2 != 0) return false;
    g.Hh = g.H / 2;
    g.row_bytes = g.W * 3;
    g.raw_pitch = aligned ? g.row_bytes : g.row_bytes + 16;
    if (g.raw_pitch / 4 > 256 || g.Hh > 256) return false;          // one TMA box: <= 256 elements per dimension
    g.cpr = g.row_bytes / 16;
    g.units_half = g.Hh * g.W / 16;
    g.buf_bytes = (uint32_t)g.Hh * g.row_bytes;
    g.raw_bytes = (uint32_t)g.Hh * g.raw_pitch;
    if (g.raw_bytes >= (1u << 20)) return false;                    // mbarrier tx-count range
    g.poly_pitch = (uint32_t)(d.max_patch_w + 31) / 32u;
    g.poly_rows = (uint32_t)(d.max_patch_h < g.Hh ? d.max_patch_h : g.Hh);   // only rows landing in my half
    uint32_t o = 0;
    g.o_buf = o;    o += up(g.raw_bytes, 128);
    g.o_poly = o;   o += up(g.poly_rows * g.poly_pitch * 4u, 16);
    g.o_rowpre = o; o += up(((uint32_t)g.H + 1u) * 4u, 16);
    g.o_lut = o;    o += (d.out_norm_dtype == SSAD_NORM_F32) ? 3072u : (d.out_norm_dtype == SSAD_NORM_BF16 ? 1536u : 0u);
    g.o_red = o;    o += 16u * 8u * 4u + 64u;
    g.o_par = o;    o += up((uint32_t)sizeof(SsadImageParams), 16);
    g.o_dec = o;    o += up((uint32_t)sizeof(Decisions2), 16);
    g.o_xchg = o;   o += 2u * 8u * 4u;
    g.o_bar = o;    o += 32u;
    g.o_pgn = o;    o += up((uint32_t)sizeof(Polygon), 16);
    g.total = o;
    return true;
}

// ---------------------------------------------------------------- 16-pixel unit, packed pairs
// v[c][k] = (pixel 2k, pixel 2k+1) of channel c in magic form (2^23 + value)
struct Unit2 {
    f32x2 v[3][8];
};

__device__ __forceinline__ void unit2_unpack(Unit2& u, const uint4& a, const uint4& b, const uint4& c) {
    const uint32_t w[12] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w, c.x, c.y, c.z, c.w};
    #pragma unroll
    for (int k = 0; k < 8; ++k) {
        #pragma unroll
        for (int ch = 0; ch < 3; ++ch) {
            const int b0 = 6 * k + ch, b1 = 6 * k + 3 + ch;
            u.v[ch][k] = pack2(magic_from_byte(w[b0 >> 2], b0 & 3), magic_from_byte(w[b1 >> 2], b1 & 3));
        }
    }
}
__device__ __forceinline__ void unit2_scalars(const Unit2& u, float (&m)[48]) {
    #pragma unroll
    for (int k = 0; k < 8; ++k) {
        #pragma unroll
        for (int ch = 0; ch < 3; ++ch) unpack2(u.v[ch][k], m[6 * k + ch], m[6 * k + 3 + ch]);
    }
}
__device__ __forceinline__ void unit2_pack(const Unit2& u, uint4& a, uint4& b, uint4& c) {
    float m[48];
    unit2_scalars(u, m);
    uint32_t w[12];
    #pragma unroll
    for (int i = 0; i < 12; ++i) w[i] = magic_pack4(m[4 * i], m[4 * i + 1], m[4 * i + 2], m[4 * i + 3]);
    a = make_uint4(w[0], w[1], w[2], w[3]);
    b = make_uint4(w[4], w[5], w[6], w[7]);
    c = make_uint4(w[8], w[9], w[10], w[11]);
}

template <bool CLAMP>
__device__ __forceinline__ f32x2 quantize2(f32x2 t) {
    if (CLAMP) {
        float lo, hi;
        unpack2(t, lo, hi);
        t = pack2(fminf(fmaxf(lo, 0.0f), 255.0f), fminf(fmaxf(hi, 0.0f), 255.0f));
    }
    return add2_rz(t, pack2(kMagic, kMagic));
}
template <bool CLAMP>
__device__ __forceinline__ void unit2_brightness(Unit2& u, float f) {
    const f32x2 F = pack2(f, f), NEG = pack2(-kMagic, -kMagic), MAG = pack2(kMagic, kMagic);
    #pragma unroll
    for (int ch = 0; ch < 3; ++ch) {
        #pragma unroll
        for (int k = 0; k < 8; ++k) {
            f32x2 t = mul2_rn(F, add2_rn(u.v[ch][k], NEG));
            if (CLAMP) {                                   // values >= 0: only the upper clip can bind
                float lo, hi;
                unpack2(t, lo, hi);
                t = pack2(fminf(lo, 255.0f), fminf(hi, 255.0f));
            }
            u.v[ch][k] = add2_rz(t, MAG);
        }
    }
}
template <bool CLAMP>
__device__ __forceinline__ void unit2_contrast(Unit2& u, float f, int mean) {
    const float nm = -(kMagic + (float)mean), af = (float)mean;
    const f32x2 F = pack2(f, f), NMA = pack2(nm, nm), AF = pack2(af, af);
    #pragma unroll
    for (int ch = 0; ch < 3; ++ch) {
        #pragma unroll
        for (int k = 0; k < 8; ++k)
            u.v[ch][k] = quantize2<CLAMP>(add2_rn(AF, mul2_rn(F, add2_rn(u.v[ch][k], NMA))));
    }
}
__device__ __forceinline__ uint32_t luma_acc_bits(uint32_t r, uint32_t g, uint32_t b) {
    return 19595u * r + 38470u * g + 7471u * b + 0x8000u;
}
template <bool CLAMP>
__device__ __forceinline__ void unit2_saturation(Unit2& u, float f) {
    const f32x2 F = pack2(f, f), MAG = pack2(kMagic, kMagic);
    #pragma unroll
    for (int k = 0; k < 8; ++k) {
        float r0, r1, g0, g1, b0, b1;
        unpack2(u.v[0][k], r0, r1); unpack2(u.v[1][k], g0, g1); unpack2(u.v[2][k], b0, b1);
        uint32_t a0 = luma_acc_bits(__float_as_uint(r0), __float_as_uint(g0), __float_as_uint(b0));
        uint32_t a1 = luma_acc_bits(__float_as_uint(r1), __float_as_uint(g1), __float_as_uint(b1));
        // -(2^23 + L): bit pattern 0xCB0000LL
        const f32x2 NML = pack2(__uint_as_float(__byte_perm(a0, 0xCB000000u, 0x7542)),
                                __uint_as_float(__byte_perm(a1, 0xCB000000u, 0x7542)));
        const f32x2 NLF = add2_rn(NML, MAG);                // -(float)L, exact
        #pragma unroll
        for (int ch = 0; ch < 3; ++ch) {
            f32x2 p = mul2_rn(F, add2_rn(u.v[ch][k], NML)); // alpha * (v - L)
            f32x2 t;
            asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(t) : "l"(p), "l"(NLF));   // L + p (same rounding as Pillow's A + p)
            u.v[ch][k] = quantize2<CLAMP>(t);
        }
    }
}
__device__ __forceinline__ uint32_t unit2_luma_sum(const Unit2& u) {
    uint32_t s = 0;
    #pragma unroll
    for (int k = 0; k < 8; ++k) {
        float r0, r1, g0, g1, b0, b1;
        unpack2(u.v[0][k], r0, r1); unpack2(u.v[1][k], g0, g1); unpack2(u.v[2][k], b0, b1);
        s += luma_acc_bits(__float_as_uint(r0), __float_as_uint(g0), __float_as_uint(b0)) >> 16;
        s += luma_acc_bits(__float_as_uint(r1), __float_as_uint(g1), __float_as_uint(b1)) >> 16;
    }
    return s;
}
__device__ __forceinline__ void unit2_apply(Unit2& u, int op, float f) {
    if (op == 0) {
        if (f > 1.0f) unit2_brightness<true>(u, f); else unit2_brightness<false>(u, f);
    } else {
        if (blend_needs_clip(f)) unit2_saturation<true>(u, f); else unit2_saturation<false>(u, f);
    }
}

// sum NV values over the CTA; result valid in every thread

// sum NV values over the CTA; result valid in every thread
template <int NV>
__device__ __forceinline__ void block_sum_vec(uint32_t (&v)[NV], uint32_t* red /* (NT2/32)*8 + 8 words */) {
    #pragma unroll
    for (int i = 0; i < NV; ++i)
        for (int o = 16; o > 0; o >>= 1) v[i] += __shfl_xor_sync(0xFFFFFFFFu, v[i], o);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) {
        #pragma unroll
        for (int i = 0; i < NV; ++i) red[(threadIdx.x >> 5) * 8 + i] = v[i];
    }
    __syncthreads();
    #pragma unroll
    for (int i = 0; i < NV; ++i) {
        uint32_t t = 0;
        #pragma unroll
        for (int w = 0; w < NT2 / 32; ++w) t += red[w * 8 + i];
        v[i] = t;
    }
}

// check_color_similarity(...) > 0.99 from exact integer channel sums.  The reference value is
//   cos = <a/|a|, b/|b|>,  a = (sum1/n1)/255, b = (sum2/n2)/255   (float64, dataset_generator.py:147-159)
// The scale factors cancel, so cos = <s1,s2> / sqrt(<s1,s1><s2,s2>) up to ~1e-15 relative error; only when that
// cheap value is within 1e-9 of the threshold is the reference's exact operation order evaluated.
__device__ inline bool similarity_branch(const uint32_t* s1, double n1, const uint32_t* s2, double n2, double& cosv) {
    const double a0 = s1[0], a1 = s1[1], a2 = s1[2], b0 = s2[0], b1 = s2[1], b2 = s2[2];
    const double dot = a0 * b0 + a1 * b1 + a2 * b2;
    const double nn = (a0 * a0 + a1 * a1 + a2 * a2) * (b0 * b0 + b1 * b1 + b2 * b2);
    double c = (nn > 0.0) ? dot * rsqrt(nn) : 0.0;
    if (fabs(c - 0.99) < 1e-9 || nn == 0.0) c = cosine_of_means(s1, n1, s2, n2);
    cosv = c;
    return c > 0.99;
}

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(NT2, 2)
generate_cluster_kernel(const __grid_constant__ SsadBatchDesc d, const __grid_constant__ ClusterGeom g,
                        const __grid_constant__ CUtensorMap tmap,
                        const SsadImageParams* __restrict__ params, const uint8_t* __restrict__ src,
                        const uint8_t* __restrict__ donor_bank, const uint32_t* __restrict__ mask_bits,
                        const void* __restrict__ norm_lut, uint8_t* __restrict__ out_u8, void* __restrict__ out_norm,
                        int64_t* __restrict__ y_out, SsadDebug* __restrict__ debug) {
    extern __shared__ __align__(128) uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((128u - (smem_u32(smem_raw) & 127u)) & 127u);   // TMA destinations need 128-B alignment
    uint8_t* img = smem + g.o_buf;
    uint32_t* poly = reinterpret_cast<uint32_t*>(smem + g.o_poly);
    uint32_t* rowpre = reinterpret_cast<uint32_t*>(smem + g.o_rowpre);
    uint32_t* red = reinterpret_cast<uint32_t*>(smem + g.o_red);
    SsadImageParams& P = *reinterpret_cast<SsadImageParams*>(smem + g.o_par);
    Decisions2& D = *reinterpret_cast<Decisions2*>(smem + g.o_dec);
    uint32_t* xchg = reinterpret_cast<uint32_t*>(smem + g.o_xchg);
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + g.o_bar);
    Polygon& G = *reinterpret_cast<Polygon*>(smem + g.o_pgn);

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    constexpr int NW = NT2 / 32;
    const uint32_t rank = cluster_ctarank(), peer = rank ^ 1u;
    const int cid = (int)cluster_id_x(), ncl = (int)cluster_nclusters_x();
    const int H = g.H, W = g.W, Hh = g.Hh, RB = g.row_bytes;
    const int y0 = (int)rank * Hh;
    const int npix = H * W;

    if (tid == 0) {
        mbar_init(full, 1);
        mbar_fence_init();
        tma_prefetch_desc(&tmap);
    }
    if (d.out_norm_dtype != SSAD_NORM_NONE && out_norm != nullptr) {
        const int nw = (d.out_norm_dtype == SSAD_NORM_F32) ? 768 : 384;
        for (int i = tid; i < nw; i += NT2)
            reinterpret_cast<uint32_t*>(smem + g.o_lut)[i] = reinterpret_cast<const uint32_t*>(norm_lut)[i];
    }
    __syncthreads();

    uint32_t phase = 0;
    int xs = 0;                                             // exchange slot parity
    for (int b = cid; b < d.batch; b += ncl) {
        // ---- parameters; thread 0 starts the TMA fetch of this CTA's half (the co-resident CTA computes meanwhile)
        if (tid < (int)(sizeof(SsadImageParams) / 4))
            reinterpret_cast<uint32_t*>(&P)[tid] = reinterpret_cast<const uint32_t*>(params + b)[tid];
        __syncthreads();
        const int cx0 = P.crop_off[0], cy0 = P.crop_off[1];
        const bool warped = P.warp_en != 0;
        if (tid == 0 && !warped) {
            fence_proxy_async();
            mbar_expect_tx(full, g.raw_bytes);
            // u32 elements: x = 16-byte-aligned start of the window row, in words
            tma_load_3d(img, &tmap, full, ((cx0 * 3) & ~15) >> 2, cy0 + y0, b);
        }
        const uint8_t* simg = src + (size_t)b * d.src_stride;
        const uint32_t* mrow0 = nullptr;
        if (d.mask_mode != SSAD_MASK_ALL)
            mrow0 = mask_bits + (d.mask_mode == SSAD_MASK_PER_IMAGE ? (size_t)b * d.mask_stride_words : 0);
        const bool defect = P.label != 0;
        const int rl = P.rect[0], rt = P.rect[1], pw = P.rect[2], ph = P.rect[3];

        // ================= pixel-independent decisions (overlap the TMA flight) =================
        // mask window: per-row popcounts -> prefix (coords_map is row-major, datasets.py:263-264)
        if (defect && d.mask_mode != SSAD_MASK_ALL) {
            for (int r = tid; r < H; r += NT2)
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
        if (y != 0) {
            // centres + container clamp (dataset_generator.py:104-144, 259-265): one thread per paste
            if (tid < npaste) {
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
                int sw = (y == 1) ? pw : P.scar_size[0], sh = (y == 1) ? ph : P.scar_size[1];
                int pl, pt;
                clamp_to_container(y == 1 ? d.box_poly : d.box_scar, sw, sh, cx, cy, pl, pt);
                D.paste[tid][0] = pl; D.paste[tid][1] = pt;
            }
            if (y == 1 && tid == 32) polygon_setup(G, P.poly_xy, P.poly_n, ph);
            __syncthreads();
            if (y == 1) {
                // rect2poly rows that land in my half (dataset_generator.py:63-100): one thread per canvas row
                const int pt = D.paste[0][1];
                const int v_lo = max(0, y0 - pt), v_hi = min(ph, y0 + Hh - pt);
                for (int v = v_lo + tid; v < v_hi; v += NT2) {
                    uint32_t* row = poly + (size_t)(v - v_lo) * g.poly_pitch;
                    for (int j = 0; j < (int)g.poly_pitch; ++j) row[j] = 0;
                    polygon_fill_row(G, v, pw, row);
                }
            }
        }

        // ================= x: RandomAffine gather (datasets.py:220-222) or the TMA-fetched window (:242-251) =======
        if (warped) {
            const int32_t a0 = P.warp_fix[0], a1 = P.warp_fix[1], a2 = P.warp_fix[2];
            const int32_t a3 = P.warp_fix[3], a4 = P.warp_fix[4], a5 = P.warp_fix[5];
            const uint8_t* sbase = simg + (size_t)cy0 * d.src_pitch + (size_t)cx0 * 3;
            const int spitch = (int)d.src_pitch;
            for (int yl = warp; yl < Hh; yl += NW) {
                const int yo = y0 + yl;
                const int xr = a2 + yo * a1, yr = a5 + yo * a4;
                uint8_t* orow = img + yl * RB;
                for (int xo = lane; xo < W; xo += 32) {
                    const int xin = (xr + xo * a0) >> 16;
                    const int yin = (yr + xo * a3) >> 16;
                    uint32_t r = 0, gg = 0, bl = 0;
                    if ((unsigned)xin < (unsigned)W && (unsigned)yin < (unsigned)H) {
                        const uint8_t* q = sbase + yin * spitch + xin * 3;
                        r = __ldg(q); gg = __ldg(q + 1); bl = __ldg(q + 2);
                    }
                    uint8_t* o = orow + xo * 3;
                    o[0] = (uint8_t)r; o[1] = (uint8_t)gg; o[2] = (uint8_t)bl;
                }
            }
        } else {
            mbar_wait(full, phase);
            phase ^= 1u;
            if (g.raw_pitch != RB) {
                // rows landed at pitch RB+16 starting at the aligned address; shift them down to the dense
                // layout: every thread stages its 16-byte chunks in registers, then all write
                const int m = (cx0 * 3) & 15, mo = m >> 2, sh = (m & 3) * 8;
                const int nchunks = Hh * g.cpr;
                constexpr int KMAX = 8;
                const int dr = NT2 / g.cpr, dj = NT2 - dr * g.cpr;     // chunk index += NT2  ->  (row += dr, col += dj) with carry
                int r = tid / g.cpr, j = tid - r * g.cpr;
                for (int base = 0; base < nchunks; base += KMAX * NT2) {
                    uint4 st[KMAX];
                    #pragma unroll
                    for (int k = 0; k < KMAX; ++k) {
                        const int c = base + k * NT2 + tid;
                        if (c < nchunks) {
                            const uint4* sp = reinterpret_cast<const uint4*>(img + r * g.raw_pitch) + j;
                            const uint4 lo = sp[0], hi = sp[1];
                            const uint32_t w[8] = {lo.x, lo.y, lo.z, lo.w, hi.x, hi.y, hi.z, hi.w};
                            uint32_t o[4];
                            #pragma unroll
                            for (int q = 0; q < 4; ++q) {
                                uint32_t a, bb;
                                if (mo == 0) { a = w[q]; bb = w[q + 1]; }
                                else if (mo == 1) { a = w[q + 1]; bb = w[q + 2]; }
                                else if (mo == 2) { a = w[q + 2]; bb = w[q + 3]; }
                                else { a = w[q + 3]; bb = w[q + 4]; }
                                o[q] = __funnelshift_r(a, bb, sh);
                            }
                            st[k] = make_uint4(o[0], o[1], o[2], o[3]);
                        }
                        r += dr; j += dj;
                        if (j >= g.cpr) { j -= g.cpr; ++r; }
                    }
                    __syncthreads();
                    #pragma unroll
                    for (int k = 0; k < KMAX; ++k) {
                        const int c = base + k * NT2 + tid;
                        if (c < nchunks) reinterpret_cast<uint4*>(img)[c] = st[k];
                    }
                    __syncthreads();
                }
            }
        }
        __syncthreads();

        if (y != 0) {
            // donor image (datasets.py:225-228, 253); patch pixel (u, v) = donor window pixel (rl+u, rt+v), PIL's crop
            // pads with 0 outside the window
            const uint8_t* dimg = simg;
            int dpitch = (int)d.src_pitch, dH = d.src_height, dW = d.src_width;
            if (P.donor_idx >= 0) {
                dimg = donor_bank + (size_t)P.donor_idx * d.donor_stride;
                dpitch = (int)d.donor_pitch; dH = d.donor_height; dW = d.donor_width;
            }
            const int dox = P.donor_off[0], doy = P.donor_off[1];
            const int u_hi = min(min(pw, d.donor_win_w - rl), dW - dox - rl);         // u < u_hi is inside window and image (rl >= 0)
            const int v_hiw = min(min(ph, d.donor_win_h - rt), dH - doy - rt);
            const uint8_t* dbase = dimg + (size_t)(doy + rt) * dpitch + (size_t)(dox + rl) * 3;   // patch pixel (0,0)

            // ---- partial sums: RGB of my half of x, RGB of my rows of the patch (dataset_generator.py:147-159)
            uint32_t s[6] = {0, 0, 0, 0, 0, 0};
            {
                const uint4* w4 = reinterpret_cast<const uint4*>(img);
                const int nv = (int)(g.buf_bytes / 48);                 // 48 B = 16 pixels = 3 vectors
                for (int i = tid; i < nv; i += NT2) {
                    const uint4 A = w4[3 * i], B2 = w4[3 * i + 1], C2 = w4[3 * i + 2];
                    const uint32_t w[12] = {A.x, A.y, A.z, A.w, B2.x, B2.y, B2.z, B2.w, C2.x, C2.y, C2.z, C2.w};
                    #pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        const uint32_t w0 = w[3 * q], w1 = w[3 * q + 1], w2 = w[3 * q + 2];
                        s[0] = __dp4a(w0, 0x01000001u, s[0]); s[0] = __dp4a(w1, 0x00010000u, s[0]); s[0] = __dp4a(w2, 0x00000100u, s[0]);
                        s[1] = __dp4a(w0, 0x00000100u, s[1]); s[1] = __dp4a(w1, 0x01000001u, s[1]); s[1] = __dp4a(w2, 0x00010000u, s[1]);
                        s[2] = __dp4a(w0, 0x00010000u, s[2]); s[2] = __dp4a(w1, 0x00000100u, s[2]); s[2] = __dp4a(w2, 0x01000001u, s[2]);
                    }
                }
            }
            if (P.src_variant != 2) {
                for (int v = (int)rank + 2 * warp; v < v_hiw; v += 2 * NW) {
                    const uint8_t* q = dbase + v * dpitch;
                    for (int u = lane; u < u_hi; u += 32) {
                        s[3] += __ldg(q + 3 * u); s[4] += __ldg(q + 3 * u + 1); s[5] += __ldg(q + 3 * u + 2);
                    }
                }
            }
            block_sum_vec<6>(s, red);
            if (tid == 0) {
                #pragma unroll
                for (int i = 0; i < 6; ++i) xchg[xs * 8 + i] = s[i];
            }
            cluster_arrive();
            cluster_wait();
            if (tid == 0) {
                uint32_t tot[6];
                #pragma unroll
                for (int i = 0; i < 6; ++i) tot[i] = s[i] + dsmem_ld_u32(&xchg[xs * 8 + i], peer);
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
                double cs;
                D.sim = similarity_branch(tot, (double)npix, psum, (double)np, cs);     // datasets.py:295 / 335
                D.cosine = cs;
                D.bf1 = P.bright_f[0]; D.bf2 = P.bright_f[1];
                if (D.flat && D.sim)                // brightness of a flat patch is a flat patch
                    for (int c = 0; c < 3; ++c) D.flat_rgb[c] = brightness_u8(brightness_u8(D.flat_rgb[c], D.bf1), D.bf2);
            }
            xs ^= 1;
            __syncthreads();

            const bool sim = D.sim != 0, flat = D.flat != 0;
            const float bf1 = D.bf1, bf2 = D.bf2;
            const uint32_t fr = D.flat_rgb[0], fg = D.flat_rgb[1], fb = D.flat_rgb[2];
            // colour of patch pixel (u, v): u in [0,pw), v in [0,ph)
            auto patch_colour = [&](int u, int v, uint32_t& r, uint32_t& gg, uint32_t& bl) {
                if (flat) { r = fr; gg = fg; bl = fb; return; }
                r = gg = bl = 0;
                if (u < u_hi && v < v_hiw) {
                    const uint8_t* q = dbase + v * dpitch + 3 * u;
                    r = __ldg(q); gg = __ldg(q + 1); bl = __ldg(q + 2);
                }
                if (sim) {      // two successive ImageEnhance.Brightness, each re-quantised (datasets.py:298-299)
                    r = brightness_u8(brightness_u8(r, bf1), bf2);
                    gg = brightness_u8(brightness_u8(gg, bf1), bf2);
                    bl = brightness_u8(brightness_u8(bl, bf1), bf2);
                }
            };
            if (y == 1) {
                // ---- paste_patch(x, patch, coords, mask) (dataset_generator.py:268-275), rows of my half
                const int pl = D.paste[0][0], pt = D.paste[0][1];
                const int v_lo = max(0, y0 - pt), v_hi = min(ph, y0 + Hh - pt);
                const int u_lo = max(0, -pl), u_end = min(pw, W - pl);
                for (int v = v_lo + warp; v < v_hi; v += NW) {
                    const uint32_t* prow = poly + (size_t)(v - v_lo) * g.poly_pitch;
                    uint8_t* orow = img + (pt + v - y0) * RB + pl * 3;
                    for (int u = u_lo + lane; u < u_end; u += 32) {
                        if (!((prow[u >> 5] >> (u & 31)) & 1u)) continue;
                        uint32_t r, gg, bl;
                        patch_colour(u, v, r, gg, bl);
                        uint8_t* q = orow + 3 * u;
                        q[0] = (uint8_t)r; q[1] = (uint8_t)gg; q[2] = (uint8_t)bl;
                    }
                }
            } else {
                // ---- scar: rotate(expand=True) NEAREST evaluated at paste time, k pastes in order (datasets.py:340-355)
                const int nw = P.scar_size[0], nh = P.scar_size[1];
                const int32_t a0 = P.scar_fix[0], a1 = P.scar_fix[1], a2 = P.scar_fix[2];
                const int32_t a3 = P.scar_fix[3], a4 = P.scar_fix[4], a5 = P.scar_fix[5];
                for (int k = 0; k < npaste; ++k) {
                    const int pl = D.paste[k][0], pt = D.paste[k][1];
                    const int v_lo = max(0, y0 - pt), v_hi = min(nh, y0 + Hh - pt);
                    const int u_lo = max(0, -pl), u_end = min(nw, W - pl);
                    for (int v = v_lo + warp; v < v_hi; v += NW) {
                        uint8_t* orow = img + (pt + v - y0) * RB + pl * 3;
                        const int xr = a2 + v * a1, yr = a5 + v * a4;
                        for (int u = u_lo + lane; u < u_end; u += 32) {
                            const int xin = (xr + u * a0) >> 16, yin = (yr + u * a3) >> 16;
                            if ((unsigned)xin >= (unsigned)pw || (unsigned)yin >= (unsigned)ph) continue;   // alpha 0 outside the source
                            uint32_t r, gg, bl;
                            patch_colour(xin, yin, r, gg, bl);
                            uint8_t* q = orow + 3 * u;
                            q[0] = (uint8_t)r; q[1] = (uint8_t)gg; q[2] = (uint8_t)bl;
                        }
                    }
                    __syncthreads();
                }
            }
            __syncthreads();
        }

        // ================= ColorJitter (datasets.py:391) =================
        int ops[3];
        float fac[3];
        int nops = 0, cpos = -1;
        #pragma unroll
        for (int st = 0; st < 4; ++st) {
            int fn = P.jit_order[st];
            if (fn < 3 && ((P.jit_en >> fn) & 1)) {
                if (fn == 1) cpos = nops;
                ops[nops] = fn; fac[nops] = P.jit_f[fn]; ++nops;
            }
        }
        uint4* img4 = reinterpret_cast<uint4*>(img);
        int cmean = -1;
        if (cpos >= 0) {
            uint32_t ls[1] = {0};
            for (int ui = tid; ui < g.units_half; ui += NT2) {
                Unit2 u;
                unit2_unpack(u, img4[3 * ui], img4[3 * ui + 1], img4[3 * ui + 2]);
                if (cpos > 0) {
                    unit2_apply(u, ops[0], fac[0]);
                    if (cpos > 1) unit2_apply(u, ops[1], fac[1]);
                    unit2_pack(u, img4[3 * ui], img4[3 * ui + 1], img4[3 * ui + 2]);
                }
                ls[0] += unit2_luma_sum(u);
            }
            block_sum_vec<1>(ls, red);
            if (tid == 0) xchg[xs * 8] = ls[0];
            cluster_arrive();
            cluster_wait();
            // every thread needs the mean: lane 0 of each warp fetches the peer's partial sum
            uint32_t tot = ls[0] + dsmem_ld_u32(&xchg[xs * 8], peer);
            cmean = (int)((2ull * tot + (uint32_t)npix) / (2ull * (uint32_t)npix));   // int(sum/count + 0.5)
            xs ^= 1;
        }
        const float cf = (cpos >= 0) ? fac[cpos] : 1.0f;
        const bool cclip = blend_needs_clip(cf);
        const int first_post = (cpos >= 0) ? cpos + 1 : 0;
        uint8_t* obase = out_u8 ? out_u8 + (size_t)b * d.out_u8_stride : nullptr;
        for (int ui = tid; ui < g.units_half; ui += NT2) {
            Unit2 u;
            unit2_unpack(u, img4[3 * ui], img4[3 * ui + 1], img4[3 * ui + 2]);
            if (cpos >= 0) {
                if (cclip) unit2_contrast<true>(u, cf, cmean); else unit2_contrast<false>(u, cf, cmean);
            }
            if (first_post < nops) {
                unit2_apply(u, ops[first_post], fac[first_post]);
                if (first_post + 1 < nops) unit2_apply(u, ops[first_post + 1], fac[first_post + 1]);
            }
            const size_t p0 = (size_t)y0 * W + (size_t)ui * 16;     // dense row-major half: unit -> pixel offset
            if (obase) {
                uint4 a, bq, c;
                unit2_pack(u, a, bq, c);
                uint4* o = reinterpret_cast<uint4*>(obase + p0 * 3);
                o[0] = a; o[1] = bq; o[2] = c;
            }
            if (out_norm) {
                // ToTensor + Normalize (datasets.py:392-393): exact table per channel, NCHW planes
                float m[48];
                unit2_scalars(u, m);
                if (d.out_norm_dtype == SSAD_NORM_F32) {
                    const float* lut = reinterpret_cast<const float*>(smem + g.o_lut);
                    float* ob = reinterpret_cast<float*>(out_norm) + (size_t)b * 3 * npix;
                    #pragma unroll
                    for (int c = 0; c < 3; ++c) {
                        float4* o = reinterpret_cast<float4*>(ob + (size_t)c * npix + p0);
                        #pragma unroll
                        for (int qq = 0; qq < 4; ++qq) {
                            float4 v;
                            v.x = lut[c * 256 + magic_to_u8(m[3 * (4 * qq + 0) + c])];
                            v.y = lut[c * 256 + magic_to_u8(m[3 * (4 * qq + 1) + c])];
                            v.z = lut[c * 256 + magic_to_u8(m[3 * (4 * qq + 2) + c])];
                            v.w = lut[c * 256 + magic_to_u8(m[3 * (4 * qq + 3) + c])];
                            o[qq] = v;
                        }
                    }
                } else {
                    const uint16_t* lut = reinterpret_cast<const uint16_t*>(smem + g.o_lut);
                    uint16_t* ob = reinterpret_cast<uint16_t*>(out_norm) + (size_t)b * 3 * npix;
                    #pragma unroll
                    for (int c = 0; c < 3; ++c) {
                        uint4* o = reinterpret_cast<uint4*>(ob + (size_t)c * npix + p0);
                        uint32_t w[8];
                        #pragma unroll
                        for (int qq = 0; qq < 8; ++qq) {
                            uint32_t lo = lut[c * 256 + magic_to_u8(m[3 * (2 * qq) + c])];
                            uint32_t hi = lut[c * 256 + magic_to_u8(m[3 * (2 * qq + 1) + c])];
                            w[qq] = lo | (hi << 16);
                        }
                        o[0] = make_uint4(w[0], w[1], w[2], w[3]);
                        o[1] = make_uint4(w[4], w[5], w[6], w[7]);
                    }
                }
            }
        }
        if (tid == 0 && rank == 0) {
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
                    for (int c = 0; c < 3; ++c) o.avg_rgb[c] = (int)D.avg_rgb[c];
                }
                debug[b] = o;
            }
        }
        __syncthreads();            // buffer, P and D are free for the next image
    }
    // the peer may still be reading this CTA's exchange slots
    cluster_arrive();
    cluster_wait();
}

}  // namespace ssad

// ------------------------------------------------------------------ host launch
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
    static EncodeTiledFn fn = nullptr;
    static bool tried = false;
    if (!tried) {
        tried = true;
        void* p = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    }
    return fn;
}

// returns SSAD_OK when launched, SSAD_EUNSUPPORTED when the caller should use the generic kernel
int ssad_launch_cluster(const SsadBatchDesc& d, const SsadImageParams* params, const uint8_t* src, const uint8_t* donor_bank,
                        const uint32_t* mask_bits, const void* norm_lut, uint8_t* out_u8, void* out_norm, int64_t* y_out,
                        SsadDebug* debug, cudaStream_t stream) {
    ssad::ClusterGeom g;
    // `aligned`: every window row of the batch starts on a 16-byte boundary (host promise via crop_align16);
    // the box then holds exactly the window and no realign pass runs
    const bool aligned = d.crop_align16 != 0;
    if (!ssad::make_geom(d, aligned, g)) return SSAD_EUNSUPPORTED;
    if ((reinterpret_cast<uintptr_t>(src) & 15) || (d.src_pitch & 15) || (d.src_stride & 15) || (d.src_width % 4)) return SSAD_EUNSUPPORTED;
    int dev = 0, max_smem = 0, sms = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return ssad_set_cuda_error(e);
    cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if ((long long)g.total + 128 > max_smem) return SSAD_EUNSUPPORTED;
    EncodeTiledFn enc = get_encode_fn();
    if (!enc) return SSAD_EUNSUPPORTED;
    CUtensorMap tmap;
    // rows as u32 words; the innermost extent is padded up to whole 16-byte groups so that the aligned
    // superset of any window is addressable (reads past the row end stay inside the pitch / are zero-filled)
    cuuint64_t gdim[3] = {(cuuint64_t)((d.src_width * 3 + 15) / 16 * 4), (cuuint64_t)d.src_height, (cuuint64_t)d.batch};
    if ((long long)gdim[0] * 4 > d.src_pitch) gdim[0] = (cuuint64_t)(d.src_pitch / 4);
    cuuint64_t gstr[2] = {(cuuint64_t)d.src_pitch, (cuuint64_t)d.src_stride};
    cuuint32_t box[3] = {(cuuint32_t)(g.raw_pitch / 4), (cuuint32_t)g.Hh, 1};
    cuuint32_t estr[3] = {1, 1, 1};
    CUresult cr = enc(&tmap, CU_TENSOR_MAP_DATA_TYPE_UINT32, 3, const_cast<uint8_t*>(src), gdim, gstr, box, estr,
                      CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                      CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (cr != CUDA_SUCCESS) return SSAD_EUNSUPPORTED;
    e = cudaFuncSetAttribute(ssad::generate_cluster_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)g.total + 128);
    if (e != cudaSuccess) return ssad_set_cuda_error(e);
    // two CTAs per SM when their shared memory fits twice (228 KB per SM, 1 KB reserved per CTA)
    const int ctas_per_sm = (2 * ((long long)g.total + 128 + 1024) <= 228 * 1024) ? 2 : 1;
    int nclusters = sms * ctas_per_sm / 2;
    if (nclusters > d.batch) nclusters = d.batch;
    if (nclusters < 1) nclusters = 1;
    ssad::generate_cluster_kernel<<<2 * nclusters, ssad::NT2, g.total + 128, stream>>>(d, g, tmap, params, src, donor_bank, mask_bits,
                                                                                  norm_lut, out_u8, out_norm, y_out, debug);
    e = cudaGetLastError();
    if (e != cudaSuccess) return ssad_set_cuda_error(e);
    return SSAD_OK;
}
