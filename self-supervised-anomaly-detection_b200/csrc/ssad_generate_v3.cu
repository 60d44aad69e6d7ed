// ssad_generate, B200 path v3: persistent clusters of C CTAs (C = 1, 2, 4 or 8) per image, rows split over the CTAs.
//
//   * each CTA owns H / C rows.  If they fit its shared memory they are fetched ONCE by TMA and stay resident through
//     the passes ("resident" mode: 256x256 -> C = 4, 48 KB per CTA, four CTAs of 256 threads per SM).  Otherwise the
//     rows stream through two chunk buffers ("streaming" mode: 1024x1024 -> C = 8, 16-row chunks), re-fetched per pass
//     by TMA while the previous chunk is processed; the image is only read from HBM once if it stays in the 126 MB L2
//     (at most 148 * R / C images are in flight);
//   * passes over the pixels:  P1 RGB sums of x (check_color_similarity; defect labels only)  ->  cluster reduction
//     ->  paste (polygon / scars)  ->  P2 jitter ops ahead of Contrast + sum of L  ->  cluster reduction (Contrast
//     mean)  ->  P3 Contrast + remaining ops + stores.  Whole-image sums cross the cluster through distributed
//     shared memory + barrier.cluster;
//   * jitter arithmetic on 8-pixel register units in packed f32x2 (FADD2 / FMUL2 / FADD2.RZ) -- bit-identical to
//     Pillow's separate float32 multiply and add -- at <= 64 registers per thread so that 32 warps are resident per SM;
//   * TMA box = 4-D (row segment words <= 256, segments, rows, image) so that any row length lands dense row-major.
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "ssad_device.cuh"
#include "ssad_sm100.cuh"
#include "ssad_unit8.cuh"

int ssad_set_cuda_error(cudaError_t e);

namespace ssad {
namespace v3 {

using namespace u8x;

constexpr int NT = 256;
constexpr int NWARP = NT / 32;
constexpr int MAXC = 8;

struct Geom {
    int H, W, C, rows_cta, Rc, nchunks, resident;
    int row_bytes, raw_pitch, units_chunk;
    uint32_t chunk_bytes, raw_bytes;
    int box_w, nseg;
    uint32_t poly_pitch, poly_rows;
    uint32_t o_buf0, o_buf1, o_poly, o_rowpre, o_lut, o_red, o_par, o_dec, o_xchg, o_tot, o_bar, o_pgn, total;
};

struct Dec {
    int sim, npaste, flat, cmean;
    int paste[5][2];
    uint32_t flat_rgb[3];
    uint32_t avg_rgb[3];
    float bf1, bf2;
    double cosine;
};

__device__ __forceinline__ void tma_load_4d(void* dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1, int c2, int c3) {
    asm volatile(
        "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
        ::"r"(smem_u32(dst)), "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3) : "memory");
}

// sum NV values over the CTA; result valid in every thread
template <int NV>
__device__ __forceinline__ void block_sum_vec(uint32_t (&v)[NV], uint32_t* red /* NWARP * 8 words */) {
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
        for (int w = 0; w < NWARP; ++w) t += red[w * 8 + i];
        v[i] = t;
    }
}

// shift rows that TMA landed at pitch RB + 16 (16-byte aligned superset of the window) down to the dense layout
template <int MO>
__device__ __forceinline__ void realign_rows(uint8_t* img, int rows, int cpr, int raw_pitch, int sh, int tid) {
    const int nchunks = rows * cpr;
    constexpr int K = 4;
    const int dr = NT / cpr, dj = NT - dr * cpr;           // chunk index += NT  ->  (row += dr, col += dj) with carry
    int r = tid / cpr, j = tid - r * cpr;
    for (int base = 0; base < nchunks; base += K * NT) {
        uint4 st[K];
        #pragma unroll
        for (int k = 0; k < K; ++k) {
            const int c = base + k * NT + tid;
            if (c < nchunks) {
                const uint4* sp = reinterpret_cast<const uint4*>(img + r * raw_pitch) + j;
                const uint4 lo = sp[0], hi = sp[1];
                const uint32_t w[8] = {lo.x, lo.y, lo.z, lo.w, hi.x, hi.y, hi.z, hi.w};
                st[k] = make_uint4(__funnelshift_r(w[MO], w[MO + 1], sh), __funnelshift_r(w[MO + 1], w[MO + 2], sh),
                                   __funnelshift_r(w[MO + 2], w[MO + 3], sh), __funnelshift_r(w[MO + 3], w[MO + 4], sh));
            }
            r += dr; j += dj;
            if (j >= cpr) { j -= cpr; ++r; }
        }
        __syncthreads();
        #pragma unroll
        for (int k = 0; k < K; ++k) {
            const int c = base + k * NT + tid;
            if (c < nchunks) reinterpret_cast<uint4*>(img)[c] = st[k];
        }
        __syncthreads();
    }
}

__global__ void __launch_bounds__(NT, 4)
generate_v3_kernel(const __grid_constant__ SsadBatchDesc d, const __grid_constant__ Geom g,
                   const __grid_constant__ CUtensorMap tmap,
                   const SsadImageParams* __restrict__ params, const uint8_t* __restrict__ src,
                   const uint8_t* __restrict__ donor_bank, const uint32_t* __restrict__ mask_bits,
                   const void* __restrict__ norm_lut, uint8_t* __restrict__ out_u8, void* __restrict__ out_norm,
                   int64_t* __restrict__ y_out, SsadDebug* __restrict__ debug) {
    extern __shared__ __align__(128) uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((128u - (smem_u32(smem_raw) & 127u)) & 127u);   // TMA destinations need 128-B alignment
    uint8_t* const buf0 = smem + g.o_buf0;
    uint8_t* const buf1 = smem + g.o_buf1;
    uint32_t* poly = reinterpret_cast<uint32_t*>(smem + g.o_poly);
    uint32_t* rowpre = reinterpret_cast<uint32_t*>(smem + g.o_rowpre);
    uint32_t* red = reinterpret_cast<uint32_t*>(smem + g.o_red);
    SsadImageParams& P = *reinterpret_cast<SsadImageParams*>(smem + g.o_par);
    Dec& D = *reinterpret_cast<Dec*>(smem + g.o_dec);
    uint32_t* xchg = reinterpret_cast<uint32_t*>(smem + g.o_xchg);
    uint32_t* tot = reinterpret_cast<uint32_t*>(smem + g.o_tot);
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + g.o_bar);
    Polygon& G = *reinterpret_cast<Polygon*>(smem + g.o_pgn);

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int C = g.C;
    const uint32_t rank = (C > 1) ? cluster_ctarank() : 0u;
    const int cid = (C > 1) ? (int)cluster_id_x() : (int)blockIdx.x;
    const int ncl = (C > 1) ? (int)cluster_nclusters_x() : (int)gridDim.x;
    const int H = g.H, W = g.W, RB = g.row_bytes, Rc = g.Rc, nch = g.nchunks;
    const bool resident = g.resident != 0;
    const int part_y0 = (int)rank * g.rows_cta;
    const int npix = H * W;

    if (tid == 0) {
        mbar_init(&bars[0], 1);
        mbar_init(&bars[1], 1);
        mbar_fence_init();
        tma_prefetch_desc(&tmap);
    }
    if (d.out_norm_dtype != SSAD_NORM_NONE && out_norm != nullptr) {
        const int nw = (d.out_norm_dtype == SSAD_NORM_F32) ? 768 : 384;
        for (int i = tid; i < nw; i += NT)
            reinterpret_cast<uint32_t*>(smem + g.o_lut)[i] = reinterpret_cast<const uint32_t*>(norm_lut)[i];
    }
    __syncthreads();

    auto cluster_sync = [&]() __attribute__((always_inline)) {
        if (C > 1) { cluster_arrive(); cluster_wait(); } else __syncthreads();
    };

    uint32_t ph0 = 0, ph1 = 0;                              // mbarrier phase parities of the two buffers
    int xs = 0;                                             // exchange slot parity
    for (int b = cid; b < d.batch; b += ncl) {
        if (tid < (int)(sizeof(SsadImageParams) / 4))
            reinterpret_cast<uint32_t*>(&P)[tid] = reinterpret_cast<const uint32_t*>(params + b)[tid];
        __syncthreads();
        const int cx0 = P.crop_off[0], cy0 = P.crop_off[1];
        const bool warped = P.warp_en != 0;
        const int xw0 = ((cx0 * 3) & ~15) >> 2;             // 16-byte-aligned start of the window row, in u32 elements

        // fetch q-th chunk load of this image into buffer (q & 1): chunk (q % nch) of this CTA's rows
        auto issue = [&](int q) __attribute__((always_inline)) {
            const int ch = q % nch, bi = resident ? 0 : (q & 1);
            fence_proxy_async();
            mbar_expect_tx(&bars[bi], g.raw_bytes);
            tma_load_4d(bi ? buf1 : buf0, &tmap, &bars[bi], xw0, 0, cy0 + part_y0 + ch * Rc, b);
        };
        if (tid == 0 && !warped) issue(0);

        const uint8_t* simg = src + (size_t)b * d.src_stride;
        const uint32_t* mrow0 = nullptr;
        if (d.mask_mode != SSAD_MASK_ALL)
            mrow0 = mask_bits + (d.mask_mode == SSAD_MASK_PER_IMAGE ? (size_t)b * d.mask_stride_words : 0);
        const bool defect = P.label != 0;
        const int rl = P.rect[0], rt = P.rect[1], pw = P.rect[2], ph = P.rect[3];

        // ================= pixel-independent decisions (overlap the TMA flight) =================
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
        int pv_lo = 0;                                       // first polygon row held in `poly`
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
                // rect2poly rows that land in my rows (dataset_generator.py:63-100): one thread per canvas row
                const int pt = D.paste[0][1];
                pv_lo = max(0, part_y0 - pt);
                const int v_hi = min(ph, part_y0 + g.rows_cta - pt);
                for (int v = pv_lo + tid; v < v_hi; v += NT) {
                    uint32_t* row = poly + (size_t)(v - pv_lo) * g.poly_pitch;
                    for (int j = 0; j < (int)g.poly_pitch; ++j) row[j] = 0;
                    polygon_fill_row(G, v, pw, row);
                }
            }
        }

        // jitter plan (datasets.py:391)
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
        const int npass = resident ? 1 : ((y != 0 ? 1 : 0) + (cpos >= 0 ? 1 : 0) + 1);
        const int nq = npass * nch;

        // bring chunk load q into its buffer: TMA wait (+ prefetch of q + 1) or the RandomAffine gather; returns the buffer
        auto acquire = [&](int q) __attribute__((always_inline)) -> uint8_t* {
            const int ch = q % nch, bi = resident ? 0 : (q & 1);
            uint8_t* cb = bi ? buf1 : buf0;
            const int gy0 = part_y0 + ch * Rc;
            if (warped) {
                // RandomAffine(3, scale=(1.05, 1.1)) NEAREST, fill 0 (datasets.py:220-222)
                const int32_t a0 = P.warp_fix[0], a1 = P.warp_fix[1], a2 = P.warp_fix[2];
                const int32_t a3 = P.warp_fix[3], a4 = P.warp_fix[4], a5 = P.warp_fix[5];
                const uint8_t* sbase = simg + (size_t)cy0 * d.src_pitch + (size_t)cx0 * 3;
                const int spitch = (int)d.src_pitch;
                for (int yl = warp; yl < Rc; yl += NWARP) {
                    const int yo = gy0 + yl;
                    const int xr = a2 + yo * a1, yr = a5 + yo * a4;
                    uint8_t* orow = cb + yl * RB;
                    for (int xo = lane; xo < W; xo += 32) {
                        const int xin = (xr + xo * a0) >> 16;
                        const int yin = (yr + xo * a3) >> 16;
                        uint32_t r = 0, gg = 0, bl = 0;
                        if ((unsigned)xin < (unsigned)W && (unsigned)yin < (unsigned)H) {
                            const uint8_t* q8 = sbase + yin * spitch + xin * 3;
                            r = __ldg(q8); gg = __ldg(q8 + 1); bl = __ldg(q8 + 2);
                        }
                        uint8_t* o = orow + xo * 3;
                        o[0] = (uint8_t)r; o[1] = (uint8_t)gg; o[2] = (uint8_t)bl;
                    }
                }
                __syncthreads();
                return cb;
            }
            if (!resident && tid == 0 && q + 1 < nq) issue(q + 1);     // the other buffer was released by the last __syncthreads
            if (bi) { mbar_wait(&bars[1], ph1); ph1 ^= 1u; } else { mbar_wait(&bars[0], ph0); ph0 ^= 1u; }
            if (g.raw_pitch != RB) {
                const int m = (cx0 * 3) & 15, sh = (m & 3) * 8, cpr = RB / 16;
                switch (m >> 2) {
                    case 0: realign_rows<0>(cb, Rc, cpr, g.raw_pitch, sh, tid); break;
                    case 1: realign_rows<1>(cb, Rc, cpr, g.raw_pitch, sh, tid); break;
                    case 2: realign_rows<2>(cb, Rc, cpr, g.raw_pitch, sh, tid); break;
                    default: realign_rows<3>(cb, Rc, cpr, g.raw_pitch, sh, tid); break;
                }
            }
            return cb;
        };

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

        // paste_patch / scar pastes into the rows [gy0, gy0 + Rc) held in cb (dataset_generator.py:268-275, datasets.py:340-355)
        auto paste_chunk = [&](uint8_t* cb, int gy0) __attribute__((always_inline)) {
            const bool sim = D.sim != 0, flat = D.flat != 0;
            const float bf1 = D.bf1, bf2 = D.bf2;
            const uint32_t fr = D.flat_rgb[0], fg = D.flat_rgb[1], fb = D.flat_rgb[2];
            auto patch_colour = [&](int u, int v, uint32_t& r, uint32_t& gg, uint32_t& bl) __attribute__((always_inline)) {
                if (flat) { r = fr; gg = fg; bl = fb; return; }
                r = gg = bl = 0;
                if (u < u_hi && v < v_hiw) {
                    const uint8_t* q8 = dbase + v * dpitch + 3 * u;
                    r = __ldg(q8); gg = __ldg(q8 + 1); bl = __ldg(q8 + 2);
                }
                if (sim) {      // two successive ImageEnhance.Brightness, each re-quantised (datasets.py:298-299)
                    r = brightness_u8(brightness_u8(r, bf1), bf2);
                    gg = brightness_u8(brightness_u8(gg, bf1), bf2);
                    bl = brightness_u8(brightness_u8(bl, bf1), bf2);
                }
            };
            if (y == 1) {
                const int pl = D.paste[0][0], pt = D.paste[0][1];
                const int v_lo = max(0, gy0 - pt), v_hi = min(ph, gy0 + Rc - pt);
                const int u_lo = max(0, -pl), u_end = min(pw, W - pl);
                for (int v = v_lo + warp; v < v_hi; v += NWARP) {
                    const uint32_t* prow = poly + (size_t)(v - pv_lo) * g.poly_pitch;
                    uint8_t* orow = cb + (pt + v - gy0) * RB + pl * 3;
                    for (int u = u_lo + lane; u < u_end; u += 32) {
                        if (!((prow[u >> 5] >> (u & 31)) & 1u)) continue;
                        uint32_t r, gg, bl;
                        patch_colour(u, v, r, gg, bl);
                        uint8_t* q8 = orow + 3 * u;
                        q8[0] = (uint8_t)r; q8[1] = (uint8_t)gg; q8[2] = (uint8_t)bl;
                    }
                }
            } else {
                const int nw = P.scar_size[0], nh = P.scar_size[1];
                const int32_t a0 = P.scar_fix[0], a1 = P.scar_fix[1], a2 = P.scar_fix[2];
                const int32_t a3 = P.scar_fix[3], a4 = P.scar_fix[4], a5 = P.scar_fix[5];
                for (int k = 0; k < npaste; ++k) {
                    const int pl = D.paste[k][0], pt = D.paste[k][1];
                    const int v_lo = max(0, gy0 - pt), v_hi = min(nh, gy0 + Rc - pt);
                    const int u_lo = max(0, -pl), u_end = min(nw, W - pl);
                    for (int v = v_lo + warp; v < v_hi; v += NWARP) {
                        uint8_t* orow = cb + (pt + v - gy0) * RB + pl * 3;
                        const int xr = a2 + v * a1, yr = a5 + v * a4;
                        for (int u = u_lo + lane; u < u_end; u += 32) {
                            const int xin = (xr + u * a0) >> 16, yin = (yr + u * a3) >> 16;
                            if ((unsigned)xin >= (unsigned)pw || (unsigned)yin >= (unsigned)ph) continue;   // alpha 0 outside the source
                            uint32_t r, gg, bl;
                            patch_colour(xin, yin, r, gg, bl);
                            uint8_t* q8 = orow + 3 * u;
                            q8[0] = (uint8_t)r; q8[1] = (uint8_t)gg; q8[2] = (uint8_t)bl;
                        }
                    }
                    __syncthreads();                          // later pastes overwrite earlier ones
                }
            }
            __syncthreads();
        };

        // sum `nv` partial values over the cluster: partials -> xchg slot -> barrier -> every CTA adds all ranks
        auto cluster_total = [&](const uint32_t* part, int nv) __attribute__((always_inline)) {
            if (tid == 0)
                for (int i = 0; i < nv; ++i) xchg[xs * 8 + i] = part[i];
            cluster_sync();
            if (tid < nv) {
                uint32_t t = 0;
                if (C > 1) {
                    for (int r = 0; r < C; ++r) t += dsmem_ld_u32(&xchg[xs * 8 + tid], (uint32_t)r);
                } else {
                    t = xchg[xs * 8 + tid];
                }
                tot[tid] = t;
            }
            xs ^= 1;
            __syncthreads();
        };

        int q = 0;                                           // running chunk-load index of this image
        uint8_t* cb = nullptr;
        // ================= P1: RGB sums of x + patch sums -> similarity branch (dataset_generator.py:147-159) =================
        if (y != 0) {
            uint32_t s[6] = {0, 0, 0, 0, 0, 0};
            for (int ch = 0; ch < nch; ++ch, ++q) {
                cb = acquire(q);
                const uint4* w4 = reinterpret_cast<const uint4*>(cb);
                const int nv = (int)(g.chunk_bytes / 48);                 // 48 B = 16 pixels = 3 vectors
                for (int i = tid; i < nv; i += NT) {
                    const uint4 A = w4[3 * i], B2 = w4[3 * i + 1], C2 = w4[3 * i + 2];
                    const uint32_t w[12] = {A.x, A.y, A.z, A.w, B2.x, B2.y, B2.z, B2.w, C2.x, C2.y, C2.z, C2.w};
                    #pragma unroll
                    for (int qq = 0; qq < 4; ++qq) {
                        const uint32_t w0 = w[3 * qq], w1 = w[3 * qq + 1], w2 = w[3 * qq + 2];
                        s[0] = __dp4a(w0, 0x01000001u, s[0]); s[0] = __dp4a(w1, 0x00010000u, s[0]); s[0] = __dp4a(w2, 0x00000100u, s[0]);
                        s[1] = __dp4a(w0, 0x00000100u, s[1]); s[1] = __dp4a(w1, 0x01000001u, s[1]); s[1] = __dp4a(w2, 0x00010000u, s[1]);
                        s[2] = __dp4a(w0, 0x00010000u, s[2]); s[2] = __dp4a(w1, 0x00000100u, s[2]); s[2] = __dp4a(w2, 0x01000001u, s[2]);
                    }
                }
                if (!resident) __syncthreads();              // buffer may be refilled by the load after next
            }
            if (P.src_variant != 2) {
                for (int v = (int)rank + C * warp; v < v_hiw; v += C * NWARP) {
                    const uint8_t* q8 = dbase + v * dpitch;
                    for (int u = lane; u < u_hi; u += 32) {
                        s[3] += __ldg(q8 + 3 * u); s[4] += __ldg(q8 + 3 * u + 1); s[5] += __ldg(q8 + 3 * u + 2);
                    }
                }
            }
            block_sum_vec<6>(s, red);
            cluster_total(s, 6);
            if (tid == 0) {
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
            __syncthreads();
            if (resident) paste_chunk(cb, part_y0);
        } else if (resident) {
            cb = acquire(q);
        }

        // ================= P2: ops ahead of Contrast + sum of L (ImageEnhance.Contrast mean of the current image) =================
        int cmean = -1;
        if (cpos >= 0) {
            uint32_t ls[1] = {0};
            for (int ch = 0; ch < nch; ++ch) {
                if (!resident) {
                    cb = acquire(q); ++q;
                    if (y != 0) paste_chunk(cb, part_y0 + ch * Rc);
                }
                uint2* c2 = reinterpret_cast<uint2*>(cb);
                for (int ui = tid; ui < g.units_chunk; ui += NT) {
                    Unit8 u;
                    unit_unpack(u, c2[3 * ui], c2[3 * ui + 1], c2[3 * ui + 2]);
                    if (cpos > 0) {
                        unit_apply(u, ops[0], fac[0]);
                        if (cpos > 1) unit_apply(u, ops[1], fac[1]);
                        if (resident) unit_pack(u, c2[3 * ui], c2[3 * ui + 1], c2[3 * ui + 2]);
                    }
                    ls[0] += unit_luma_sum(u);
                }
                if (!resident) __syncthreads();
            }
            block_sum_vec<1>(ls, red);
            cluster_total(ls, 1);
            cmean = (int)((2ull * tot[0] + (uint32_t)npix) / (2ull * (uint32_t)npix));   // int(sum / count + 0.5)
        }

        // ================= P3: Contrast + the remaining ops, stores =================
        const float cf = (cpos >= 0) ? fac[cpos] : 1.0f;
        const bool cclip = blend_needs_clip(cf);
        const int first_post = (cpos >= 0) ? cpos + 1 : 0;
        uint8_t* obase = out_u8 ? out_u8 + (size_t)b * d.out_u8_stride : nullptr;
        for (int ch = 0; ch < nch; ++ch) {
            const int gy0 = part_y0 + ch * Rc;
            if (!resident) {
                cb = acquire(q); ++q;
                if (y != 0) paste_chunk(cb, gy0);
            }
            const uint2* c2 = reinterpret_cast<const uint2*>(cb);
            for (int ui = tid; ui < g.units_chunk; ui += NT) {
                Unit8 u;
                unit_unpack(u, c2[3 * ui], c2[3 * ui + 1], c2[3 * ui + 2]);
                if (cpos >= 0) {
                    if (!resident && cpos > 0) {
                        unit_apply(u, ops[0], fac[0]);
                        if (cpos > 1) unit_apply(u, ops[1], fac[1]);
                    }
                    if (cclip) unit_contrast<true>(u, cf, cmean); else unit_contrast<false>(u, cf, cmean);
                }
                if (first_post < nops) {
                    unit_apply(u, ops[first_post], fac[first_post]);
                    if (first_post + 1 < nops) unit_apply(u, ops[first_post + 1], fac[first_post + 1]);
                }
                const size_t p0 = (size_t)gy0 * W + (size_t)ui * 8;     // dense row-major rows: unit -> pixel offset
                if (obase) {
                    uint2 a, bq, c;
                    unit_pack(u, a, bq, c);
                    uint2* o = reinterpret_cast<uint2*>(obase + p0 * 3);
                    o[0] = a; o[1] = bq; o[2] = c;
                }
                if (out_norm) {
                    // ToTensor + Normalize (datasets.py:392-393): exact table per channel, NCHW planes
                    float m[24];
                    unit_scalars(u, m);
                    if (d.out_norm_dtype == SSAD_NORM_F32) {
                        const float* lut = reinterpret_cast<const float*>(smem + g.o_lut);
                        float* ob = reinterpret_cast<float*>(out_norm) + (size_t)b * 3 * npix;
                        #pragma unroll
                        for (int c = 0; c < 3; ++c) {
                            float4* o = reinterpret_cast<float4*>(ob + (size_t)c * npix + p0);
                            #pragma unroll
                            for (int qq = 0; qq < 2; ++qq) {
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
                            uint32_t w[4];
                            #pragma unroll
                            for (int qq = 0; qq < 4; ++qq) {
                                uint32_t lo = lut[c * 256 + magic_to_u8(m[3 * (2 * qq) + c])];
                                uint32_t hi = lut[c * 256 + magic_to_u8(m[3 * (2 * qq + 1) + c])];
                                w[qq] = lo | (hi << 16);
                            }
                            *reinterpret_cast<uint4*>(ob + (size_t)c * npix + p0) = make_uint4(w[0], w[1], w[2], w[3]);
                        }
                    }
                }
            }
            if (!resident) __syncthreads();
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
        __syncthreads();            // buffers, P and D are free for the next image
    }
    // peers may still be reading this CTA's exchange slots
    if (C > 1) { cluster_arrive(); cluster_wait(); }
}

static inline uint32_t up(uint32_t v, uint32_t a) { return (v + a - 1) / a * a; }

static bool make_geom(const SsadBatchDesc& d, bool aligned, int force_c, uint32_t chunk_budget, Geom& g) {
    g.H = d.height; g.W = d.width;
    if (g.W % 16 != 0) return false;
    g.row_bytes = g.W * 3;
    g.raw_pitch = aligned ? g.row_bytes : g.row_bytes + 16;
    const int raw_words = g.raw_pitch / 4;
    // TMA box: one segment of <= 256 words, or (image-level, whole rows) several equal segments
    g.nseg = 1;
    while (raw_words % g.nseg != 0 || raw_words / g.nseg > 256 || (raw_words / g.nseg) % 4 != 0) {
        if (++g.nseg > 64) return false;
    }
    if (g.nseg > 1 && (!aligned || d.src_width != d.width || d.src_pitch != 3ll * d.width)) return false;
    g.box_w = raw_words / g.nseg;
    // cluster size: the smallest that lets a CTA keep its rows resident at four CTAs per SM; else eight
    const uint32_t small = 50 * 1024, big = 200 * 1024;
    int C = 0;
    for (int c = 1; c <= MAXC; c *= 2) {
        if (force_c && c != force_c) continue;
        if (g.H % c) continue;
        if ((uint32_t)(g.H / c) * g.raw_pitch <= small || force_c) { C = c; break; }
    }
    if (!C) {
        if (g.H % MAXC) return false;
        C = MAXC;
    }
    g.C = C;
    g.rows_cta = g.H / C;
    if ((uint32_t)g.rows_cta * g.raw_pitch <= big && g.rows_cta <= 256) {
        g.resident = 1; g.Rc = g.rows_cta;
    } else {
        if (!aligned) return false;
        g.resident = 0;
        g.Rc = 0;
        for (int r = g.rows_cta; r >= 1; --r)
            if (g.rows_cta % r == 0 && (uint32_t)r * g.raw_pitch <= chunk_budget && r <= 256) { g.Rc = r; break; }
        if (!g.Rc) return false;
    }
    g.nchunks = g.rows_cta / g.Rc;
    g.chunk_bytes = (uint32_t)g.Rc * g.row_bytes;
    g.raw_bytes = (uint32_t)g.Rc * g.raw_pitch;
    if (g.raw_bytes >= (1u << 20) || g.chunk_bytes % 48 != 0) return false;
    g.units_chunk = g.Rc * g.W / 8;
    g.poly_pitch = (uint32_t)(d.max_patch_w + 31) / 32u;
    g.poly_rows = (uint32_t)(d.max_patch_h < g.rows_cta ? d.max_patch_h : g.rows_cta);
    uint32_t o = 0;
    g.o_buf0 = o;   o += up(g.raw_bytes, 128);
    g.o_buf1 = o;   o += g.resident ? 0u : up(g.raw_bytes, 128);
    g.o_poly = o;   o += up(g.poly_rows * g.poly_pitch * 4u, 16);
    g.o_rowpre = o; o += (d.mask_mode == SSAD_MASK_ALL) ? 16u : up(((uint32_t)g.H + 1u) * 4u, 16);
    g.o_lut = o;    o += (d.out_norm_dtype == SSAD_NORM_F32) ? 3072u : (d.out_norm_dtype == SSAD_NORM_BF16 ? 1536u : 0u);
    g.o_red = o;    o += (uint32_t)NWARP * 8u * 4u;
    g.o_par = o;    o += up((uint32_t)sizeof(SsadImageParams), 16);
    g.o_dec = o;    o += up((uint32_t)sizeof(Dec), 16);
    g.o_xchg = o;   o += 2u * 8u * 4u;
    g.o_tot = o;    o += 8u * 4u;
    g.o_bar = o;    o += 16u;
    g.o_pgn = o;    o += up((uint32_t)sizeof(Polygon), 16);
    g.total = o;
    return true;
}

}  // namespace v3
}  // namespace ssad

// ------------------------------------------------------------------ host launch
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn v3_encode_fn() {
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

// returns SSAD_OK when launched, SSAD_EUNSUPPORTED when the caller should use another kernel
int ssad_launch_v3(const SsadBatchDesc& d, const SsadImageParams* params, const uint8_t* src, const uint8_t* donor_bank,
                   const uint32_t* mask_bits, const void* norm_lut, uint8_t* out_u8, void* out_norm, int64_t* y_out,
                   SsadDebug* debug, cudaStream_t stream) {
    using namespace ssad::v3;
    Geom g;
    const bool aligned = d.crop_align16 != 0;
    int force_c = 0;
    if (const char* fc = getenv("SSAD_V3_C")) force_c = atoi(fc);
    uint32_t chunk_budget = 48 * 1024;                      // streaming mode: bytes per chunk buffer (two buffers per CTA)
    if (const char* ck = getenv("SSAD_V3_CHUNK_KB")) { int v = atoi(ck); if (v > 0) chunk_budget = (uint32_t)v * 1024u; }
    if (!make_geom(d, aligned, force_c, chunk_budget, g)) return SSAD_EUNSUPPORTED;
    if ((reinterpret_cast<uintptr_t>(src) & 15) || (d.src_pitch & 15) || (d.src_stride & 15) || (d.src_width % 4)) return SSAD_EUNSUPPORTED;
    int dev = 0, max_smem = 0, sms = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return ssad_set_cuda_error(e);
    cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int smem_bytes = (int)g.total + 128;
    if (smem_bytes > max_smem) return SSAD_EUNSUPPORTED;
    EncodeTiledFn enc = v3_encode_fn();
    if (!enc) return SSAD_EUNSUPPORTED;
    CUtensorMap tmap;
    cuuint64_t gdim[4];
    cuuint64_t gstr[3];
    if (g.nseg == 1) {
        // rows as u32 words; the innermost extent is padded up to whole 16-byte groups so that the aligned superset
        // of any window is addressable (reads past the row end stay inside the pitch / are zero-filled)
        gdim[0] = (cuuint64_t)((d.src_width * 3 + 15) / 16 * 4);
        if ((long long)gdim[0] * 4 > d.src_pitch) gdim[0] = (cuuint64_t)(d.src_pitch / 4);
        gdim[1] = 1;
        gstr[0] = (cuuint64_t)d.src_pitch;
    } else {
        gdim[0] = (cuuint64_t)g.box_w;
        gdim[1] = (cuuint64_t)g.nseg;
        gstr[0] = (cuuint64_t)g.box_w * 4;
    }
    gdim[2] = (cuuint64_t)d.src_height; gdim[3] = (cuuint64_t)d.batch;
    gstr[1] = (cuuint64_t)d.src_pitch; gstr[2] = (cuuint64_t)d.src_stride;
    cuuint32_t box[4] = {(cuuint32_t)g.box_w, (cuuint32_t)g.nseg, (cuuint32_t)g.Rc, 1};
    cuuint32_t estr[4] = {1, 1, 1, 1};
    CUresult cr = enc(&tmap, CU_TENSOR_MAP_DATA_TYPE_UINT32, 4, const_cast<uint8_t*>(src), gdim, gstr, box, estr,
                      CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                      CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (cr != CUDA_SUCCESS) return SSAD_EUNSUPPORTED;
    e = cudaFuncSetAttribute(generate_v3_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes);
    if (e != cudaSuccess) return ssad_set_cuda_error(e);

    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.blockDim = dim3(NT, 1, 1);
    cfg.dynamicSmemBytes = (size_t)smem_bytes;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = (unsigned)g.C; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    // resident clusters: ask the occupancy calculator (shared memory, registers, GPC shapes)
    int nclusters = 0;
    cfg.gridDim = dim3((unsigned)(g.C * sms), 1, 1);
    e = cudaOccupancyMaxActiveClusters(&nclusters, generate_v3_kernel, &cfg);
    if (e != cudaSuccess || nclusters < 1) {
        cudaGetLastError();
        int per_sm = 0;
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, generate_v3_kernel, NT, (size_t)smem_bytes);
        nclusters = per_sm * sms / g.C;
    }
    if (const char* fr = getenv("SSAD_V3_CLUSTERS")) { int v = atoi(fr); if (v > 0) nclusters = v; }
    if (nclusters > d.batch) nclusters = d.batch;
    if (nclusters < 1) nclusters = 1;
    cfg.gridDim = dim3((unsigned)(g.C * nclusters), 1, 1);
    e = cudaLaunchKernelEx(&cfg, generate_v3_kernel, d, g, tmap, params, src, donor_bank, mask_bits, norm_lut, out_u8, out_norm,
                           y_out, debug);
    if (e != cudaSuccess) return ssad_set_cuda_error(e);
    e = cudaGetLastError();
    if (e != cudaSuccess) return ssad_set_cuda_error(e);
    return SSAD_OK;
}
