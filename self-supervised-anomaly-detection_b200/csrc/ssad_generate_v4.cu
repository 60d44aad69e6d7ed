// ssad_generate, B200 path v4: a pipeline of barrier-free streaming kernels over the whole batch.
//
// PretextTaskDataset.__getitem__ (datasets.py:209-394) has two whole-image dependencies -- the mean colour of x for
// check_color_similarity (dataset_generator.py:147-159) and the mean of L for ImageEnhance.Contrast -- and everything
// else is pointwise.  Instead of keeping an image inside one SM / cluster and synchronising its CTAs at those two
// points (v3, cluster kernel: half the issue slots were lost at barriers), the dependencies sit at KERNEL boundaries:
//
//   decide     one CTA per image     mask-window prefix -> coords_map[randint] centres -> container clamp ->
//                                    rect2poly bit rows; ColorJitter plan; label after the coverage rule
//   warp       (RandomAffine only)   x = NEAREST affine gather of the source, materialised once (+ RGB sums of x)
//   rgbsum     units, no barriers    RGB sums of x (defect labels, not warped)
//   patchsum   8 CTAs per image      RGB sums of the would-be cut (variants cut / average)
//   similar    one thread per image  cosine > 0.99 -> brightness branch; flat colours
//   luma       units, no barriers    paste on the fly -> ops ahead of Contrast -> sum of L
//   final      units, no barriers    paste on the fly -> whole ColorJitter chain -> uint8 HWC / normalised NCHW stores
//
// A unit = 8 consecutive pixels of a row = 24 bytes = 3 x 64-bit coalesced loads per thread (byte-misaligned crop
// windows: 4 loads + funnel shifts); the three streaming kernels have no shared memory and no barriers, every warp
// runs on its own, partial sums go to per-image accumulators with one atomic per warp.  Any image size works the
// same way (nothing has to fit an SM), work items are 512-unit slices of the flattened (image, unit) space, so the
// grid balances to within one slice.  HBM sees x up to three times (L2 absorbs part of it: 126 MB) -- the kernels
// are issue-bound, not bandwidth-bound, so trading bytes for barrier-free issue is the right way round.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "ssad_unit8.cuh"

int ssad_set_cuda_error(cudaError_t e);

namespace ssad {
namespace v4 {

using namespace u8x;

constexpr int NT = 256;
constexpr int UPT = 2;                 // units per thread per work item
constexpr int ITEM_UNITS = NT * UPT;

struct ImgDec {
    int32_t y, npaste, n_coords, sim;
    int32_t paste[5][2];
    int32_t flat;
    uint32_t flat_rgb[3];
    float bf1, bf2;
    int32_t nops, cpos;
    int32_t ops[3];
    float fac[3];
    int32_t avg_rgb[3];
    int32_t warped;
    double cosine;
    uint32_t sums[8];                  // 0-2 RGB of x, 3-5 RGB of the cut, 6 sum of L
};

struct Shape {
    int H, W, upr, upi;                // units per row / per image
    int upr_shift, upi_shift;          // log2 when a power of two, else -1
    long long total_units;
    int nitems;
    uint32_t poly_pitch, poly_rows;
    size_t o_dec, o_poly, o_x, total;  // workspace offsets
};

__device__ __forceinline__ void split(int g_hi_lo, int div, int shift, int& q, int& r) {
    if (shift >= 0) { q = g_hi_lo >> shift; r = g_hi_lo & (div - 1); }
    else { q = g_hi_lo / div; r = g_hi_lo - q * div; }
}

// 24 bytes at p (any alignment) -> 6 words
__device__ __forceinline__ void load24(const uint8_t* __restrict__ p, uint32_t (&w)[6]) {
    const uintptr_t a = reinterpret_cast<uintptr_t>(p);
    const uint32_t mis = (uint32_t)(a & 7u);
    const uint2* q = reinterpret_cast<const uint2*>(a - mis);
    if (mis == 0) {
        const uint2 A = __ldg(q), B = __ldg(q + 1), C = __ldg(q + 2);
        w[0] = A.x; w[1] = A.y; w[2] = B.x; w[3] = B.y; w[4] = C.x; w[5] = C.y;
    } else {
        const uint2 A = __ldg(q), B = __ldg(q + 1), C = __ldg(q + 2), D = __ldg(q + 3);
        const uint32_t sh = (mis & 3u) * 8u;
        if (mis < 4) {
            w[0] = __funnelshift_r(A.x, A.y, sh); w[1] = __funnelshift_r(A.y, B.x, sh); w[2] = __funnelshift_r(B.x, B.y, sh);
            w[3] = __funnelshift_r(B.y, C.x, sh); w[4] = __funnelshift_r(C.x, C.y, sh); w[5] = __funnelshift_r(C.y, D.x, sh);
        } else {
            w[0] = __funnelshift_r(A.y, B.x, sh); w[1] = __funnelshift_r(B.x, B.y, sh); w[2] = __funnelshift_r(B.y, C.x, sh);
            w[3] = __funnelshift_r(C.x, C.y, sh); w[4] = __funnelshift_r(C.y, D.x, sh); w[5] = __funnelshift_r(D.x, D.y, sh);
        }
    }
}

// replace pixel p (0..7, compile time) of the unit's words
template <int PIX>
__device__ __forceinline__ void put_pixel(uint32_t (&w)[6], uint32_t r, uint32_t g, uint32_t b) {
    constexpr int b0 = 3 * PIX, b1 = b0 + 1, b2 = b0 + 2;
    constexpr uint32_t keep = 0x3210u;
    w[b0 >> 2] = __byte_perm(w[b0 >> 2], r, (keep & ~(0xFu << (4 * (b0 & 3)))) | (4u << (4 * (b0 & 3))));
    w[b1 >> 2] = __byte_perm(w[b1 >> 2], g, (keep & ~(0xFu << (4 * (b1 & 3)))) | (4u << (4 * (b1 & 3))));
    w[b2 >> 2] = __byte_perm(w[b2 >> 2], b, (keep & ~(0xFu << (4 * (b2 & 3)))) | (4u << (4 * (b2 & 3))));
}

struct Ctx {
    SsadBatchDesc d;
    Shape s;
    const SsadImageParams* params;
    const uint8_t* src;
    const uint8_t* donor_bank;
    const uint32_t* mask_bits;
    const void* norm_lut;
    uint8_t* out_u8;
    void* out_norm;
    int64_t* y_out;
    SsadDebug* debug;
    ImgDec* dec;
    uint32_t* polytab;
    uint8_t* xbuf;
};

// address of unit (row yy, first pixel x0) of image b's working image x
__device__ __forceinline__ const uint8_t* unit_ptr(const Ctx& c, int b, bool warped, const SsadImageParams& P, int yy, int x0) {
    if (warped) return c.xbuf + ((size_t)b * c.s.H + yy) * (size_t)(c.s.W * 3) + (size_t)x0 * 3;
    return c.src + (size_t)b * c.d.src_stride + (size_t)(P.crop_off[1] + yy) * c.d.src_pitch + (size_t)(P.crop_off[0] + x0) * 3;
}

// donor / cut source of image b (datasets.py:225-228, 253): patch pixel (u, v) = donor window pixel (rl + u, rt + v)
struct Donor {
    const uint8_t* base;       // patch pixel (0, 0)
    int pitch, u_hi, v_hi;     // u < u_hi, v < v_hi: inside the donor window and the donor image (PIL's crop pads with 0)
};
__device__ __forceinline__ Donor donor_of(const Ctx& c, int b, const SsadImageParams& P) {
    const uint8_t* dimg = c.src + (size_t)b * c.d.src_stride;
    int dpitch = (int)c.d.src_pitch, dH = c.d.src_height, dW = c.d.src_width;
    if (P.donor_idx >= 0) {
        dimg = c.donor_bank + (size_t)P.donor_idx * c.d.donor_stride;
        dpitch = (int)c.d.donor_pitch; dH = c.d.donor_height; dW = c.d.donor_width;
    }
    const int rl = P.rect[0], rt = P.rect[1], pw = P.rect[2], ph = P.rect[3];
    const int dox = P.donor_off[0], doy = P.donor_off[1];
    Donor D;
    D.pitch = dpitch;
    D.u_hi = min(min(pw, c.d.donor_win_w - rl), dW - dox - rl);
    D.v_hi = min(min(ph, c.d.donor_win_h - rt), dH - doy - rt);
    D.base = dimg + (size_t)(doy + rt) * dpitch + (size_t)(dox + rl) * 3;
    return D;
}

// colour of patch pixel (u, v) after the variant / similarity-brightness decisions (datasets.py:270-299)
__device__ __forceinline__ void patch_colour(const ImgDec& E, const Donor& dn, int u, int v, uint32_t& r, uint32_t& g, uint32_t& bl) {
    if (E.flat) { r = E.flat_rgb[0]; g = E.flat_rgb[1]; bl = E.flat_rgb[2]; return; }
    r = g = bl = 0;
    if (u < dn.u_hi && v < dn.v_hi) {
        const uint8_t* q = dn.base + (size_t)v * dn.pitch + 3 * u;
        r = __ldg(q); g = __ldg(q + 1); bl = __ldg(q + 2);
    }
    if (E.sim) {        // two successive ImageEnhance.Brightness, each re-quantised (datasets.py:298-299)
        r = brightness_u8(brightness_u8(r, E.bf1), E.bf2);
        g = brightness_u8(brightness_u8(g, E.bf1), E.bf2);
        bl = brightness_u8(brightness_u8(bl, E.bf1), E.bf2);
    }
}

// paste_patch / scar pastes (dataset_generator.py:268-275, datasets.py:340-355) applied to the 8 pixels of a unit
__device__ __forceinline__ void paste_unit(const Ctx& c, int b, const ImgDec& E, const SsadImageParams& P, int yy, int x0,
                                           uint32_t (&w)[6]) {
    const int pw = P.rect[2], ph = P.rect[3];
    if (E.y == 1) {
        const int pl = E.paste[0][0], pt = E.paste[0][1];
        const int v = yy - pt;
        if (v < 0 || v >= ph || x0 + 8 <= pl || x0 >= pl + pw) return;
        const uint32_t* prow = c.polytab + ((size_t)b * c.s.poly_rows + v) * c.s.poly_pitch;
        const Donor dn = donor_of(c, b, P);
        #define SSAD_PASTE1(PIX)                                                                              \
            { const int u = x0 + PIX - pl;                                                                    \
              if (u >= 0 && u < pw && ((__ldg(prow + (u >> 5)) >> (u & 31)) & 1u)) {                           \
                  uint32_t r, g, bl; patch_colour(E, dn, u, v, r, g, bl); put_pixel<PIX>(w, r, g, bl); } }
        SSAD_PASTE1(0) SSAD_PASTE1(1) SSAD_PASTE1(2) SSAD_PASTE1(3) SSAD_PASTE1(4) SSAD_PASTE1(5) SSAD_PASTE1(6) SSAD_PASTE1(7)
        #undef SSAD_PASTE1
    } else {
        const int nw = P.scar_size[0], nh = P.scar_size[1];
        bool any = false;
        for (int k = 0; k < E.npaste; ++k) {
            const int pl = E.paste[k][0], v = yy - E.paste[k][1];
            any |= !(v < 0 || v >= nh || x0 + 8 <= pl || x0 >= pl + nw);
        }
        if (!any) return;
        const Donor dn = donor_of(c, b, P);
        const int32_t a0 = P.scar_fix[0], a1 = P.scar_fix[1], a2 = P.scar_fix[2];
        const int32_t a3 = P.scar_fix[3], a4 = P.scar_fix[4], a5 = P.scar_fix[5];
        for (int k = 0; k < E.npaste; ++k) {                 // in order: later pastes overwrite earlier ones
            const int pl = E.paste[k][0], v = yy - E.paste[k][1];
            if (v < 0 || v >= nh || x0 + 8 <= pl || x0 >= pl + nw) continue;
            const int xr = a2 + v * a1, yr = a5 + v * a4;
            #define SSAD_PASTE2(PIX)                                                                          \
                { const int u = x0 + PIX - pl;                                                                \
                  if (u >= 0 && u < nw) {                                                                     \
                      const int xin = (xr + u * a0) >> 16, yin = (yr + u * a3) >> 16;                         \
                      if ((unsigned)xin < (unsigned)pw && (unsigned)yin < (unsigned)ph) {                     \
                          uint32_t r, g, bl; patch_colour(E, dn, xin, yin, r, g, bl); put_pixel<PIX>(w, r, g, bl); } } }
            SSAD_PASTE2(0) SSAD_PASTE2(1) SSAD_PASTE2(2) SSAD_PASTE2(3) SSAD_PASTE2(4) SSAD_PASTE2(5) SSAD_PASTE2(6) SSAD_PASTE2(7)
            #undef SSAD_PASTE2
        }
    }
}

__device__ __forceinline__ void rgb_sums(const uint32_t (&w)[6], uint32_t (&s)[3]) {
    #pragma unroll
    for (int q = 0; q < 2; ++q) {
        const uint32_t w0 = w[3 * q], w1 = w[3 * q + 1], w2 = w[3 * q + 2];
        s[0] = __dp4a(w0, 0x01000001u, s[0]); s[0] = __dp4a(w1, 0x00010000u, s[0]); s[0] = __dp4a(w2, 0x00000100u, s[0]);
        s[1] = __dp4a(w0, 0x00000100u, s[1]); s[1] = __dp4a(w1, 0x01000001u, s[1]); s[1] = __dp4a(w2, 0x00010000u, s[1]);
        s[2] = __dp4a(w0, 0x00010000u, s[2]); s[2] = __dp4a(w1, 0x00000100u, s[2]); s[2] = __dp4a(w2, 0x01000001u, s[2]);
    }
}

// add per-thread partial sums to image b's accumulators: one atomic per warp when the warp works on one image
template <int NV>
__device__ __forceinline__ void warp_accumulate(uint32_t (&v)[NV], int b, bool active, ImgDec* dec, int first) {
    const uint32_t FULLM = 0xFFFFFFFFu;
    const int b0 = __shfl_sync(FULLM, b, 0);
    if (__all_sync(FULLM, b == b0)) {
        if (!__any_sync(FULLM, active)) return;
        #pragma unroll
        for (int i = 0; i < NV; ++i) {
            uint32_t t = active ? v[i] : 0u;
            for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(FULLM, t, o);
            if ((threadIdx.x & 31) == 0 && t) atomicAdd(&dec[b0].sums[first + i], t);
        }
    } else if (active) {
        #pragma unroll
        for (int i = 0; i < NV; ++i)
            if (v[i]) atomicAdd(&dec[b].sums[first + i], v[i]);
    }
}

// ------------------------------------------------------------------ decide: one CTA per image
__global__ void __launch_bounds__(128)
decide_kernel(const __grid_constant__ Ctx c) {
    extern __shared__ __align__(16) uint32_t rowpre[];       // H + 1 words
    __shared__ SsadImageParams P;
    __shared__ Polygon G;
    __shared__ int s_paste[5][2];
    const int tid = threadIdx.x, b = blockIdx.x;
    const SsadBatchDesc& d = c.d;
    const int H = c.s.H, W = c.s.W;
    if (tid < (int)(sizeof(SsadImageParams) / 4))
        reinterpret_cast<uint32_t*>(&P)[tid] = reinterpret_cast<const uint32_t*>(c.params + b)[tid];
    __syncthreads();
    const int cx0 = P.crop_off[0], cy0 = P.crop_off[1];
    const uint32_t* mrow0 = nullptr;
    if (d.mask_mode != SSAD_MASK_ALL)
        mrow0 = c.mask_bits + (d.mask_mode == SSAD_MASK_PER_IMAGE ? (size_t)b * d.mask_stride_words : 0);
    const bool defect = P.label != 0;
    const int pw = P.rect[2], ph = P.rect[3];
    if (defect && d.mask_mode != SSAD_MASK_ALL) {
        // mask window: per-row popcounts -> prefix (coords_map is row-major, datasets.py:263-264)
        for (int r = tid; r < H; r += blockDim.x)
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
    const int n_coords = (d.mask_mode == SSAD_MASK_ALL) ? H * W : (defect ? (int)rowpre[H] : 0);
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
            s_paste[tid][0] = pl; s_paste[tid][1] = pt;
        }
        if (y == 1 && tid == 32) polygon_setup(G, P.poly_xy, P.poly_n, ph);
        __syncthreads();
        if (y == 1) {
            // rect2poly bit rows (dataset_generator.py:63-100): one thread per canvas row
            for (int v = tid; v < ph && v < (int)c.s.poly_rows; v += blockDim.x) {
                uint32_t* row = c.polytab + ((size_t)b * c.s.poly_rows + v) * c.s.poly_pitch;
                for (int j = 0; j < (int)c.s.poly_pitch; ++j) row[j] = 0;
                polygon_fill_row(G, v, pw, row);
            }
        }
    }
    if (tid == 0) {
        ImgDec E;
        memset(&E, 0, sizeof(E));
        E.y = y; E.npaste = npaste; E.n_coords = n_coords;
        for (int k = 0; k < npaste; ++k) { E.paste[k][0] = s_paste[k][0]; E.paste[k][1] = s_paste[k][1]; }
        E.bf1 = P.bright_f[0]; E.bf2 = P.bright_f[1];
        E.warped = P.warp_en != 0;
        // ColorJitter plan (datasets.py:391): ops in fn_idx order, position of Contrast
        E.cpos = -1;
        for (int st = 0; st < 4; ++st) {
            int fn = P.jit_order[st];
            if (fn < 3 && ((P.jit_en >> fn) & 1)) {
                if (fn == 1) E.cpos = E.nops;
                E.ops[E.nops] = fn; E.fac[E.nops] = P.jit_f[fn]; ++E.nops;
            }
        }
        c.dec[b] = E;
        if (c.y_out) c.y_out[b] = y;
    }
}

// ------------------------------------------------------------------ work items: 512-unit slices of (image, unit)
#define SSAD_FOR_UNITS(c)                                                                                        \
    for (int item = blockIdx.x; item < (c).s.nitems; item += gridDim.x)                                          \
        for (int k_ = 0; k_ < UPT; ++k_)

struct UnitPos { int b, yy, x0; bool valid; };
__device__ __forceinline__ UnitPos unit_pos(const Ctx& c, int item, int k) {
    UnitPos p;
    const long long g = (long long)item * ITEM_UNITS + k * NT + threadIdx.x;
    p.valid = g < c.s.total_units;
    const long long gg = p.valid ? g : c.s.total_units - 1;
    int u;
    if (c.s.upi_shift >= 0) { p.b = (int)(gg >> c.s.upi_shift); u = (int)(gg & (c.s.upi - 1)); }
    else { p.b = (int)(gg / c.s.upi); u = (int)(gg - (long long)p.b * c.s.upi); }
    int cu;
    split(u, c.s.upr, c.s.upr_shift, p.yy, cu);
    p.x0 = cu * 8;
    return p;
}

// ---- RandomAffine(3, scale=(1.05, 1.1)) NEAREST, fill 0 (datasets.py:220-222): materialise x (+ its RGB sums)
__global__ void __launch_bounds__(NT)
warp_kernel(const __grid_constant__ Ctx c) {
    SSAD_FOR_UNITS(c) {
        const UnitPos p = unit_pos(c, item, k_);
        const SsadImageParams& P = c.params[p.b];
        const bool act = p.valid && P.warp_en != 0;
        uint32_t s[3] = {0, 0, 0};
        if (act) {
            const int W = c.s.W, H = c.s.H;
            const int32_t a0 = P.warp_fix[0], a1 = P.warp_fix[1], a2 = P.warp_fix[2];
            const int32_t a3 = P.warp_fix[3], a4 = P.warp_fix[4], a5 = P.warp_fix[5];
            const uint8_t* sbase = c.src + (size_t)p.b * c.d.src_stride + (size_t)P.crop_off[1] * c.d.src_pitch + (size_t)P.crop_off[0] * 3;
            const int spitch = (int)c.d.src_pitch;
            const int xr = a2 + p.yy * a1 + p.x0 * a0, yr = a5 + p.yy * a4 + p.x0 * a3;
            uint32_t w[6] = {0, 0, 0, 0, 0, 0};
            #define SSAD_GATHER(PIX)                                                                          \
                { const int xin = (xr + PIX * a0) >> 16, yin = (yr + PIX * a3) >> 16;                         \
                  if ((unsigned)xin < (unsigned)W && (unsigned)yin < (unsigned)H) {                           \
                      const uint8_t* q = sbase + yin * spitch + xin * 3;                                      \
                      put_pixel<PIX>(w, __ldg(q), __ldg(q + 1), __ldg(q + 2)); } }
            SSAD_GATHER(0) SSAD_GATHER(1) SSAD_GATHER(2) SSAD_GATHER(3) SSAD_GATHER(4) SSAD_GATHER(5) SSAD_GATHER(6) SSAD_GATHER(7)
            #undef SSAD_GATHER
            uint2* o = reinterpret_cast<uint2*>(c.xbuf + ((size_t)p.b * H + p.yy) * (size_t)(W * 3) + (size_t)p.x0 * 3);
            o[0] = make_uint2(w[0], w[1]); o[1] = make_uint2(w[2], w[3]); o[2] = make_uint2(w[4], w[5]);
            if (c.dec[p.b].y != 0) rgb_sums(w, s);
        }
        warp_accumulate<3>(s, p.b, act, c.dec, 0);
    }
}

// ---- RGB sums of x for check_color_similarity (defect labels, images that were not materialised by warp_kernel)
__global__ void __launch_bounds__(NT)
rgbsum_kernel(const __grid_constant__ Ctx c) {
    SSAD_FOR_UNITS(c) {
        const UnitPos p = unit_pos(c, item, k_);
        const ImgDec& E = c.dec[p.b];
        const bool act = p.valid && E.y != 0 && !E.warped;
        uint32_t s[3] = {0, 0, 0};
        if (act) {
            uint32_t w[6];
            load24(unit_ptr(c, p.b, false, c.params[p.b], p.yy, p.x0), w);
            rgb_sums(w, s);
        }
        warp_accumulate<3>(s, p.b, act, c.dec, 0);
    }
}

// ---- RGB sums of the would-be cut (dataset_generator.py:202-207 / 147-159): grid (8, batch)
__global__ void __launch_bounds__(128)
patchsum_kernel(const __grid_constant__ Ctx c) {
    const int b = blockIdx.y;
    const ImgDec& E = c.dec[b];
    const SsadImageParams& P = c.params[b];
    if (E.y == 0 || P.src_variant == 2) return;
    const Donor dn = donor_of(c, b, P);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
    uint32_t s[3] = {0, 0, 0};
    for (int v = blockIdx.x * nwarp + warp; v < dn.v_hi; v += gridDim.x * nwarp) {
        const uint8_t* q = dn.base + (size_t)v * dn.pitch;
        for (int u = lane; u < dn.u_hi; u += 32) {
            s[0] += __ldg(q + 3 * u); s[1] += __ldg(q + 3 * u + 1); s[2] += __ldg(q + 3 * u + 2);
        }
    }
    warp_accumulate<3>(s, b, true, c.dec, 3);
}

// ---- check_color_similarity + variant colours (datasets.py:270-299): one thread per image
__global__ void __launch_bounds__(128)
similar_kernel(const __grid_constant__ Ctx c) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= c.d.batch) return;
    ImgDec& E = c.dec[b];
    if (E.y == 0) return;
    const SsadImageParams& P = c.params[b];
    const uint32_t np = (uint32_t)(P.rect[2] * P.rect[3]);
    uint32_t xs[3] = {E.sums[0], E.sums[1], E.sums[2]};
    uint32_t psum[3] = {E.sums[3], E.sums[4], E.sums[5]};
    int flat = 0;
    uint32_t frgb[3] = {0, 0, 0};
    if (P.src_variant == 1) {           // colour = int(mean) (dataset_generator.py:202-207)
        flat = 1;
        for (int k = 0; k < 3; ++k) { frgb[k] = psum[k] / np; E.avg_rgb[k] = (int)frgb[k]; psum[k] = frgb[k] * np; }
    } else if (P.src_variant == 2) {    // random colour (dataset_generator.py:194-199)
        flat = 1;
        for (int k = 0; k < 3; ++k) { frgb[k] = P.rnd_rgb[k]; psum[k] = frgb[k] * np; }
    }
    double cs;
    const bool sim = similarity_branch(xs, (double)(c.s.H * c.s.W), psum, (double)np, cs);     // datasets.py:295 / 335
    if (flat && sim)                    // brightness of a flat patch is a flat patch
        for (int k = 0; k < 3; ++k) frgb[k] = brightness_u8(brightness_u8(frgb[k], E.bf1), E.bf2);
    E.flat = flat; E.sim = sim ? 1 : 0; E.cosine = cs;
    for (int k = 0; k < 3; ++k) E.flat_rgb[k] = frgb[k];
}

// ---- sum of L of the image ImageEnhance.Contrast sees: paste -> ops ahead of Contrast -> convert('L')
__global__ void __launch_bounds__(NT, 3)
luma_kernel(const __grid_constant__ Ctx c) {
    SSAD_FOR_UNITS(c) {
        const UnitPos p = unit_pos(c, item, k_);
        const ImgDec& E = c.dec[p.b];
        const bool act = p.valid && E.cpos >= 0;
        uint32_t s[1] = {0};
        if (act) {
            const SsadImageParams& P = c.params[p.b];
            uint32_t w[6];
            load24(unit_ptr(c, p.b, E.warped != 0, P, p.yy, p.x0), w);
            if (E.y != 0) paste_unit(c, p.b, E, P, p.yy, p.x0, w);
            Unit8 u;
            unit_unpack(u, make_uint2(w[0], w[1]), make_uint2(w[2], w[3]), make_uint2(w[4], w[5]));
            const int cpos = E.cpos;
            if (cpos > 0) {
                unit_apply(u, E.ops[0], E.fac[0]);
                if (cpos > 1) unit_apply(u, E.ops[1], E.fac[1]);
            }
            s[0] = unit_luma_sum(u);
        }
        warp_accumulate<1>(s, p.b, act, c.dec, 6);
    }
}

// ---- paste -> ColorJitter chain -> stores (datasets.py:391-393)
__global__ void __launch_bounds__(NT, 3)
final_kernel(const __grid_constant__ Ctx c) {
    const int npix = c.s.H * c.s.W;
    SSAD_FOR_UNITS(c) {
        const UnitPos p = unit_pos(c, item, k_);
        if (!p.valid) continue;
        const ImgDec& E = c.dec[p.b];
        const SsadImageParams& P = c.params[p.b];
        uint32_t w[6];
        load24(unit_ptr(c, p.b, E.warped != 0, P, p.yy, p.x0), w);
        if (E.y != 0) paste_unit(c, p.b, E, P, p.yy, p.x0, w);
        Unit8 u;
        unit_unpack(u, make_uint2(w[0], w[1]), make_uint2(w[2], w[3]), make_uint2(w[4], w[5]));
        const int nops = E.nops, cpos = E.cpos;
        int cmean = -1;
        int next = 0;
        if (cpos >= 0) {
            if (cpos > 0) {
                unit_apply(u, E.ops[0], E.fac[0]);
                if (cpos > 1) unit_apply(u, E.ops[1], E.fac[1]);
            }
            // int(mean + 0.5) with mean = sum / count in float64 (ImageEnhance.py Contrast) == integer rounding
            cmean = (int)((2ull * E.sums[6] + (uint32_t)npix) / (2ull * (uint32_t)npix));
            const float cf = E.fac[cpos];
            if (blend_needs_clip(cf)) unit_contrast<true>(u, cf, cmean); else unit_contrast<false>(u, cf, cmean);
            next = cpos + 1;
        }
        if (next < nops) {
            unit_apply(u, E.ops[next], E.fac[next]);
            if (next + 1 < nops) unit_apply(u, E.ops[next + 1], E.fac[next + 1]);
        }
        const size_t p0 = (size_t)p.yy * c.s.W + p.x0;
        if (c.out_u8) {
            uint2 a, bq, cq;
            unit_pack(u, a, bq, cq);
            uint2* o = reinterpret_cast<uint2*>(c.out_u8 + (size_t)p.b * c.d.out_u8_stride + p0 * 3);
            o[0] = a; o[1] = bq; o[2] = cq;
        }
        if (c.out_norm) {
            // ToTensor + Normalize (datasets.py:392-393): exact table per channel, NCHW planes
            float m[24];
            unit_scalars(u, m);
            if (c.d.out_norm_dtype == SSAD_NORM_F32) {
                const float* lut = reinterpret_cast<const float*>(c.norm_lut);
                float* ob = reinterpret_cast<float*>(c.out_norm) + (size_t)p.b * 3 * npix;
                #pragma unroll
                for (int ch = 0; ch < 3; ++ch) {
                    float4* o = reinterpret_cast<float4*>(ob + (size_t)ch * npix + p0);
                    #pragma unroll
                    for (int qq = 0; qq < 2; ++qq) {
                        float4 v;
                        v.x = __ldg(lut + ch * 256 + magic_to_u8(m[3 * (4 * qq + 0) + ch]));
                        v.y = __ldg(lut + ch * 256 + magic_to_u8(m[3 * (4 * qq + 1) + ch]));
                        v.z = __ldg(lut + ch * 256 + magic_to_u8(m[3 * (4 * qq + 2) + ch]));
                        v.w = __ldg(lut + ch * 256 + magic_to_u8(m[3 * (4 * qq + 3) + ch]));
                        o[qq] = v;
                    }
                }
            } else {
                const uint16_t* lut = reinterpret_cast<const uint16_t*>(c.norm_lut);
                uint16_t* ob = reinterpret_cast<uint16_t*>(c.out_norm) + (size_t)p.b * 3 * npix;
                #pragma unroll
                for (int ch = 0; ch < 3; ++ch) {
                    uint32_t wv[4];
                    #pragma unroll
                    for (int qq = 0; qq < 4; ++qq) {
                        uint32_t lo = __ldg(lut + ch * 256 + magic_to_u8(m[3 * (2 * qq) + ch]));
                        uint32_t hi = __ldg(lut + ch * 256 + magic_to_u8(m[3 * (2 * qq + 1) + ch]));
                        wv[qq] = lo | (hi << 16);
                    }
                    *reinterpret_cast<uint4*>(ob + (size_t)ch * npix + p0) = make_uint4(wv[0], wv[1], wv[2], wv[3]);
                }
            }
        }
        if (c.debug && p.yy == 0 && p.x0 == 0) {
            SsadDebug o;
            memset(&o, 0, sizeof(o));
            o.n_coords = E.n_coords;
            o.label_out = E.y;
            o.contrast_mean = cmean;
            if (E.y != 0) {
                o.sim = E.sim;
                o.cosine = E.cosine;
                for (int k = 0; k < E.npaste; ++k) { o.paste[k][0] = E.paste[k][0]; o.paste[k][1] = E.paste[k][1]; }
                for (int k = 0; k < 3; ++k) o.avg_rgb[k] = E.avg_rgb[k];
            }
            c.debug[p.b] = o;
        }
    }
}

static int ilog2_exact(int v) {
    if (v <= 0 || (v & (v - 1))) return -1;
    int s = 0;
    while ((1 << s) < v) ++s;
    return s;
}

static bool make_shape(const SsadBatchDesc& d, Shape& s) {
    s.H = d.height; s.W = d.width;
    if (s.W % 8 != 0) return false;
    s.upr = s.W / 8;
    s.upi = s.upr * s.H;
    s.upr_shift = ilog2_exact(s.upr);
    s.upi_shift = ilog2_exact(s.upi);
    s.total_units = (long long)d.batch * s.upi;
    const long long nitems = (s.total_units + ITEM_UNITS - 1) / ITEM_UNITS;
    if (nitems > 0x7FFFFFFFll / ITEM_UNITS) return false;
    s.nitems = (int)nitems;
    s.poly_pitch = (uint32_t)(d.max_patch_w + 31) / 32u;
    s.poly_rows = (uint32_t)d.max_patch_h;
    size_t o = 0;
    s.o_dec = o;  o += ((size_t)d.batch * sizeof(ImgDec) + 255) / 256 * 256;
    s.o_poly = o; o += ((size_t)d.batch * s.poly_rows * s.poly_pitch * 4 + 255) / 256 * 256;
    s.o_x = o;    o += d.any_warp ? ((size_t)d.batch * s.H * s.W * 3 + 255) / 256 * 256 : 0;
    s.total = o;
    return true;
}

}  // namespace v4
}  // namespace ssad

size_t ssad_v4_workspace(const SsadBatchDesc& d) {
    ssad::v4::Shape s;
    if (!ssad::v4::make_shape(d, s)) return 0;
    return s.total;
}

// returns SSAD_OK when launched, SSAD_EUNSUPPORTED when the caller should use another kernel
int ssad_launch_v4(const SsadBatchDesc& d, const SsadImageParams* params, const uint8_t* src, const uint8_t* donor_bank,
                   const uint32_t* mask_bits, const void* norm_lut, uint8_t* out_u8, void* out_norm, int64_t* y_out,
                   SsadDebug* debug, void* workspace, size_t workspace_bytes, cudaStream_t stream) {
    using namespace ssad::v4;
    Ctx c;
    memset(&c, 0, sizeof(c));
    if (!make_shape(d, c.s)) return SSAD_EUNSUPPORTED;
    if ((reinterpret_cast<uintptr_t>(src) & 15) || (d.src_pitch & 7) || (d.src_stride & 7)) return SSAD_EUNSUPPORTED;
    if (!workspace || workspace_bytes < c.s.total) return SSAD_EWORKSPACE;
    if (reinterpret_cast<uintptr_t>(workspace) & 255) return SSAD_EALIGN;
    c.d = d; c.params = params; c.src = src; c.donor_bank = donor_bank; c.mask_bits = mask_bits; c.norm_lut = norm_lut;
    c.out_u8 = out_u8; c.out_norm = out_norm; c.y_out = y_out; c.debug = debug;
    uint8_t* ws = static_cast<uint8_t*>(workspace);
    c.dec = reinterpret_cast<ImgDec*>(ws + c.s.o_dec);
    c.polytab = reinterpret_cast<uint32_t*>(ws + c.s.o_poly);
    c.xbuf = d.any_warp ? ws + c.s.o_x : nullptr;

    static int sms = 0;
    if (!sms) {
        int dev = 0;
        cudaError_t e = cudaGetDevice(&dev);
        if (e != cudaSuccess) return ssad_set_cuda_error(e);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    }
    const int B = d.batch;
    const size_t dec_smem = ((size_t)c.s.H + 1) * 4;
    if (dec_smem > 40 * 1024) return SSAD_EUNSUPPORTED;
    // streaming kernels: persistent grid, a multiple of the SM count (8 / 3 CTAs of 256 threads per SM)
    const int grid_light = min(c.s.nitems, sms * 8), grid_heavy = min(c.s.nitems, sms * 3);
    decide_kernel<<<B, 128, dec_smem, stream>>>(c);
    if (d.any_warp) warp_kernel<<<grid_light, NT, 0, stream>>>(c);
    rgbsum_kernel<<<grid_light, NT, 0, stream>>>(c);
    patchsum_kernel<<<dim3(8, B), 128, 0, stream>>>(c);
    similar_kernel<<<(B + 127) / 128, 128, 0, stream>>>(c);
    luma_kernel<<<grid_heavy, NT, 0, stream>>>(c);
    final_kernel<<<grid_heavy, NT, 0, stream>>>(c);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return ssad_set_cuda_error(e);
    return SSAD_OK;
}
