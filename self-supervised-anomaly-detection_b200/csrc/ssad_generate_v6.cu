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

namespace ssad {
namespace v6 {

using namespace u8x;

constexpr int NT = 256;
constexpr int NWARP = NT / 32;
constexpr int MAXC = 8;
// table after Saturation: index = bits(2^23 + t, rounded toward zero) - 0x4B000000 + TB_OFF.  t < 0 lands in the
// binade below 2^23 (half steps): offsets down to -2 * 64; t < 255 * kSatMax above.  Valid for Saturation factors <= kSatMax.
constexpr int TB_OFF = 128;
constexpr int TB_SIZE = TB_OFF + 320;                  // t < 255 * kSatMax
constexpr float kSatMax = 1.25f;

struct Geom {
    int H, W, C, rows_cta;
    int upr, upr_shift;                 // units per row (log2 when a power of two, else -1)
    uint32_t cov_pitch, cov_rows;       // coverage canvas: words per row, rows
    uint32_t canvas_pitch;              // colour canvas (global): words per row
    size_t canvas_stride, o_canvas, ws_total;   // words per image; workspace offsets / size in bytes
    uint32_t o_rowpre, o_cov, o_flags, o_lut, total;
};

// one stage program: bytes -> bytes
enum { PK_COPY = 0, PK_BYTELUT = 1, PK_SAT = 2 };
enum { SA_UNPACK = 0, SA_LUT = 1 };
enum { SB_PACK = 0, SB_CLAMP = 1, SB_LUT = 2 };
struct Prog {
    int32_t kind, sa, sb;
    float fs;                           // Saturation factor
};
struct Tables {                         // shared memory, per program
    uint32_t a[256];                    // SA_LUT: magic float (2^23 + run(v)) of the per-channel run ahead of Saturation
    uint8_t ab[256];                    // PK_BYTELUT: run(v)
    uint8_t b[TB_SIZE];                 // SB_LUT: run(clip(t)) of the per-channel run after Saturation
};

// per-image plan, one copy per CTA (shared memory)
struct Plan {
    int32_t y, npaste, n_coords, sim;
    int32_t paste[5][2];
    int32_t bw, bh;                     // canvas size: (pw, ph) polygon | (nw, nh) rotated scar
    int32_t flat;
    uint32_t flat_rgb[3];
    int32_t avg_rgb[3];
    float bf1, bf2;
    int32_t nops, cpos, cmean;
    int32_t ops[3];
    float fac[3];
    int32_t s_in_pre;                   // Saturation precedes Contrast: pass 1 stores its result
    Prog pre, fin;                      // ops ahead of Contrast | what pass 2 still has to do
    double cosine;
};

// ------------------------------------------------------------------ per-channel ops, scalar (table fill)
__device__ __forceinline__ uint32_t op_brightness(uint32_t v, float f) {       // ImageEnhance.Brightness: blend(0, v, f)
    const float m = magic_from_u8(v);
    return magic_to_u8(f > 1.0f ? magic_brightness<true>(m, f) : magic_brightness<false>(m, f));
}
__device__ __forceinline__ uint32_t op_contrast(uint32_t v, float f, int mean) {  // ImageEnhance.Contrast: blend(mean, v, f)
    const float m = magic_from_u8(v), mA = magic_from_u8((uint32_t)mean), aF = (float)mean;
    return magic_to_u8(blend_needs_clip(f) ? magic_blend<true>(m, mA, aF, f) : magic_blend<false>(m, mA, aF, f));
}
// run of per-channel ops ops[i0 .. i1) applied to v
__device__ __forceinline__ uint32_t run_ops(uint32_t v, const int32_t* ops, const float* fac, int i0, int i1, int mean) {
    for (int i = i0; i < i1; ++i) v = (ops[i] == 0) ? op_brightness(v, fac[i]) : op_contrast(v, fac[i], mean);
    return v;
}

// program for ops[i0 .. i1) (at most one Saturation inside) + its tables
__device__ inline void build_prog(Prog& G, Tables& T, const int32_t* ops, const float* fac, int i0, int i1, int mean, int tid) {
    int is = -1;
    for (int i = i0; i < i1; ++i) if (ops[i] == 2) is = i;
    if (i1 <= i0) { if (tid == 0) { G.kind = PK_COPY; G.sa = G.sb = 0; G.fs = 1.0f; } return; }
    if (is < 0) {
        if (tid == 0) { G.kind = PK_BYTELUT; G.sa = G.sb = 0; G.fs = 1.0f; }
        for (int v = tid; v < 256; v += NT) T.ab[v] = (uint8_t)run_ops((uint32_t)v, ops, fac, i0, i1, mean);
        return;
    }
    const float fs = fac[is];
    const int sa = (is > i0) ? SA_LUT : SA_UNPACK;
    const int sb = (is + 1 < i1) ? SB_LUT : SB_CLAMP;          // (the clamp form also serves factors <= 1)
    if (tid == 0) { G.kind = PK_SAT; G.sa = sa; G.sb = sb; G.fs = fs; }
    if (sa == SA_LUT)
        for (int v = tid; v < 256; v += NT) T.a[v] = kMagicBits | run_ops((uint32_t)v, ops, fac, i0, is, mean);
    if (sb == SB_LUT)
        for (int j = tid; j < TB_SIZE; j += NT) {
            const int t = j - TB_OFF;
            const uint32_t v = (uint32_t)min(max(t, 0), 255);       // Blend.c clip (negative half steps -> 0)
            T.b[j] = (uint8_t)run_ops(v, ops, fac, is + 1, i1, mean);
        }
}

// ------------------------------------------------------------------ small helpers
__device__ __forceinline__ void st_cluster_u32(const void* local_ptr, uint32_t rank, uint32_t v) {
    uint32_t remote;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(smem_u32(local_ptr)), "r"(rank));
    asm volatile("st.shared::cluster.u32 [%0], %1;" ::"r"(remote), "r"(v) : "memory");
}

__device__ __forceinline__ f32x2 fma2_rn(f32x2 a, f32x2 b, f32x2 c) {
    f32x2 r;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
    return r;
}

// 24 bytes at p (any alignment) -> 6 words.  nc = read-only path (source images); plain loads for the x buffer.
template <bool NC>
__device__ __forceinline__ uint2 ld8(const uint2* q) {
    if (NC) return __ldg(q);
    return *q;
}
template <bool NC>
__device__ __forceinline__ void load24(const uint8_t* __restrict__ p, uint32_t (&w)[6]) {
    const uintptr_t a = reinterpret_cast<uintptr_t>(p);
    const uint32_t mis = (uint32_t)(a & 7u);
    const uint2* q = reinterpret_cast<const uint2*>(a - mis);
    if (mis == 0) {
        const uint2 A = ld8<NC>(q), B = ld8<NC>(q + 1), C = ld8<NC>(q + 2);
        w[0] = A.x; w[1] = A.y; w[2] = B.x; w[3] = B.y; w[4] = C.x; w[5] = C.y;
    } else {
        const uint2 A = ld8<NC>(q), B = ld8<NC>(q + 1), C = ld8<NC>(q + 2), D = ld8<NC>(q + 3);
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

// the same load split in two for software pipelining: raw 8-byte-aligned loads now, the byte shift later
struct Raw { uint2 a, b, c, d; };
__device__ __forceinline__ void ld_raw(const uint8_t* q /* 8-byte aligned */, bool four, bool nc, Raw& r) {
    const uint2* p = reinterpret_cast<const uint2*>(q);
    r.d = make_uint2(0u, 0u);
    (void)nc;        // one flavour of load: the x buffer is written by this kernel, so no ld.global.nc
    r.a = p[0]; r.b = p[1]; r.c = p[2]; if (four) r.d = p[3];
}
__device__ __forceinline__ void raw_to_words(const Raw& r, uint32_t mis, uint32_t (&w)[6]) {
    if (mis == 0) {
        w[0] = r.a.x; w[1] = r.a.y; w[2] = r.b.x; w[3] = r.b.y; w[4] = r.c.x; w[5] = r.c.y;
    } else {
        const uint32_t sh = (mis & 3u) * 8u;
        if (mis < 4) {
            w[0] = __funnelshift_r(r.a.x, r.a.y, sh); w[1] = __funnelshift_r(r.a.y, r.b.x, sh); w[2] = __funnelshift_r(r.b.x, r.b.y, sh);
            w[3] = __funnelshift_r(r.b.y, r.c.x, sh); w[4] = __funnelshift_r(r.c.x, r.c.y, sh); w[5] = __funnelshift_r(r.c.y, r.d.x, sh);
        } else {
            w[0] = __funnelshift_r(r.a.y, r.b.x, sh); w[1] = __funnelshift_r(r.b.x, r.b.y, sh); w[2] = __funnelshift_r(r.b.y, r.c.x, sh);
            w[3] = __funnelshift_r(r.c.x, r.c.y, sh); w[4] = __funnelshift_r(r.c.y, r.d.x, sh); w[5] = __funnelshift_r(r.d.x, r.d.y, sh);
        }
    }
}

// replace pixel PIX (0..7, compile time) of the unit's words
template <int PIX>
__device__ __forceinline__ void put_pixel(uint32_t (&w)[6], uint32_t r, uint32_t g, uint32_t b) {
    constexpr int b0 = 3 * PIX, b1 = b0 + 1, b2 = b0 + 2;
    constexpr uint32_t keep = 0x3210u;
    w[b0 >> 2] = __byte_perm(w[b0 >> 2], r, (keep & ~(0xFu << (4 * (b0 & 3)))) | (4u << (4 * (b0 & 3))));
    w[b1 >> 2] = __byte_perm(w[b1 >> 2], g, (keep & ~(0xFu << (4 * (b1 & 3)))) | (4u << (4 * (b1 & 3))));
    w[b2 >> 2] = __byte_perm(w[b2 >> 2], b, (keep & ~(0xFu << (4 * (b2 & 3)))) | (4u << (4 * (b2 & 3))));
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

// sum of Convert.c's L over the 8 pixels of a unit, straight from the packed bytes:
// L = (19595 R + 38470 G + 7471 B + 0x8000) >> 16 with two 16x8-bit dot products per pixel.
__device__ __forceinline__ uint32_t dp2a_lo(uint32_t a16x2, uint32_t b8x4, uint32_t c) {
    uint32_t d;
    asm("dp2a.lo.u32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a16x2), "r"(b8x4), "r"(c));
    return d;
}
__device__ __forceinline__ uint32_t dp2a_hi(uint32_t a16x2, uint32_t b8x4, uint32_t c) {
    uint32_t d;
    asm("dp2a.hi.u32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a16x2), "r"(b8x4), "r"(c));
    return d;
}
__device__ __forceinline__ void luma_words(const uint32_t (&w)[6], uint32_t (&L)[8]) {
    constexpr uint32_t kRG = (38470u << 16) | 19595u, kB0 = 7471u;          // pixel at bytes 0,1,2 of a word
    constexpr uint32_t k0R = (19595u << 16), kGB = (7471u << 16) | 38470u;  // pixel at bytes 1,2,3 of a word
    #pragma unroll
    for (int q = 0; q < 2; ++q) {
        const uint32_t w0 = w[3 * q], w1 = w[3 * q + 1], w2 = w[3 * q + 2];
        const uint32_t p1 = __byte_perm(w0, w1, 0x0543);     // pixel 1: byte 3 of w0, bytes 0,1 of w1
        const uint32_t p2 = __byte_perm(w1, w2, 0x0432);     // pixel 2: bytes 2,3 of w1, byte 0 of w2
        L[4 * q + 0] = dp2a_hi(kB0, w0, dp2a_lo(kRG, w0, 0x8000u)) >> 16;
        L[4 * q + 1] = dp2a_hi(kB0, p1, dp2a_lo(kRG, p1, 0x8000u)) >> 16;
        L[4 * q + 2] = dp2a_hi(kB0, p2, dp2a_lo(kRG, p2, 0x8000u)) >> 16;
        L[4 * q + 3] = dp2a_hi(kGB, w2, dp2a_lo(k0R, w2, 0x8000u)) >> 16;
    }
}
__device__ __forceinline__ uint32_t luma_sum_words(const uint32_t (&w)[6]) {
    uint32_t L[8];
    luma_words(w, L);
    return ((L[0] + L[1]) + (L[2] + L[3])) + ((L[4] + L[5]) + (L[6] + L[7]));
}
// ... leaving out the pixels whose bit is set in `skip`
__device__ __forceinline__ uint32_t luma_sum_words_skip(const uint32_t (&w)[6], uint32_t skip) {
    uint32_t L[8];
    luma_words(w, L);
    uint32_t s = 0;
    #pragma unroll
    for (int p = 0; p < 8; ++p) s += ((skip >> p) & 1u) ? 0u : L[p];
    return s;
}

// sum NV values over the CTA; result valid in every thread
template <int NV>
__device__ __forceinline__ void block_sum_vec(uint32_t (&v)[NV], uint32_t* red /* NWARP * 8 + 8 words */) {
    #pragma unroll
    for (int i = 0; i < NV; ++i)
        for (int o = 16; o > 0; o >>= 1) v[i] += __shfl_xor_sync(0xFFFFFFFFu, v[i], o);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) {
        #pragma unroll
        for (int i = 0; i < NV; ++i) red[(threadIdx.x >> 5) * 8 + i] = v[i];
    }
    __syncthreads();
    if (threadIdx.x < NV) {
        uint32_t t = 0;
        #pragma unroll
        for (int w = 0; w < NWARP; ++w) t += red[w * 8 + threadIdx.x];
        red[NWARP * 8 + threadIdx.x] = t;
    }
    __syncthreads();
    #pragma unroll
    for (int i = 0; i < NV; ++i) v[i] = red[NWARP * 8 + i];
}

// ------------------------------------------------------------------ donor / cut source, patch colours
// donor / cut source of image b (datasets.py:225-228, 253): patch pixel (u, v) = donor window pixel (rl + u, rt + v)
struct Donor {
    const uint8_t* base;       // patch pixel (0, 0)
    int pitch, u_hi, v_hi;     // u < u_hi, v < v_hi: inside the donor window and the donor image (PIL's crop pads with 0)
};
__device__ __forceinline__ Donor donor_of(const SsadBatchDesc& d, const uint8_t* src, const uint8_t* donor_bank, int b,
                                          const SsadImageParams& P) {
    const uint8_t* dimg = src + (size_t)b * d.src_stride;
    int dpitch = (int)d.src_pitch, dH = d.src_height, dW = d.src_width;
    if (P.donor_idx >= 0) {
        dimg = donor_bank + (size_t)P.donor_idx * d.donor_stride;
        dpitch = (int)d.donor_pitch; dH = d.donor_height; dW = d.donor_width;
    }
    const int rl = P.rect[0], rt = P.rect[1], pw = P.rect[2], ph = P.rect[3];
    const int dox = P.donor_off[0], doy = P.donor_off[1];
    Donor D;
    D.pitch = dpitch;
    D.u_hi = min(min(pw, d.donor_win_w - rl), dW - dox - rl);
    D.v_hi = min(min(ph, d.donor_win_h - rt), dH - doy - rt);
    D.base = dimg + (size_t)(doy + rt) * dpitch + (size_t)(dox + rl) * 3;
    return D;
}

// colour of patch pixel (u, v) after the variant / similarity-brightness decisions (datasets.py:270-299)
struct Colour {
    int flat, sim;
    uint32_t frgb[3];
    float bf1, bf2;
};
__device__ __forceinline__ void patch_colour(const Colour& K, const Donor& dn, int u, int v, uint32_t& r, uint32_t& g, uint32_t& bl) {
    if (K.flat) { r = K.frgb[0]; g = K.frgb[1]; bl = K.frgb[2]; return; }
    r = g = bl = 0;
    if (u < dn.u_hi && v < dn.v_hi) {
        const uint8_t* q = dn.base + (size_t)v * dn.pitch + 3 * u;
        r = __ldg(q); g = __ldg(q + 1); bl = __ldg(q + 2);
    }
    if (K.sim) {        // two successive ImageEnhance.Brightness, each re-quantised (datasets.py:298-299)
        r = brightness_u8(brightness_u8(r, K.bf1), K.bf2);
        g = brightness_u8(brightness_u8(g, K.bf1), K.bf2);
        bl = brightness_u8(brightness_u8(bl, K.bf1), K.bf2);
    }
}

// ------------------------------------------------------------------ paste on the fly
// 8 coverage bits of canvas row `crow` starting at canvas x = u0 (u0 may be negative or run past bw: those bits are 0)
__device__ __forceinline__ uint32_t cov_bits8(const uint32_t* __restrict__ crow, int u0, int pitch_words) {
    const int lo = max(u0, 0);
    const int wi = lo >> 5;
    uint32_t a = crow[wi];
    uint32_t b = (wi + 1 < pitch_words) ? crow[wi + 1] : 0u;
    uint32_t bits = __funnelshift_r(a, b, lo & 31) & 0xFFu;          // bits for canvas x = lo .. lo + 7
    return (u0 < 0) ? (bits << (-u0)) & 0xFFu : bits;                 // pixel PIX <-> canvas x = u0 + PIX
}

// ------------------------------------------------------------------ stages on 8-pixel units
__device__ __forceinline__ uint32_t min_relu_s16x2(uint32_t a, uint32_t b) {
    uint32_t d;
    asm("min.relu.s16x2 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b));
    return d;
}
__device__ __forceinline__ uint32_t lds_u32(uint32_t addr) {
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr));
    return v;
}
__device__ __forceinline__ uint32_t lds_u8(uint32_t addr) {
    uint32_t v;
    asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(addr));
    return v;
}

// stage A: bytes -> magic floats, plain (PRMT) or through the table of the run ahead of Saturation
template <int SA>
__device__ __forceinline__ void stage_in(Unit8& u, const uint32_t (&w)[6], uint32_t ta /* smem address of Tables::a */) {
    if (SA == SA_UNPACK) {
        unit_unpack(u, make_uint2(w[0], w[1]), make_uint2(w[2], w[3]), make_uint2(w[4], w[5]));
    } else {
        #pragma unroll
        for (int k = 0; k < 4; ++k) {
            #pragma unroll
            for (int ch = 0; ch < 3; ++ch) {
                const int b0 = 6 * k + ch, b1 = 6 * k + 3 + ch;
                const uint32_t i0 = __byte_perm(w[b0 >> 2], 0, 0x4440 + (b0 & 3)), i1 = __byte_perm(w[b1 >> 2], 0, 0x4440 + (b1 & 3));
                u.v[ch][k] = pack2(__uint_as_float(lds_u32(ta + i0 * 4u)), __uint_as_float(lds_u32(ta + i1 * 4u)));
            }
        }
    }
}

// ImageEnhance.Color on a unit, result left unclipped as RZ(2^23 + t)
__device__ __forceinline__ void unit_saturation_raw(Unit8& u, float f) {
    const f32x2 F = pack2(f, f), MAG = pack2(kMagic, kMagic);
    #pragma unroll
    for (int k = 0; k < 4; ++k) {
        float r0, r1, g0, g1, b0, b1;
        unpack2(u.v[0][k], r0, r1); unpack2(u.v[1][k], g0, g1); unpack2(u.v[2][k], b0, b1);
        const uint32_t a0 = luma_bits(__float_as_uint(r0), __float_as_uint(g0), __float_as_uint(b0));
        const uint32_t a1 = luma_bits(__float_as_uint(r1), __float_as_uint(g1), __float_as_uint(b1));
        // -(2^23 + L): bit pattern 0xCB0000LL
        const f32x2 NML = pack2(__uint_as_float(__byte_perm(a0, 0xCB000000u, 0x7542)),
                                __uint_as_float(__byte_perm(a1, 0xCB000000u, 0x7542)));
        const f32x2 NLF = add2_rn(NML, MAG);                // -(float)L, exact
        #pragma unroll
        for (int ch = 0; ch < 3; ++ch) {
            const f32x2 p = mul2_rn(F, add2_rn(u.v[ch][k], NML));     // alpha * (v - L)
            f32x2 t;
            asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(t) : "l"(p), "l"(NLF));   // L + p (same rounding as Pillow's A + p)
            u.v[ch][k] = add2_rz(t, MAG);
        }
    }
}

// stage B: magic floats -> bytes
template <int SB>
__device__ __forceinline__ void stage_out(const Unit8& u, uint32_t (&w)[6], uint32_t tbc /* smem address of Tables::b + TB_OFF - 0x4B000000 */) {
    float m[24];
    unit_scalars(u, m);
    #pragma unroll
    for (int i = 0; i < 6; ++i) {
        const uint32_t q0 = __float_as_uint(m[4 * i]), q1 = __float_as_uint(m[4 * i + 1]);
        const uint32_t q2 = __float_as_uint(m[4 * i + 2]), q3 = __float_as_uint(m[4 * i + 3]);
        if (SB == SB_PACK) {
            w[i] = __byte_perm(__byte_perm(q0, q1, 0x0040), __byte_perm(q2, q3, 0x0040), 0x5410);
        } else if (SB == SB_CLAMP) {
            // low halves as s16: in range 0x00vv, above 0x01vv, negative 0xFFvv -> min(.., 255) then relu
            const uint32_t lo = min_relu_s16x2(__byte_perm(q0, q1, 0x5410), 0x00FF00FFu);
            const uint32_t hi = min_relu_s16x2(__byte_perm(q2, q3, 0x5410), 0x00FF00FFu);
            w[i] = __byte_perm(lo, hi, 0x6420);
        } else {
            const uint32_t a = lds_u8(q0 + tbc), b = lds_u8(q1 + tbc), c = lds_u8(q2 + tbc), d = lds_u8(q3 + tbc);
            w[i] = __byte_perm(__byte_perm(a, b, 0x0040), __byte_perm(c, d, 0x0040), 0x5410);
        }
    }
}

// bytes -> bytes through one 256-entry table
__device__ __forceinline__ void unit_bytelut(uint32_t (&w)[6], uint32_t tab /* smem address of Tables::ab */) {
    #pragma unroll
    for (int i = 0; i < 6; ++i) {
        const uint32_t x = w[i];
        const uint32_t a = lds_u8(tab + (x & 0xFFu)), b = lds_u8(tab + __byte_perm(x, 0, 0x4441));
        const uint32_t c = lds_u8(tab + __byte_perm(x, 0, 0x4442)), d = lds_u8(tab + (x >> 24));
        w[i] = __byte_perm(__byte_perm(a, b, 0x0040), __byte_perm(c, d, 0x0040), 0x5410);
    }
}

template <int SA, int SB>
__device__ __forceinline__ void unit_sat_prog(uint32_t (&w)[6], float fs, uint32_t ta, uint32_t tbc) {
    Unit8 u;
    stage_in<SA>(u, w, ta);
    unit_saturation_raw(u, fs);
    stage_out<SB>(u, w, tbc);
}

// the same program on one pixel (paste colours, correction pass) -- identical float sequence
__device__ __forceinline__ void pixel_prog(const Prog& G, const Tables& T, uint32_t& r, uint32_t& g, uint32_t& b) {
    if (G.kind == PK_COPY) return;
    if (G.kind == PK_BYTELUT) { r = T.ab[r]; g = T.ab[g]; b = T.ab[b]; return; }
    float m[3];
    if (G.sa == SA_LUT) { m[0] = __uint_as_float(T.a[r]); m[1] = __uint_as_float(T.a[g]); m[2] = __uint_as_float(T.a[b]); }
    else { m[0] = magic_from_u8(r); m[1] = magic_from_u8(g); m[2] = magic_from_u8(b); }
    const uint32_t acc = luma_bits(__float_as_uint(m[0]), __float_as_uint(m[1]), __float_as_uint(m[2]));
    const float nml = __uint_as_float(__byte_perm(acc, 0xCB000000u, 0x7542));
    const float nlf = __fadd_rn(nml, kMagic);
    uint32_t q[3];
    #pragma unroll
    for (int c = 0; c < 3; ++c) {
        const float p = __fmul_rn(G.fs, __fadd_rn(m[c], nml));
        q[c] = __float_as_uint(__fadd_rz(__fsub_rn(p, nlf), kMagic));
    }
    #pragma unroll
    for (int c = 0; c < 3; ++c) {
        const int t = (int)(q[c] - kMagicBits);
        if (G.sb == SB_LUT) q[c] = T.b[t + TB_OFF];
        else q[c] = (uint32_t)min(max(t, 0), 255);
    }
    r = q[0]; g = q[1]; b = q[2];
}

// program kinds the unit loops are specialised for
enum { LK_COPY = 0, LK_BYTELUT = 1, LK_SAT_UC = 2, LK_SAT_UL = 3, LK_SAT_LC = 4, LK_SAT_LL = 5 };
__device__ __forceinline__ int loop_kind(const Prog& G) {
    if (G.kind == PK_COPY) return LK_COPY;
    if (G.kind == PK_BYTELUT) return LK_BYTELUT;
    return LK_SAT_UC + (G.sa == SA_LUT ? 2 : 0) + (G.sb == SB_LUT ? 1 : 0);
}
struct ProgRegs {              // a program's constants in registers
    float fs;
    uint32_t ta, tab, tbc;     // shared-memory addresses: Tables::a, Tables::ab, Tables::b + TB_OFF - 0x4B000000
};
template <int LK>
__device__ __forceinline__ void apply_prog(uint32_t (&w)[6], const ProgRegs& R) {
    if (LK == LK_COPY) return;
    if (LK == LK_BYTELUT) { unit_bytelut(w, R.tab); return; }
    if (LK == LK_SAT_UC) unit_sat_prog<SA_UNPACK, SB_CLAMP>(w, R.fs, R.ta, R.tbc);
    if (LK == LK_SAT_UL) unit_sat_prog<SA_UNPACK, SB_LUT>(w, R.fs, R.ta, R.tbc);
    if (LK == LK_SAT_LC) unit_sat_prog<SA_LUT, SB_CLAMP>(w, R.fs, R.ta, R.tbc);
    if (LK == LK_SAT_LL) unit_sat_prog<SA_LUT, SB_LUT>(w, R.fs, R.ta, R.tbc);
}

struct Boxes {                 // paste geometry of one image, in registers inside the unit loops
    const uint32_t* cov;       // coverage bit rows of the canvas (shared memory)
    const uint32_t* canvas;    // packed (r, g, b, L) per canvas pixel (global, written by this CTA)
    int pitch, cpitch, bw, bh, npaste;
};

// union of the coverage of all pastes over the 8 pixels of a unit (pass 1: those pixels are replaced later)
__device__ __forceinline__ uint32_t covered_bits8(const Plan& D, const Boxes& B, uint32_t rowflag, int yy, int x0) {
    uint32_t m = 0;
    for (int k = 0; k < B.npaste; ++k) {
        if (!((rowflag >> k) & 1u)) continue;
        const int u0 = x0 - D.paste[k][0], v = yy - D.paste[k][1];
        if (u0 + 8 <= 0 || u0 >= B.bw) continue;
        m |= cov_bits8(B.cov + (size_t)v * B.pitch, u0, B.pitch);
    }
    return m;
}

// replace pixel PIX of the unit by the (r, g, b) bytes of a packed canvas word
template <int PIX>
__device__ __forceinline__ void put_pixel_c(uint32_t (&w)[6], uint32_t c) {
    constexpr int b0 = 3 * PIX, b1 = b0 + 1, b2 = b0 + 2;
    constexpr uint32_t keep = 0x3210u;
    w[b0 >> 2] = __byte_perm(w[b0 >> 2], c, (keep & ~(0xFu << (4 * (b0 & 3)))) | (4u << (4 * (b0 & 3))));
    w[b1 >> 2] = __byte_perm(w[b1 >> 2], c, (keep & ~(0xFu << (4 * (b1 & 3)))) | (5u << (4 * (b1 & 3))));
    w[b2 >> 2] = __byte_perm(w[b2 >> 2], c, (keep & ~(0xFu << (4 * (b2 & 3)))) | (6u << (4 * (b2 & 3))));
}

// paste_patch / scar pastes (dataset_generator.py:268-275, datasets.py:340-355) applied to the 8 pixels of a unit:
// covered pixels take the canvas colour; pastes in order, later ones overwrite earlier ones
__device__ __forceinline__ void paste_unit(const Plan& D, const Boxes& B, uint32_t rowflag, int yy, int x0, uint32_t (&w)[6]) {
    for (int k = 0; k < B.npaste; ++k) {
        if (!((rowflag >> k) & 1u)) continue;
        const int u0 = x0 - D.paste[k][0], v = yy - D.paste[k][1];
        if (u0 + 8 <= 0 || u0 >= B.bw) continue;
        const uint32_t bits = cov_bits8(B.cov + (size_t)v * B.pitch, u0, B.pitch);
        if (!bits) continue;
        const uint32_t* crow = B.canvas + (ptrdiff_t)v * B.cpitch + u0;
        #define SSAD_PASTE(PIX) if (bits & (1u << PIX)) put_pixel_c<PIX>(w, crow[PIX]);
        SSAD_PASTE(0) SSAD_PASTE(1) SSAD_PASTE(2) SSAD_PASTE(3) SSAD_PASTE(4) SSAD_PASTE(5) SSAD_PASTE(6) SSAD_PASTE(7)
        #undef SSAD_PASTE
    }
}

// u in [0, n) with 0 <= base + u * step < limit (int64: no wrap) -> [lo, hi] inclusive (empty: lo > hi)
__device__ __forceinline__ long long floor_div(long long a, long long b) {     // b > 0
    long long q = a / b;
    return (a % b < 0) ? q - 1 : q;
}
__device__ __forceinline__ void affine_span(long long base, long long step, long long limit, int& lo, int& hi) {
    if (step == 0) { if (!(base >= 0 && base < limit)) { lo = 1; hi = 0; } return; }
    long long l, h;
    if (step > 0) { l = -floor_div(base, step); h = floor_div(limit - 1 - base, step); }      // ceil(-base / step)
    else { const long long sp = -step; h = floor_div(base, sp); l = floor_div(base - limit, sp) + 1; }
    if (l > lo) lo = (int)min(l, (long long)0x3FFFFFFF);
    if (h < hi) hi = (int)max(h, (long long)-1);
}

// ------------------------------------------------------------------ the unit loops
// A CTA's slab = rows [r0, r0 + nrows) x upr units; thread t takes units t, t + NT, ... (row-major).
struct Slab {
    int r0, nunits, upr, tid;
    int dr, dc;                // NT units further = dr rows and dc units (with one carry at most)
    int rr0, cu0;              // this thread's first unit
};
__device__ __forceinline__ Slab make_slab(int r0, int nunits, int upr, int tid) {
    Slab S;
    S.r0 = r0; S.nunits = nunits; S.upr = upr; S.tid = tid;
    S.dr = NT / upr; S.dc = NT - S.dr * upr;
    S.rr0 = tid / upr; S.cu0 = tid - S.rr0 * upr;
    return S;
}
// rr / cu = row / unit-in-row of the current unit; lin = its dense index (rr * upr + cu = t, t + NT, ...)
#define SSAD_SLAB_LOOP(S)                                                                                     \
    for (int lin = (S).tid, rr = (S).rr0, cu = (S).cu0; lin < (S).nunits;                                     \
         lin += NT, rr += (S).dr, cu += (S).dc, rr += (cu >= (S).upr) ? 1 : 0, cu -= (cu >= (S).upr) ? (S).upr : 0)

// Software-pipelined walk over a thread's units: the loads of the next unit are issued before the current one is
// processed.  Inside the body: w[6] = the unit's bytes, lin = its dense index in the slab, rr = its row in the slab,
// yy / x0 = its image row / first pixel.  (Rows are 8-byte periodic: pitch % 8 == 0, so `mis` is one value per slab.)
#define SSAD_PIPELINED_LOOP(S, in0, in_pitch, in_nc)                                                          \
    const uint32_t mis_ = (uint32_t)(reinterpret_cast<uintptr_t>(in0) & 7u);                                  \
    const uint8_t* const in0a_ = (in0) - mis_;                                                                \
    Raw nx_;                                                                                                  \
    int nlin_ = (S).tid, nrr_ = (S).rr0, ncu_ = (S).cu0;                                                      \
    if (nlin_ < (S).nunits) ld_raw(in0a_ + nrr_ * (in_pitch) + ncu_ * 24, mis_ != 0, in_nc, nx_);             \
    while (nlin_ < (S).nunits)
#define SSAD_PIPELINED_STEP(S, in_pitch, in_nc)                                                               \
    const Raw cur_ = nx_;                                                                                     \
    const int lin = nlin_, rr = nrr_, yy = (S).r0 + nrr_, x0 = ncu_ * 8;                                      \
    nlin_ += NT; nrr_ += (S).dr; ncu_ += (S).dc;                                                              \
    if (ncu_ >= (S).upr) { ncu_ -= (S).upr; ++nrr_; }                                                         \
    if (nlin_ < (S).nunits) ld_raw(in0a_ + nrr_ * (in_pitch) + ncu_ * 24, mis_ != 0, in_nc, nx_);             \
    uint32_t w[6];                                                                                            \
    raw_to_words(cur_, mis_, w);

// pass 1: RGB sums of x; the program ahead of Contrast; sum of L over the pixels no paste replaces; keep the bytes
template <int LK>
__device__ __forceinline__ void pass1_loop(const Slab& S, const uint8_t* in, int in_pitch, bool in_nc, uint8_t* xb, int xpitch,
                                           bool store, bool need_rgb, bool need_luma, const ProgRegs& R, const Plan& D,
                                           const Boxes& BX, const uint8_t* flags, uint32_t (&acc)[7]) {
    uint32_t s0 = 0, s1 = 0, s2 = 0, sl = 0;
    const uint8_t* const in0 = in + (size_t)S.r0 * in_pitch;
    uint8_t* const xb0 = xb + (size_t)S.r0 * xpitch;           // dense rows: unit `lin` of the slab sits at lin * 24
    SSAD_PIPELINED_LOOP(S, in0, in_pitch, in_nc) {
        SSAD_PIPELINED_STEP(S, in_pitch, in_nc)
        if (need_rgb) {
            uint32_t s[3] = {s0, s1, s2};
            rgb_sums(w, s);
            s0 = s[0]; s1 = s[1]; s2 = s[2];
        }
        if (need_luma) {
            apply_prog<LK>(w, R);
            // sum of L of what Contrast will see, minus the pixels a paste replaces (their L comes from the canvas)
            const uint32_t rf = D.y != 0 ? (uint32_t)flags[rr] : 0u;
            const uint32_t skip = rf ? covered_bits8(D, BX, rf, yy, x0) : 0u;
            sl += skip ? luma_sum_words_skip(w, skip) : luma_sum_words(w);
            if (store) {                        // the x buffer keeps pre(x): pass 2 starts from there
                uint2* const o = reinterpret_cast<uint2*>(xb0 + lin * 24);
                o[0] = make_uint2(w[0], w[1]); o[1] = make_uint2(w[2], w[3]); o[2] = make_uint2(w[4], w[5]);
            }
        }
    }
    acc[0] += s0; acc[1] += s1; acc[2] += s2; acc[3] += sl;
}

// pass 2: paste on the fly, the rest of the chain, stores
template <int LK>
__device__ __forceinline__ void pass2_loop(const Slab& S, const uint8_t* in, int in_pitch, bool in_nc, const ProgRegs& R,
                                           const Plan& D, const Boxes& BX, const uint8_t* flags, int W, int npix,
                                           uint8_t* obase, void* onorm, int norm_dtype, const uint8_t* lutbase) {
    const uint8_t* const in0 = in + (size_t)S.r0 * in_pitch;
    const size_t pbase = (size_t)S.r0 * W;                     // dense outputs: unit `lin` starts at pixel pbase + lin * 8
    SSAD_PIPELINED_LOOP(S, in0, in_pitch, in_nc) {
        SSAD_PIPELINED_STEP(S, in_pitch, in_nc)
        if (D.y != 0) {
            const uint32_t rf = flags[rr];
            if (rf) paste_unit(D, BX, rf, yy, x0, w);
        }
        apply_prog<LK>(w, R);
        const size_t p0 = pbase + (size_t)lin * 8;
        if (obase) {
            uint2* o = reinterpret_cast<uint2*>(obase + p0 * 3);
            o[0] = make_uint2(w[0], w[1]); o[1] = make_uint2(w[2], w[3]); o[2] = make_uint2(w[4], w[5]);
        }
        if (onorm) {
            // ToTensor + Normalize (datasets.py:392-393): exact table per channel, NCHW planes
            uint32_t e[24];
            #pragma unroll
            for (int i = 0; i < 24; ++i) e[i] = (w[i >> 2] >> (8 * (i & 3))) & 0xFFu;      // byte i = pixel i / 3, channel i % 3
            if (norm_dtype == SSAD_NORM_F32) {
                const float* lut = reinterpret_cast<const float*>(lutbase);
                float* ob = reinterpret_cast<float*>(onorm);
                #pragma unroll
                for (int ch = 0; ch < 3; ++ch) {
                    float4* o = reinterpret_cast<float4*>(ob + (size_t)ch * npix + p0);
                    #pragma unroll
                    for (int qq = 0; qq < 2; ++qq) {
                        float4 v;
                        v.x = lut[ch * 256 + e[3 * (4 * qq + 0) + ch]];
                        v.y = lut[ch * 256 + e[3 * (4 * qq + 1) + ch]];
                        v.z = lut[ch * 256 + e[3 * (4 * qq + 2) + ch]];
                        v.w = lut[ch * 256 + e[3 * (4 * qq + 3) + ch]];
                        o[qq] = v;
                    }
                }
            } else {
                const uint16_t* lut = reinterpret_cast<const uint16_t*>(lutbase);
                uint16_t* ob = reinterpret_cast<uint16_t*>(onorm);
                #pragma unroll
                for (int ch = 0; ch < 3; ++ch) {
                    uint32_t wv[4];
                    #pragma unroll
                    for (int qq = 0; qq < 4; ++qq) {
                        uint32_t lo = lut[ch * 256 + e[3 * (2 * qq) + ch]];
                        uint32_t hi = lut[ch * 256 + e[3 * (2 * qq + 1) + ch]];
                        wv[qq] = lo | (hi << 16);
                    }
                    *reinterpret_cast<uint4*>(ob + (size_t)ch * npix + p0) = make_uint4(wv[0], wv[1], wv[2], wv[3]);
                }
            }
        }
    }
}

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
