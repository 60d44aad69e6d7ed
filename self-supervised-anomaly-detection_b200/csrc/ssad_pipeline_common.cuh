// Device code of the generator pipeline (ssad_generate_pipeline.cu):
// per-image plan and tables, 8-pixel-unit stages (tables, Saturation, luma / RGB sums), paste on the fly, the two unit loops.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "ssad_device.cuh"
#include "ssad_sm100.cuh"
#include "ssad_unit8.cuh"

namespace ssad {
namespace v6 {

using namespace u8x;

constexpr int NT = 256;
constexpr int NWARP = NT / 32;
// table after Saturation: index = bits(2^23 + t, rounded toward zero) - 0x4B000000 + TB_OFF.  t < 0 lands in the
// binade below 2^23 (half steps): offsets down to -2 * 255 (kSatMax - 1); t < 255 * kSatMax above.  Valid for Saturation
// factors <= kSatMax = 2: the whole range of torchvision's ColorJitter(saturation <= 1).
constexpr int TB_OFF = 512;
constexpr int TB_SIZE = TB_OFF + 512;                  // t < 255 * kSatMax
constexpr float kSatMax = 2.0f;

// one stage program: bytes -> bytes
enum { PK_COPY = 0, PK_BYTELUT = 1, PK_SAT = 2 };
enum { SA_UNPACK = 0, SA_LUT = 1 };
enum { SB_PACK = 0, SB_CLAMP = 1, SB_LUT = 2 };
struct Prog {
    int32_t kind, sa, sb;
    float fs;                           // Saturation factor
};
struct alignas(16) Tables {             // shared memory, per program
    uint8_t a[256];                     // SA_LUT: run(v) of the per-channel run ahead of Saturation (a byte table: random
                                        // lookups meet two words per bank at most)
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
        for (int v = tid; v < 256; v += NT) T.a[v] = (uint8_t)run_ops((uint32_t)v, ops, fac, i0, is, mean);
    if (sb == SB_LUT)
        for (int j = tid; j < TB_SIZE; j += NT) {
            const int t = j - TB_OFF;
            const uint32_t v = (uint32_t)min(max(t, 0), 255);       // Blend.c clip (negative half steps -> 0)
            T.b[j] = (uint8_t)run_ops(v, ops, fac, is + 1, i1, mean);
        }
}

// ------------------------------------------------------------------ small helpers
// 24 bytes at any alignment -> 6 words, split in two for software pipelining: raw 8-byte-aligned loads now, the byte
// shift later
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
                u.v[ch][k] = pack2(__uint_as_float(kMagicBits | lds_u8(ta + i0)), __uint_as_float(kMagicBits | lds_u8(ta + i1)));
            }
        }
    }
}

// ImageEnhance.Color on a unit, result left unclipped as RZ(2^23 + t)
__device__ __forceinline__ void unit_saturation_raw(Unit8& u, float f) {
    const f32x2 MAG = pack2(kMagic, kMagic);
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
            // alpha * (v - L), rounded on its own as Blend.c rounds it: two scalar mul.rn -- ptxas contracts a packed
            // mul.rn.f32x2 followed by a packed add / sub into one FFMA2 (a single rounding), which differs from Pillow
            // whenever alpha (v - L) lies within an ulp of an integer
            float d0, d1;
            unpack2(add2_rn(u.v[ch][k], NML), d0, d1);
            const f32x2 p = pack2(__fmul_rn(f, d0), __fmul_rn(f, d1));
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
    if (G.sa == SA_LUT) { m[0] = magic_from_u8(T.a[r]); m[1] = magic_from_u8(T.a[g]); m[2] = magic_from_u8(T.a[b]); }
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

}  // namespace v6
}  // namespace ssad

