// ssad_generate: the kernel pipeline (six kernels per chunk of the batch).
//
// Round 1's single-kernel design ran the whole of PretextTaskDataset.__getitem__ (datasets.py:209-394) for one image inside
// one cluster; ncu showed that kernel issue-bound with 39 % of every CTA's life spent in its short serial phases
// (geometry, decisions, canvas, tables) and a 1.7-wave tail (profiles/history/).  Here each phase is its own grid, sized
// for what it does, and the two passes over the pixels are pure streaming kernels -- no paste, mask or flag logic inside
// their loops:
//
//   plan     CTA / image      parameters -> centres -> container clamp -> paste boxes; rect2poly / rotated-scar coverage
//                             bit rows; the jitter plan and the tables of the ops ahead of Contrast; the image joins the
//                             work lists of the two passes, one list per loop kind.  Reads the mask bits, no pixel.
//   pass 1   CTA / slab       (RandomAffine gather) ; RGB sums of x ; ops ahead of Contrast ; sum of L of EVERY pixel ;
//                             RGB sums of the would-be cut -> atomicAdd into the image's accumulators.  Leaves in the
//                             x buffer what pass 2 starts from: pre(x) when Saturation precedes Contrast, else x itself
//                             when something will be pasted onto it.
//   canvas   NC CTAs / image  cosine > 0.99 decision, flat colours; the patch canvas (packed RGB + L); the pastes written
//                             INTO the x buffer, later pastes over earlier ones; sum of L(new) - L(old) -> atomicAdd
//   finish   CTA / image      Contrast mean -> tables of Contrast and the ops after it; labels, debug records
//   pass 2   CTA / slab       x buffer (or the source when nothing touched it) -> rest of the chain -> stores
//
// CTAs of the passes are ordered by loop kind (the lists), so an SM runs one or two loop bodies at a time: the six
// specialised bodies together are far larger than the instruction caches.
// The per-image state (plan, tables, accumulators), the coverage rows and the lists live in the caller's workspace.
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <mutex>

#include "ssad_pipeline_common.cuh"

int ssad_set_cuda_error(cudaError_t e);

namespace ssad {
namespace v7 {

using namespace v6;
using namespace u8x;

constexpr int NKIND = 6;         // LK_COPY .. LK_SAT_LL
constexpr int kCanvasRows = 512; // canvas rows one CTA of the canvas kernel can own

struct ImgState {                // per image, global memory
    Plan D;
    uint32_t acc[8];             // 0-2 RGB of x, 3 sum of L of pre(x), 4-6 RGB of the cut, 7 sum of L(new) - L(old) over the pastes
    int32_t sk[4];               // loop kind and position in that kind's list: pass 1 (kind NKIND: nothing to do), pass 2
    Tables TP, TF;
};
struct Sched {                   // images per loop kind in the two passes (zeroed before the plan kernel); [0][NKIND] counts
    uint32_t count[2][8];        // the images pass 1 skips
};
// One slot per image and pass, ordered by loop kind: what a CTA of a pass needs to start streaming.
struct alignas(16) Hdr {
    int32_t b;                   // image, -1: no work
    int32_t kind;                // LK_*
    int32_t flags;               // HF_*
    int32_t cx0, cy0;            // crop window
    float fs;                    // Saturation factor of the pass's program
    int32_t si;                  // source image (the record's src_index, or the image itself)
    uint32_t rows;               // rows [lo, hi) of the working image the pastes touch: lo | hi << 16 (0: none)
};
enum { HF_WARP = 1, HF_PRE = 2, HF_DEFECT = 4, HF_LUMA = 8 };

struct Geo7 {
    int H, W, upr, rows_slab, nslab, ncanvas, batch;
    uint32_t cov_pitch, cov_rows, canvas_pitch;
    size_t canvas_stride, cov_stride;                       // words per image
    size_t o_canvas, o_state, o_cov, o_sched, o_hdr, ws_total;   // byte offsets in the workspace (x buffer at 0)
    int fast;                                               // NT % upr == 0 and every slab holds NT * UPT units
    int nslab_shift;                                        // log2(nslab), or -1
    uint32_t inv_upr;                                       // ceil(2^16 / upr): tid / upr == tid * inv_upr >> 16 for tid < 256
    uint32_t smem_plan;                                     // dynamic shared memory of the plan kernel (row prefix)
    int src_indexed;                                        // records carry their source image (else: image b of the chunk)
};

// Programmatic dependent launch: every kernel lets its successor's CTAs start as soon as its own have all started
// (they fill the SMs the tail of this grid leaves idle), and waits for its predecessor's memory before its first read.
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

template <typename T>
__device__ __forceinline__ void copy_words(T* dst, const T* src, int tid) {
    static_assert(sizeof(T) % 4 == 0, "word copy");
    uint32_t* d = reinterpret_cast<uint32_t*>(dst);
    const uint32_t* s = reinterpret_cast<const uint32_t*>(src);
    for (int i = tid; i < (int)(sizeof(T) / 4); i += NT) d[i] = s[i];
}

// loop kind of the program build_prog makes of ops[i0 .. i1)
__device__ __forceinline__ int kind_of_ops(const int32_t* ops, int i0, int i1) {
    if (i1 <= i0) return LK_COPY;
    int is = -1;
    for (int i = i0; i < i1; ++i) if (ops[i] == 2) is = i;
    if (is < 0) return LK_BYTELUT;
    return LK_SAT_UC + ((is > i0) ? 2 : 0) + ((is + 1 < i1) ? 1 : 0);
}

// ------------------------------------------------------------------ unit loops
// Fast path: NT % upr == 0 (a thread keeps its unit column: the byte alignment is a loop invariant, pointers advance by
// constants) and the slab holds exactly NT * UPT units (no bounds test).  Every thread announces all its units to the
// L2 up front (prefetch.global.L2), then keeps one unit of loads in flight in registers.
#ifndef SSAD_V7_UPT_FAST
#define SSAD_V7_UPT_FAST 8
#endif
constexpr int UPT = SSAD_V7_UPT_FAST;
// resident CTAs per SM the pass kernels are compiled for (4 x 256 threads = 64 registers per thread)
// pixels per thread and round in the canvas kernel's paste loops (their loads are issued together)
#ifndef SSAD_V7_NB_POLY
#define SSAD_V7_NB_POLY 2
#endif
#ifndef SSAD_V7_NB_SCAR
#define SSAD_V7_NB_SCAR 2
#endif
#ifndef SSAD_V7_P1_MINB
#define SSAD_V7_P1_MINB 4
#endif
#ifndef SSAD_V7_P2_MINB
#define SSAD_V7_P2_MINB 4
#endif

#ifdef SSAD_V7_PREFETCH_L1
__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L1 [%0];" ::"l"(p)); }
#else
__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }
#endif

// sum of Convert.c's L over the 8 pixels of a unit: two 16x8-bit dot products per pixel, straight from the words
__device__ __forceinline__ uint32_t luma_sum_fast(const uint32_t (&w)[6]) {
    constexpr uint32_t kRG = (38470u << 16) | 19595u, kB0 = 7471u, kR3 = 19595u << 16, kGB = (7471u << 16) | 38470u;
    uint32_t s = 0;
    #pragma unroll
    for (int q = 0; q < 2; ++q) {
        const uint32_t w0 = w[3 * q], w1 = w[3 * q + 1], w2 = w[3 * q + 2];
        s += dp2a_hi(kB0, w0, dp2a_lo(kRG, w0, 0x8000u)) >> 16;        // bytes 0,1,2 of w0
        s += dp2a_lo(kGB, w1, dp2a_hi(kR3, w0, 0x8000u)) >> 16;        // byte 3 of w0, bytes 0,1 of w1
        s += dp2a_lo(kB0, w2, dp2a_hi(kRG, w1, 0x8000u)) >> 16;        // bytes 2,3 of w1, byte 0 of w2
        s += dp2a_hi(kGB, w2, dp2a_lo(kR3, w2, 0x8000u)) >> 16;        // bytes 1,2,3 of w2
    }
    return s;
}

__device__ __forceinline__ void store24(uint8_t* p, const uint32_t (&w)[6]) {
    uint2* const o = reinterpret_cast<uint2*>(p);
    o[0] = make_uint2(w[0], w[1]); o[1] = make_uint2(w[2], w[3]); o[2] = make_uint2(w[4], w[5]);
}

// what pass 1 does to one unit: sums of x, (store x), program ahead of Contrast, sum of L, (store pre(x))
template <int LK>
__device__ __forceinline__ void pass1_unit(uint32_t (&w)[6], uint8_t* o, bool store_raw, bool store_pre, bool need_rgb,
                                           bool need_luma, const ProgRegs& R, uint32_t& s0, uint32_t& s1, uint32_t& s2,
                                           uint32_t& sl) {
    if (need_rgb) {
        uint32_t s[3] = {s0, s1, s2};
        rgb_sums(w, s);
        s0 = s[0]; s1 = s[1]; s2 = s[2];
    }
    if (store_raw) store24(o, w);
    if (need_luma) {
        apply_prog<LK>(w, R);
        sl += luma_sum_fast(w);
        if (store_pre) store24(o, w);
    }
}

// ToTensor + Normalize (datasets.py:392-393) of one unit: exact table per channel, NCHW planes
__device__ __forceinline__ void store_norm(const uint32_t (&w)[6], void* onorm, int norm_dtype, const uint8_t* lutbase, int npix,
                                           size_t p0) {
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
                __stcs(o + qq, v);               // never re-read on the GPU: evict-first keeps the x buffer in the L2 (+0.5 %)
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
            __stcs(reinterpret_cast<uint4*>(ob + (size_t)ch * npix + p0), make_uint4(wv[0], wv[1], wv[2], wv[3]));
        }
    }
}

struct FastCol {                // a thread's unit column in the fast path
    const uint8_t* ip;          // 8-byte aligned load address of the current unit
    size_t istep;
    int rr0, cu, dr;
    uint32_t mis;
};
__device__ __forceinline__ FastCol fast_col(const uint8_t* in0, int in_pitch, const Geo7& g, int tid) {
    FastCol F;
    const int upr = g.upr;
    F.rr0 = (int)(((uint32_t)tid * g.inv_upr) >> 16); F.cu = tid - F.rr0 * upr; F.dr = g.rows_slab / UPT;   // NT / upr
    const uint8_t* ip = in0 + (size_t)F.rr0 * in_pitch + F.cu * 24;
    F.mis = (uint32_t)(reinterpret_cast<uintptr_t>(ip) & 7u);
    F.ip = ip - F.mis;
    F.istep = (size_t)F.dr * in_pitch;
    #pragma unroll
    for (int it = 1; it < UPT; ++it) prefetch_l2(F.ip + it * F.istep + 16);
    return F;
}

// F / nx: the thread's column and its first unit's loads, already in flight (fast_col + ld_raw)
template <int LK>
__device__ __forceinline__ void pass1_fast(FastCol& F, Raw& nx, uint8_t* xb0 /* slab row 0 */, int xpitch, bool store_raw,
                                           bool store_pre, bool need_rgb, bool need_luma, const ProgRegs& R, uint32_t (&acc)[7]) {
    uint8_t* op = xb0 + (size_t)F.rr0 * xpitch + F.cu * 24;
    const size_t ostep = (size_t)F.dr * xpitch;
    uint32_t s0 = 0, s1 = 0, s2 = 0, sl = 0;
    const bool four = F.mis != 0;
    #pragma unroll 2
    for (int it = 0; it < UPT; ++it) {
        const Raw cur = nx;
        F.ip += F.istep;
        if (it + 1 < UPT) ld_raw(F.ip, four, false, nx);
        uint32_t w[6];
        raw_to_words(cur, F.mis, w);
        pass1_unit<LK>(w, op, store_raw, store_pre, need_rgb, need_luma, R, s0, s1, s2, sl);
        op += ostep;
    }
    acc[0] += s0; acc[1] += s1; acc[2] += s2; acc[3] += sl;
}

template <int LK>
__device__ __forceinline__ void pass2_fast(FastCol& F, Raw& nx, int r0, int W, int npix, const ProgRegs& R, uint8_t* obase,
                                           void* onorm, int norm_dtype, const uint8_t* lutbase) {
    int p0 = (r0 + F.rr0) * W + F.cu * 8;                  // first pixel of the current unit inside the image
    const int pstep = F.dr * W;
    uint8_t* op = obase ? obase + (size_t)p0 * 3 : nullptr;
    const size_t ostep = (size_t)pstep * 3;
    const bool four = F.mis != 0;
    #pragma unroll 2
    for (int it = 0; it < UPT; ++it) {
        const Raw cur = nx;
        F.ip += F.istep;
        if (it + 1 < UPT) ld_raw(F.ip, four, false, nx);
        uint32_t w[6];
        raw_to_words(cur, F.mis, w);
        apply_prog<LK>(w, R);
        if (op) { store24(op, w); op += ostep; }
        if (onorm) store_norm(w, onorm, norm_dtype, lutbase, npix, (size_t)p0);
        p0 += pstep;
    }
}

// General path: any width that is a multiple of 8, ragged last slab.
template <int LK>
__device__ __forceinline__ void pass1_gen(const Slab& S, const uint8_t* in, int in_pitch, uint8_t* xb, int xpitch, bool store_raw,
                                          bool store_pre, bool need_rgb, bool need_luma, const ProgRegs& R, uint32_t (&acc)[7]) {
    uint32_t s0 = 0, s1 = 0, s2 = 0, sl = 0;
    const uint8_t* const in0 = in + (size_t)S.r0 * in_pitch;
    uint8_t* const xb0 = xb + (size_t)S.r0 * xpitch;           // dense rows: unit `lin` of the slab sits at lin * 24
    SSAD_PIPELINED_LOOP(S, in0, in_pitch, false) {
        SSAD_PIPELINED_STEP(S, in_pitch, false)
        (void)rr; (void)yy; (void)x0;
        pass1_unit<LK>(w, xb0 + (size_t)lin * 24, store_raw, store_pre, need_rgb, need_luma, R, s0, s1, s2, sl);
    }
    acc[0] += s0; acc[1] += s1; acc[2] += s2; acc[3] += sl;
}

template <int LK>
__device__ __forceinline__ void pass2_gen(const Slab& S, const uint8_t* in, int in_pitch, const ProgRegs& R, int W, int npix,
                                          uint8_t* obase, void* onorm, int norm_dtype, const uint8_t* lutbase) {
    const uint8_t* const in0 = in + (size_t)S.r0 * in_pitch;
    const size_t pbase = (size_t)S.r0 * W;                     // dense outputs: unit `lin` starts at pixel pbase + lin * 8
    SSAD_PIPELINED_LOOP(S, in0, in_pitch, false) {
        SSAD_PIPELINED_STEP(S, in_pitch, false)
        (void)rr; (void)yy; (void)x0;
        apply_prog<LK>(w, R);
        const size_t p0 = pbase + (size_t)lin * 8;
        if (obase) store24(obase + p0 * 3, w);
        if (onorm) store_norm(w, onorm, norm_dtype, lutbase, npix, p0);
    }
}

// the loop of pass 1 / pass 2 for one slab, by kind
#define SSAD_V7_BY_KIND(kind, CALL)                                                     \
    switch (kind) {                                                                     \
        case LK_COPY:    CALL(LK_COPY); break;                                          \
        case LK_BYTELUT: CALL(LK_BYTELUT); break;                                       \
        case LK_SAT_UC:  CALL(LK_SAT_UC); break;                                        \
        case LK_SAT_UL:  CALL(LK_SAT_UL); break;                                        \
        case LK_SAT_LC:  CALL(LK_SAT_LC); break;                                        \
        default:         CALL(LK_SAT_LL); break;                                        \
    }

// ================================================================ plan
__global__ void __launch_bounds__(NT)
plan_v7_kernel(const __grid_constant__ SsadBatchDesc d, const __grid_constant__ Geo7 g,
               const SsadImageParams* __restrict__ params, const uint32_t* __restrict__ mask_bits,
               ImgState* __restrict__ state, uint32_t* __restrict__ cov_all, Sched* __restrict__ sched) {
    extern __shared__ __align__(16) uint8_t smem[];
    __shared__ SsadImageParams P;
    __shared__ Plan D;
    __shared__ Polygon G;
    __shared__ Tables TP;
    __shared__ Polygon quads[31];                // label 3: the polygons of the wide segments (at most 32 points)
    __shared__ LineBox LB;
    uint32_t* const rowpre = reinterpret_cast<uint32_t*>(smem);
    const int tid = threadIdx.x, b = blockIdx.x;
    const int H = g.H, W = g.W, npix = H * W;
    uint32_t* const cov = cov_all + (size_t)b * g.cov_stride;
    pdl_launch_dependents();
    ImgState* const S = state + b;

    if (tid < (int)(sizeof(SsadImageParams) / 4))
        reinterpret_cast<uint32_t*>(&P)[tid] = reinterpret_cast<const uint32_t*>(params + b)[tid];
    if (tid < 8) S->acc[tid] = 0;
    __syncthreads();

    const int cx0 = P.crop_off[0], cy0 = P.crop_off[1];
    const uint32_t* mrow0 = nullptr;
    if (d.mask_mode != SSAD_MASK_ALL)
        mrow0 = mask_bits + (d.mask_mode == SSAD_MASK_PER_IMAGE ? (size_t)(d.src_indexed ? (int)P.src_index : b) * d.mask_stride_words : 0);
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
    // label 3 "long line" (datasets.py:357-388): a flat colour behind the coverage of the poly-line, one "paste" of its
    // bounding box at the box's own position
    const int16_t* const lxy = (y == 3 && d.line_xy) ? d.line_xy + (size_t)b * d.line_max * 2 : nullptr;
    const int ln = lxy ? min((int)P.line_n, d.line_max) : 0;
    if (y == 3) {
        if (tid == 0) {
            LB = line_box(lxy, ln, d.line_width, W, H);
            if ((uint32_t)LB.w > g.cov_pitch * 32u || (uint32_t)LB.h > g.cov_rows) LB.w = LB.h = 0;    // (cannot happen: the host sizes the rows)
            D.paste[0][0] = LB.x0; D.paste[0][1] = LB.y0;
        }
        __syncthreads();
    }
    const int npaste = (y == 0) ? 0 : (y == 1 ? 1 : (y == 3 ? (LB.w > 0 ? 1 : 0) : P.scar_k));
    const int bw = (y == 1) ? pw : (y == 3 ? LB.w : P.scar_size[0]), bh = (y == 1) ? ph : (y == 3 ? LB.h : P.scar_size[1]);
    if (y != 3 && tid < npaste) {
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
    } else if (tid < 5 && !(y == 3 && tid == 0)) {
        D.paste[tid][0] = 0; D.paste[tid][1] = 0;
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
        D.fin.kind = PK_COPY; D.fin.sa = D.fin.sb = 0; D.fin.fs = 1.0f;
    }
    __syncthreads();
    build_prog(D.pre, TP, D.ops, D.fac, 0, max(D.cpos, 0), 0, tid);      // ops ahead of Contrast (never Contrast itself)
    const int pitchw = (int)g.cov_pitch;
    if (y == 1) {
        // rect2poly bit rows (dataset_generator.py:63-100)
        for (int v = tid; v < ph; v += NT) {
            uint32_t* row = cov + (size_t)v * pitchw;
            for (int j = 0; j < pitchw; ++j) row[j] = 0;
            polygon_fill_row(G, v, pw, row);
        }
    } else if (y == 2) {
        // alpha of scar.rotate(angle, expand=True) (datasets.py:344): per canvas row one span of u, in closed form
        for (int v = tid; v < bh; v += NT) {
            uint32_t* row = cov + (size_t)v * pitchw;
            for (int j = 0; j < pitchw; ++j) row[j] = 0;
            int lo = 0, hi = bw - 1;
            affine_span((long long)P.scar_fix[2] + (long long)v * P.scar_fix[1], P.scar_fix[0], (long long)pw << 16, lo, hi);
            affine_span((long long)P.scar_fix[5] + (long long)v * P.scar_fix[4], P.scar_fix[3], (long long)ph << 16, lo, hi);
            if (lo <= hi) row_set_bits(row, lo, hi);
        }
    } else if (y == 3 && npaste) {
        for (int i = tid; i < bh * pitchw; i += NT) cov[i] = 0;
        __syncthreads();
        line_rasterize(lxy, ln, d.line_width, W, H, LB, cov, pitchw, quads, tid, NT);
    }
    if (tid == 0) {
        // work lists: pass 1 runs the ops ahead of Contrast (and may have nothing to do at all), pass 2 the rest
        const bool need1 = P.warp_en != 0 || y != 0 || D.cpos >= 0;
        const int k1 = need1 ? kind_of_ops(D.ops, 0, max(D.cpos, 0)) : NKIND;
        const int k2 = kind_of_ops(D.ops, D.s_in_pre ? D.cpos : 0, D.nops);
        S->sk[0] = k1; S->sk[1] = (int32_t)atomicAdd(&sched->count[0][k1], 1u);
        S->sk[2] = k2; S->sk[3] = (int32_t)atomicAdd(&sched->count[1][k2], 1u);
    }
    __syncthreads();
    copy_words(&S->D, &D, tid);
    copy_words(&S->TP, &TP, tid);
}


// ================================================================ lists
// CTA-wide sums of NV counters with one REDUX per counter and warp: thread i < NV returns the total of counter i.
// `red` must not be in use by other warps of the CTA (it is written without a leading barrier).
template <int NV>
__device__ __forceinline__ uint32_t block_sum_to(uint32_t (&v)[NV], uint32_t* red /* NWARP * 8 words */) {
    #pragma unroll
    for (int i = 0; i < NV; ++i) v[i] = __reduce_add_sync(0xFFFFFFFFu, v[i]);
    if ((threadIdx.x & 31) == 0) {
        #pragma unroll
        for (int i = 0; i < NV; ++i) red[(threadIdx.x >> 5) * 8 + i] = v[i];
    }
    __syncthreads();
    uint32_t t = 0;
    if (threadIdx.x < NV) {
        #pragma unroll
        for (int w = 0; w < NWARP; ++w) t += red[w * 8 + threadIdx.x];
    }
    return t;
}

// slot of every image in the kind-ordered lists of the two passes (thread per image)
__global__ void __launch_bounds__(NT)
lists_v7_kernel(const __grid_constant__ Geo7 g, const SsadImageParams* __restrict__ params,
                const ImgState* __restrict__ state, const Sched* __restrict__ sched, Hdr* __restrict__ hdr) {
    const int b = blockIdx.x * NT + threadIdx.x;
    if (b >= g.batch) return;
    pdl_launch_dependents();
    pdl_wait();
    const ImgState* const S = state + b;
    const SsadImageParams* const p = params + b;
    const Plan& D = S->D;
    const int flags = (p->warp_en ? HF_WARP : 0) | (D.s_in_pre ? HF_PRE : 0) | (D.y != 0 ? HF_DEFECT : 0) | (D.cpos >= 0 ? HF_LUMA : 0);
    #pragma unroll
    for (int ps = 0; ps < 2; ++ps) {
        const int kind = S->sk[2 * ps], pos = S->sk[2 * ps + 1];
        // longest loop bodies first (the kinds are numbered by cost), so the short ones fill the tail of the grid;
        // images with nothing to do in this pass (kind == NKIND) sit behind them all
        uint32_t rank = (uint32_t)pos;
#ifdef SSAD_V7_LIGHT_FIRST
        for (int k = 0; k < kind; ++k) rank += sched->count[ps][k];
#else
        for (int k = 0; k < NKIND; ++k) if (kind == NKIND || k > kind) rank += sched->count[ps][k];
#endif
        Hdr h;
        h.b = (kind < NKIND) ? b : -1;
        h.kind = kind; h.flags = flags; h.cx0 = p->crop_off[0]; h.cy0 = p->crop_off[1];
        float fs = 1.0f;                 // factor of the Saturation inside the pass's program, if any
        const int i0 = (ps == 0) ? 0 : (D.s_in_pre ? D.cpos : 0), i1 = (ps == 0) ? max(D.cpos, 0) : D.nops;
        for (int i = i0; i < i1; ++i) if (D.ops[i] == 2) fs = D.fac[i];
        h.fs = fs; h.si = g.src_indexed ? (int32_t)p->src_index : b;
        // rows any paste touches: only those slabs of an un-warped defect image go through the x buffer
        int lo = g.H, hi = 0;
        for (int k = 0; k < D.npaste; ++k) { lo = min(lo, max(D.paste[k][1], 0)); hi = max(hi, min(D.paste[k][1] + D.bh, g.H)); }
        h.rows = (lo < hi) ? ((uint32_t)lo | ((uint32_t)hi << 16)) : 0u;
        hdr[(size_t)ps * g.batch + rank] = h;
    }
}

// ================================================================ pass 1
__global__ void __launch_bounds__(NT, SSAD_V7_P1_MINB)
pass1_v7_kernel(const __grid_constant__ SsadBatchDesc d, const __grid_constant__ Geo7 g,
                const SsadImageParams* __restrict__ params, const uint8_t* __restrict__ src,
                const uint8_t* __restrict__ donor_bank, ImgState* __restrict__ state, const Hdr* __restrict__ hdr,
                uint8_t* xbuf) {
    __shared__ Tables TP;
    __shared__ uint32_t red[NWARP * 8 + 8];
    const int tid = threadIdx.x;
    const uint32_t j = g.nslab_shift >= 0 ? blockIdx.x >> g.nslab_shift : blockIdx.x / (unsigned)g.nslab;
    const int slab = (int)(blockIdx.x - j * (unsigned)g.nslab);
    pdl_launch_dependents();
    pdl_wait();
    const int4 h0 = __ldg(reinterpret_cast<const int4*>(hdr + j));                  // b, kind, flags, cx0
    const int b = h0.x, kind = h0.y, flags = h0.z;
    if (b < 0) return;                           // images with nothing to do in pass 1 sit behind the lists
    const int4 h1 = __ldg(reinterpret_cast<const int4*>(hdr + j) + 1);              // cy0, fs, source image
    const int H = g.H, W = g.W, npix = H * W;
    ImgState* const S = state + b;
    const bool warped = (flags & HF_WARP) != 0, defect = (flags & HF_DEFECT) != 0, need_luma = (flags & HF_LUMA) != 0;
    const bool store_pre = (flags & HF_PRE) != 0;                    // pass 2 starts from pre(x)
    const int r0 = slab * g.rows_slab, r1 = min(H, r0 + g.rows_slab);
    const bool pasted = r0 < (int)((uint32_t)h1.w >> 16) && r1 > (int)((uint32_t)h1.w & 0xFFFFu);     // a paste touches this slab
    const bool store_raw = defect && !store_pre && !warped && pasted;      // the pastes are written into x
    const int nunits = (r1 - r0) * g.upr;
    const uint8_t* const sbase = src + (size_t)h1.z * d.src_stride + (size_t)h1.x * d.src_pitch + (size_t)h0.w * 3;
    const int spitch = (int)d.src_pitch;
    uint8_t* const xb = xbuf + (size_t)b * npix * 3;
    const int xpitch = W * 3;
    // x as the passes read it (each thread re-reads only what it wrote itself)
    const uint8_t* const xin_base = warped ? xb : sbase;
    const int xin_pitch = warped ? xpitch : spitch;

    // the first loads of the stream leave before the tables arrive
    FastCol F;
    Raw nx;
    const bool fast = g.fast != 0;
    if (fast && !warped) {
        F = fast_col(xin_base + (size_t)r0 * xin_pitch, xin_pitch, g, tid);
        ld_raw(F.ip, F.mis != 0, false, nx);
    }
    const SsadImageParams& P = params[b];        // read in place (L1 broadcast): only the gather and the cut sums need it
    if (kind != LK_COPY) {                       // (uniform over the CTA)
        copy_words(&TP, &S->TP, tid);
        __syncthreads();
    }

    uint32_t acc[7] = {0, 0, 0, 0, 0, 0, 0};          // 0-2 RGB of x, 3 sum of L, 4-6 RGB of the cut
    if (warped) {
        const Slab SL = make_slab(r0, nunits, g.upr, tid);
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
            store24(xb + (size_t)r0 * xpitch + (size_t)lin * 24, w);
        }
        if (!defect && !need_luma) return;       // only the gather was needed
        if (fast) {
            F = fast_col(xin_base + (size_t)r0 * xin_pitch, xin_pitch, g, tid);
            ld_raw(F.ip, F.mis != 0, false, nx);
        }
    }
    {
        ProgRegs R;
        R.fs = __int_as_float(h1.y); R.ta = smem_u32(TP.a); R.tab = smem_u32(TP.ab); R.tbc = smem_u32(TP.b) + (uint32_t)TB_OFF - kMagicBits;
        const bool need_rgb = defect;
        if (fast) {
            uint8_t* const xb0 = xb + (size_t)r0 * xpitch;
            #define SSAD_CALL(LK) pass1_fast<LK>(F, nx, xb0, xpitch, store_raw, store_pre, need_rgb, need_luma, R, acc)
            SSAD_V7_BY_KIND(kind, SSAD_CALL)
            #undef SSAD_CALL
        } else {
            const Slab SL = make_slab(r0, nunits, g.upr, tid);
            #define SSAD_CALL(LK) pass1_gen<LK>(SL, xin_base, xin_pitch, xb, xpitch, store_raw, store_pre, need_rgb, need_luma, R, acc)
            SSAD_V7_BY_KIND(kind, SSAD_CALL)
            #undef SSAD_CALL
        }
        // RGB sums of the would-be cut (dataset_generator.py:202-207 / 147-159): rows dealt over the image's warps
        if (defect && P.src_variant != 2) {
            // the cut box dealt pixel by pixel over the threads of the image's slabs, four loads in flight per channel
            const Donor dn = donor_of(d, src, donor_bank, h1.z, P);
            const int bw = max(dn.u_hi, 0), total = bw * max(dn.v_hi, 0);
            const uint32_t inv_bw = 0xFFFFFFFFu / (uint32_t)max(bw, 2) + 1u;
            const int stride = g.nslab * NT;
            for (int i0 = slab * NT + tid; i0 < total; i0 += 4 * stride) {
                uint32_t r[4], gg[4], bl[4];
                #pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const int idx = i0 + q * stride;
                    r[q] = gg[q] = bl[q] = 0;
                    if (idx < total) {
                        const int v = (bw < 2) ? idx : (int)__umulhi((uint32_t)idx, inv_bw), u = idx - v * bw;
                        const uint8_t* p = dn.base + (size_t)v * dn.pitch + 3 * u;
                        r[q] = __ldg(p); gg[q] = __ldg(p + 1); bl[q] = __ldg(p + 2);
                    }
                }
                acc[4] += (r[0] + r[1]) + (r[2] + r[3]); acc[5] += (gg[0] + gg[1]) + (gg[2] + gg[3]); acc[6] += (bl[0] + bl[1]) + (bl[2] + bl[3]);
            }
        }
    }
    const uint32_t tot = block_sum_to<7>(acc, red);
    if (tid < 7 && tot != 0) atomicAdd(&S->acc[tid], tot);
}

// ---- 8-pixel units in the canvas kernel (polygon / poly-line paste)
// Convert.c L of the 8 pixels of a unit
__device__ __forceinline__ void luma8(const uint32_t (&w)[6], uint32_t (&L)[8]) {
    constexpr uint32_t kRG = (38470u << 16) | 19595u, kB0 = 7471u;          // pixel at bytes 0,1,2 of a word
    constexpr uint32_t k0R = (19595u << 16), kGB = (7471u << 16) | 38470u;  // pixel at bytes 1,2,3 of a word
    #pragma unroll
    for (int q = 0; q < 2; ++q) {
        const uint32_t w0 = w[3 * q], w1 = w[3 * q + 1], w2 = w[3 * q + 2];
        const uint32_t p1 = __byte_perm(w0, w1, 0x0543), p2 = __byte_perm(w1, w2, 0x0432);
        L[4 * q + 0] = dp2a_hi(kB0, w0, dp2a_lo(kRG, w0, 0x8000u)) >> 16;
        L[4 * q + 1] = dp2a_hi(kB0, p1, dp2a_lo(kRG, p1, 0x8000u)) >> 16;
        L[4 * q + 2] = dp2a_hi(kB0, p2, dp2a_lo(kRG, p2, 0x8000u)) >> 16;
        L[4 * q + 3] = dp2a_hi(kGB, w2, dp2a_lo(k0R, w2, 0x8000u)) >> 16;
    }
}
// 8 coverage bits of canvas row `crow` for canvas x = u0 .. u0 + 7 (u0 may be negative; bits beyond the row are zero)
__device__ __forceinline__ uint32_t cov_bits8(const uint32_t* __restrict__ crow, int u0, int pitch_words) {
    const int lo = max(u0, 0);
    const int wi = lo >> 5;
    const uint32_t a = wi < pitch_words ? crow[wi] : 0u;
    const uint32_t b = (wi + 1 < pitch_words) ? crow[wi + 1] : 0u;
    const uint32_t bits = __funnelshift_r(a, b, lo & 31) & 0xFFu;      // canvas x = lo .. lo + 7
    return (u0 < 0) ? ((u0 > -8) ? (bits << (-u0)) & 0xFFu : 0u) : bits;   // pixel p of the unit <-> canvas x = u0 + p
}

// One paste of a polygon patch (or the flat colour of a poly-line) into the x buffer, an 8-pixel unit at a time:
// coverage bits -> byte masks, the cut / flat colour as 24 bytes, the old pixels as 24 bytes, a byte-wise select, and
// the change of the sum of L that Contrast will see (through the `pre` program of the image, kind LK).
template <int LK>
__device__ __forceinline__ uint32_t polygon_units(int total, const uint16_t* row_cu0, const uint32_t* row_pre, int nrows, int my_lo,
                                                  const uint32_t* __restrict__ cov, int pitchw, int pl, int pt, int W,
                                                  uint8_t* xb, const Donor& dn, bool flat, const uint32_t (&fw)[3], bool cut_sim,
                                                  uint32_t simtab_addr, const uint32_t (*bmask)[3], bool want_l, const ProgRegs& R, int tid) {
    constexpr bool SP = LK >= LK_SAT_UC;          // Saturation precedes Contrast: the x buffer holds pre(x)
    uint32_t dl = 0;
    for (int t = tid; t < total; t += NT) {
        int lo = 0, hi = max(nrows - 1, 0);       // the row of item t
        while (lo < hi) {
            const int mid = (lo + hi + 1) >> 1;
            if ((int)row_pre[mid] <= t) lo = mid; else hi = mid - 1;
        }
        const int v = my_lo + lo, cu = (int)row_cu0[lo] + t - (int)row_pre[lo];
        const int X0 = cu * 8, u0 = X0 - pl;
        const uint32_t m8 = cov_bits8(cov + (size_t)v * pitchw, u0, pitchw);
        if (!m8) continue;
        uint8_t* const q = xb + ((size_t)(pt + v) * W + X0) * 3;
        uint32_t O[6], N[6];
        {
            const uint2* qo = reinterpret_cast<const uint2*>(q);
            const uint2 a = qo[0], b = qo[1], c = qo[2];
            O[0] = a.x; O[1] = a.y; O[2] = b.x; O[3] = b.y; O[4] = c.x; O[5] = c.y;
        }
        if (flat) {
            N[0] = fw[0]; N[1] = fw[1]; N[2] = fw[2]; N[3] = fw[0]; N[4] = fw[1]; N[5] = fw[2];
        } else {
            if (u0 >= 0 && u0 + 8 <= dn.u_hi && v + 1 < dn.v_hi) {  // the whole unit lies inside the cut (and a row follows:
                                                                    // the aligned 32-byte read may run 8 bytes past the unit)
                const uint8_t* c = dn.base + (size_t)v * dn.pitch + 3 * u0;
                const uint32_t mis = (uint32_t)(reinterpret_cast<uintptr_t>(c) & 7u);
                Raw r;
                ld_raw(c - mis, mis != 0, true, r);
                raw_to_words(r, mis, N);
            } else {                                               // patch edge: PIL's crop pads with 0 outside the donor
                #pragma unroll
                for (int i = 0; i < 6; ++i) N[i] = 0u;
                for (int p = 0; p < 8; ++p) {
                    const int u = u0 + p;
                    if (!((m8 >> p) & 1u) || u < 0 || u >= dn.u_hi || v >= dn.v_hi) continue;
                    const uint8_t* c = dn.base + (size_t)v * dn.pitch + 3 * u;
                    #pragma unroll
                    for (int ch = 0; ch < 3; ++ch) {
                        const int k = 3 * p + ch;
                        N[k >> 2] |= (uint32_t)__ldg(c + ch) << (8 * (k & 3));
                    }
                }
            }
            if (cut_sim) unit_bytelut(N, simtab_addr);
        }
        uint32_t M[6];
        {
            const uint32_t n0 = m8 & 15u, n1 = m8 >> 4;
            M[0] = bmask[n0][0]; M[1] = bmask[n0][1]; M[2] = bmask[n0][2];
            M[3] = bmask[n1][0]; M[4] = bmask[n1][1]; M[5] = bmask[n1][2];
        }
        if (want_l) {
            uint32_t Np[6], Op[6], Ln[8], Lo[8];
            #pragma unroll
            for (int i = 0; i < 6; ++i) { Np[i] = N[i]; Op[i] = O[i]; }
            apply_prog<LK>(Np, R);                               // pre(new)
            if (!SP) apply_prog<LK>(Op, R);                      // the x buffer holds x itself: pre(old)
            luma8(Np, Ln);
            luma8(Op, Lo);
            #pragma unroll
            for (int p = 0; p < 8; ++p) if ((m8 >> p) & 1u) dl += Ln[p] - Lo[p];
            if (SP) {
                #pragma unroll
                for (int i = 0; i < 6; ++i) N[i] = Np[i];
            }
        }
        uint2* qs = reinterpret_cast<uint2*>(q);
        qs[0] = make_uint2((O[0] & ~M[0]) | (N[0] & M[0]), (O[1] & ~M[1]) | (N[1] & M[1]));
        qs[1] = make_uint2((O[2] & ~M[2]) | (N[2] & M[2]), (O[3] & ~M[3]) | (N[3] & M[3]));
        qs[2] = make_uint2((O[4] & ~M[4]) | (N[4] & M[4]), (O[5] & ~M[5]) | (N[5] & M[5]));
    }
    return dl;
}

// ================================================================ canvas
__global__ void __launch_bounds__(NT)
canvas_v7_kernel(const __grid_constant__ SsadBatchDesc d, const __grid_constant__ Geo7 g,
                 const SsadImageParams* __restrict__ params, const uint8_t* __restrict__ src,
                 const uint8_t* __restrict__ donor_bank, ImgState* __restrict__ state,
                 const uint32_t* __restrict__ cov_all, uint32_t* __restrict__ canvas_all, uint8_t* xbuf) {
    __shared__ SsadImageParams P;
    __shared__ Plan D;
    __shared__ Tables TP;
    __shared__ uint8_t simtab[256];
    __shared__ uint32_t red[NWARP * 8 + 8];
    __shared__ uint16_t row_lo[kCanvasRows];     // first covered column of this CTA's canvas rows (polygon: first unit column)
    __shared__ uint32_t row_pre[kCanvasRows + 1];   // covered-span pixels (polygon: 8-pixel units) in the rows before
    __shared__ uint32_t bmask[16][3];            // byte masks of the four pixels of half a unit, by their coverage bits
    const int tid = threadIdx.x;
    const int b = (int)(blockIdx.x / (unsigned)g.ncanvas), part = (int)(blockIdx.x - (unsigned)b * g.ncanvas);
    const int H = g.H, W = g.W, npix = H * W;
    ImgState* const S = state + b;
    pdl_launch_dependents();
    pdl_wait();
    const int y0 = S->D.y;                       // (uniform over the CTA)
    if (y0 == 0) return;
    const int nparts = g.ncanvas;                // CTAs that share the image's canvas
    if (part >= nparts) return;
    if (tid < (int)(sizeof(SsadImageParams) / 4))
        reinterpret_cast<uint32_t*>(&P)[tid] = reinterpret_cast<const uint32_t*>(params + b)[tid];
    copy_words(&D, &S->D, tid);
    copy_words(&TP, &S->TP, tid);
    __syncthreads();
    const int y = D.y, npaste = D.npaste, bw = D.bw, bh = D.bh, cpos = D.cpos;
    const bool s_in_pre = D.s_in_pre != 0;
    const int pw = P.rect[2], ph = P.rect[3];
    const uint32_t* const cov = cov_all + (size_t)b * g.cov_stride;
    uint32_t* const canvas = canvas_all + (size_t)b * g.canvas_stride;
    uint8_t* const xb = xbuf + (size_t)b * npix * 3;
    const int pitchw = (int)g.cov_pitch;
    const Donor dn = donor_of(d, src, donor_bank, g.src_indexed ? (int)P.src_index : b, P);

    // ---- check_color_similarity + variant colours (every CTA of the image decides the same; part 0 records it)
    if (tid == 0) {
        const uint32_t np = (uint32_t)(pw * ph);
        uint32_t xs[3] = {S->acc[0], S->acc[1], S->acc[2]};
        uint32_t psum[3] = {S->acc[4], S->acc[5], S->acc[6]};
        int flat = 0;
        uint32_t frgb[3] = {0, 0, 0};
        int avg[3] = {0, 0, 0};
        if (y0 == 3) {                      // the line's colour, no similarity branch (datasets.py:369-388)
            flat = 1;
            for (int k = 0; k < 3; ++k) { frgb[k] = P.rnd_rgb[k]; psum[k] = 0; }
        } else if (P.src_variant == 1) {    // colour = int(mean) (dataset_generator.py:202-207)
            flat = 1;
            for (int k = 0; k < 3; ++k) { frgb[k] = psum[k] / np; avg[k] = (int)frgb[k]; psum[k] = frgb[k] * np; }
        } else if (P.src_variant == 2) {    // random colour (dataset_generator.py:194-199)
            flat = 1;
            for (int k = 0; k < 3; ++k) { frgb[k] = P.rnd_rgb[k]; psum[k] = frgb[k] * np; }
        }
        double cs = 0.0;
        const bool sim = (y0 != 3) && similarity_branch(xs, (double)npix, psum, (double)np, cs);     // datasets.py:295 / 335
        if (flat && sim)                    // brightness of a flat patch is a flat patch
            for (int k = 0; k < 3; ++k) frgb[k] = brightness_u8(brightness_u8(frgb[k], D.bf1), D.bf2);
        D.flat = flat; D.sim = sim ? 1 : 0; D.cosine = cs;
        for (int k = 0; k < 3; ++k) { D.flat_rgb[k] = frgb[k]; D.avg_rgb[k] = avg[k]; }
        if (part == 0) {
            S->D.flat = flat; S->D.sim = sim ? 1 : 0; S->D.cosine = cs;
            for (int k = 0; k < 3; ++k) { S->D.flat_rgb[k] = frgb[k]; S->D.avg_rgb[k] = avg[k]; }
        }
    }
    __syncthreads();
    Colour K;
    K.flat = D.flat; K.sim = D.sim; K.bf1 = D.bf1; K.bf2 = D.bf2;
    K.frgb[0] = D.flat_rgb[0]; K.frgb[1] = D.flat_rgb[1]; K.frgb[2] = D.flat_rgb[2];

    // ---- patch canvas: canvas pixel (u, v) -> packed (r, g, b, L): the colour that is pasted (after the `pre` program
    // when the x buffer holds pre(x)) and the L Contrast will see there.  This CTA owns a share of the canvas rows that
    // land inside the image; the (row, column) box is dealt pixel by pixel over the threads.
    const bool cut_sim = !K.flat && K.sim;
    if (cut_sim) {
        for (int v = tid; v < 256; v += NT) simtab[v] = (uint8_t)brightness_u8(brightness_u8((uint32_t)v, K.bf1), K.bf2);
        __syncthreads();
    }
    int cv_lo = bh, cv_hi = 0;
    for (int k = 0; k < npaste; ++k) {
        const int pt = D.paste[k][1];
        const int v_lo = max(0, -pt), v_hi = min(bh, H - pt);
        if (v_lo < v_hi) { cv_lo = min(cv_lo, v_lo); cv_hi = max(cv_hi, v_hi); }
    }
    const int per = (max(cv_hi - cv_lo, 0) + nparts - 1) / nparts;
    const int my_lo = cv_lo + part * per, my_hi = min(cv_hi, my_lo + per);
    if (my_lo >= my_hi) return;
    const int cpitch = (int)g.canvas_pitch;
    const bool want_l = cpos >= 0;
    uint32_t dl = 0;                             // sum of L(new) - L(old) over the pixels this CTA pastes
    // Only the covered span [first set bit, last set bit] of every canvas row is visited (a rotated scar covers a third
    // of its box): spans -> prefix over the rows -> item idx of the flattened spans -> (row, column) by bisection.
    const int nrows = my_hi - my_lo;
    for (int r = tid; r < nrows; r += NT) {
        const uint32_t* crow = cov + (size_t)(my_lo + r) * pitchw;
        int lo = -1, hi = -1;
        for (int j = 0; j < pitchw; ++j) {
            const uint32_t wv = crow[j];
            if (wv) { if (lo < 0) lo = 32 * j + __ffs(wv) - 1; hi = 32 * j + 31 - __clz(wv); }
        }
        lo = min(max(lo, 0), max(bw - 1, 0)); hi = min(hi, bw - 1);
        if (y != 2) {
            // polygon / poly-line: the 8-pixel units of the image row the covered span touches
            const int Xlo = max(D.paste[0][0] + lo, 0), Xhi = min(D.paste[0][0] + hi, W - 1);
            const bool any = hi >= lo && hi >= 0 && Xlo <= Xhi;
            row_lo[r] = (uint16_t)(Xlo >> 3);
            row_pre[r + 1] = any ? (uint32_t)((Xhi >> 3) - (Xlo >> 3) + 1) : 0u;
        } else {
            row_lo[r] = (uint16_t)lo;
            row_pre[r + 1] = (hi >= lo && hi >= 0) ? (uint32_t)(hi - lo + 1) : 0u;
        }
    }
    if (tid < 48) {
        const int n = tid / 3, j = tid - 3 * n;
        uint32_t m = 0;
        for (int qb = 0; qb < 4; ++qb) if ((n >> ((4 * j + qb) / 3)) & 1) m |= 0xFFu << (8 * qb);
        bmask[n][j] = m;
    }
    __syncthreads();
    if (tid < 32) {
        uint32_t carry = 0;
        for (int base = 0; base < nrows; base += 32) {
            const int r = base + tid;
            uint32_t sv = (r < nrows) ? row_pre[r + 1] : 0u;
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, sv, o);
                if (tid >= o) sv += t;
            }
            if (r < nrows) row_pre[r + 1] = carry + sv;
            carry += __shfl_sync(0xFFFFFFFFu, sv, 31);
        }
        if (tid == 0) row_pre[0] = 0;
    }
    __syncthreads();
    const int span_total = (int)row_pre[nrows];
    // Items are dealt to the warps in contiguous ranges (lanes = consecutive items), so a warp's items climb through
    // the rows a span piece at a time: `seek` bisects to the row of a range's first item, `locate` walks on from the
    // row of the previous item (items must not decrease between two seeks).
    const int warp = tid >> 5, lane = tid & 31;
    int row_at = 0;
    auto seek = [&](int idx) {
        int lo = 0, hi = max(nrows - 1, 0);
        while (lo < hi) {
            const int mid = (lo + hi + 1) >> 1;
            if ((int)row_pre[mid] <= idx) lo = mid; else hi = mid - 1;
        }
        row_at = lo;
    };
    auto locate = [&](int idx, int& u, int& v) {
        while ((int)row_pre[row_at + 1] <= idx) ++row_at;
        v = my_lo + row_at; u = (int)row_lo[row_at] + idx - (int)row_pre[row_at];
    };
    auto warp_range = [&](int total, int& w_lo, int& w_hi) {
        const int per_w = ((total + NWARP - 1) / NWARP + 31) & ~31;
        w_lo = min(total, warp * per_w); w_hi = min(total, w_lo + per_w);
    };
    if (y == 1 || y == 3) {
        // polygon (or the poly-line of label 3: a flat colour): one paste, straight into the x buffer (paste_patch,
        // dataset_generator.py:268-275), 8 pixels = 24 bytes at a time (see polygon_units)
        uint32_t fw[3];                          // the flat colour as three words of repeating R, G, B bytes
        {
            const uint32_t r = K.frgb[0], gg = K.frgb[1], bl = K.frgb[2];
            fw[0] = r | (gg << 8) | (bl << 16) | (r << 24);
            fw[1] = gg | (bl << 8) | (r << 16) | (gg << 24);
            fw[2] = bl | (r << 8) | (gg << 16) | (bl << 24);
        }
        ProgRegs R;
        R.fs = D.pre.fs; R.ta = smem_u32(TP.a); R.tab = smem_u32(TP.ab); R.tbc = smem_u32(TP.b) + (uint32_t)TB_OFF - kMagicBits;
        const int kind = want_l ? loop_kind(D.pre) : LK_COPY;
        #define SSAD_CALL(LK) dl = polygon_units<LK>(span_total, row_lo, row_pre, nrows, my_lo, cov, pitchw, D.paste[0][0], D.paste[0][1], \
                                                     W, xb, dn, K.flat != 0, fw, cut_sim, smem_u32(simtab), bmask, want_l, R, tid)
        SSAD_V7_BY_KIND(kind, SSAD_CALL)
        #undef SSAD_CALL
    } else {
        int w_lo, w_hi;
        warp_range(span_total, w_lo, w_hi);
        if (w_lo < w_hi) seek(w_lo);
        for (int idx = w_lo + lane; idx < w_hi; idx += 32) {
            int u, v;
            locate(idx, u, v);
            if (!((cov[(size_t)v * pitchw + (u >> 5)] >> (u & 31)) & 1u)) continue;
            uint32_t r, gg, bl;
            if (K.flat) { r = K.frgb[0]; gg = K.frgb[1]; bl = K.frgb[2]; }
            else {
                int su = u, sv = v;          // source pixel of the cut: itself (polygon) | inverse rotation (scar)
                if (y == 2) {
                    su = (P.scar_fix[2] + v * P.scar_fix[1] + u * P.scar_fix[0]) >> 16;
                    sv = (P.scar_fix[5] + v * P.scar_fix[4] + u * P.scar_fix[3]) >> 16;
                }
                r = gg = bl = 0;
                if (su < dn.u_hi && sv < dn.v_hi) {
                    const uint8_t* q = dn.base + (size_t)sv * dn.pitch + 3 * su;
                    r = __ldg(q); gg = __ldg(q + 1); bl = __ldg(q + 2);
                }
                if (cut_sim) { r = simtab[r]; gg = simtab[gg]; bl = simtab[bl]; }
            }
            uint32_t L = 0;
            if (want_l) {
                uint32_t pr = r, pg = gg, pb = bl;
                pixel_prog(D.pre, TP, pr, pg, pb);
                L = luma_u8(pr, pg, pb);
                if (s_in_pre) { r = pr; gg = pg; bl = pb; }
            }
            if (y == 1) {
                // one paste: straight into the x buffer (paste_patch, dataset_generator.py:268-275)
                const int X = D.paste[0][0] + u, Y = D.paste[0][1] + v;
                if (X < 0 || X >= W) continue;
                uint8_t* const q = xb + ((size_t)Y * W + X) * 3;
                if (want_l) {
                    uint32_t pr = q[0], pg = q[1], pb = q[2];
                    if (!s_in_pre) pixel_prog(D.pre, TP, pr, pg, pb);      // the x buffer holds x itself
                    dl += L - luma_u8(pr, pg, pb);
                }
                q[0] = (uint8_t)r; q[1] = (uint8_t)gg; q[2] = (uint8_t)bl;
                continue;
            }
            canvas[(size_t)v * cpitch + u] = r | (gg << 8) | (bl << 16) | (L << 24);
        }
    }
    if (y == 2) __syncthreads();
    // ---- paste_patch / scar pastes (dataset_generator.py:268-275, datasets.py:340-355) into the x buffer: every image
    // pixel is written by the LAST paste that covers it, so no two threads meet; the sum of L Contrast sees moves by
    // L(new) - L(old)
    // (all pastes x canvas rows of this CTA x columns) dealt over the threads, NB items per thread and round with
    // their loads issued together
    if (y == 2) {
        constexpr int NB = SSAD_V7_NB_SCAR;
        const int box = max(span_total, 1), total = npaste * span_total;
        int w_lo, w_hi;
        warp_range(total, w_lo, w_hi);
        int k = 0, kbase = 0;                    // paste of the current item and its first id (ids only grow)
        if (w_lo < w_hi) { k = (w_lo + lane) / box; kbase = k * box; seek(min(w_lo + lane, w_hi - 1) - kbase); }
        for (int id0 = w_lo + lane; id0 < w_hi; id0 += 32 * NB) {
            bool on[NB];
            uint8_t* qx[NB];
            const uint32_t* qc[NB];
            #pragma unroll
            for (int j = 0; j < NB; ++j) {
                const int id = min(id0 + j * 32, w_hi - 1);
                while (id >= kbase + box) { kbase += box; ++k; row_at = 0; }     // next paste: back to the first row
                const int idx = id - kbase;
                int u, v;
                locate(idx, u, v);
                const int X = D.paste[k][0] + u, Y = D.paste[k][1] + v;
                bool ok = (id0 + j * 32 < w_hi) && X >= 0 && X < W && Y >= 0 && Y < H &&
                          ((cov[(size_t)v * pitchw + (u >> 5)] >> (u & 31)) & 1u);
                if (ok)                          // a later paste overwrites this pixel: it is written there
                    for (int jj = k + 1; jj < npaste; ++jj) {
                        const int uj = X - D.paste[jj][0], vj = Y - D.paste[jj][1];
                        if (uj >= 0 && uj < bw && vj >= 0 && vj < bh && ((cov[(size_t)vj * pitchw + (uj >> 5)] >> (uj & 31)) & 1u)) ok = false;
                    }
                on[j] = ok;
                qc[j] = ok ? canvas + (size_t)v * cpitch + u : nullptr;
                qx[j] = ok ? xb + ((size_t)Y * W + X) * 3 : nullptr;
            }
            uint32_t c[NB], xr[NB], xg[NB], xbl[NB];
            #pragma unroll
            for (int j = 0; j < NB; ++j) {
                c[j] = xr[j] = xg[j] = xbl[j] = 0;
                if (on[j]) {
                    c[j] = *qc[j];
                    if (want_l) { xr[j] = qx[j][0]; xg[j] = qx[j][1]; xbl[j] = qx[j][2]; }
                }
            }
            #pragma unroll
            for (int j = 0; j < NB; ++j) {
                if (!on[j]) continue;
                if (want_l) {
                    uint32_t pr = xr[j], pg = xg[j], pb = xbl[j];
                    if (!s_in_pre) pixel_prog(D.pre, TP, pr, pg, pb);      // the x buffer holds x itself
                    dl += (c[j] >> 24) - luma_u8(pr, pg, pb);
                }
                qx[j][0] = (uint8_t)c[j]; qx[j][1] = (uint8_t)(c[j] >> 8); qx[j][2] = (uint8_t)(c[j] >> 16);
            }
        }
    }
    if (!want_l) return;
    uint32_t v1[1] = {dl};
    const uint32_t tot = block_sum_to<1>(v1, red);
    if (tid == 0 && tot != 0) atomicAdd(&S->acc[7], tot);
}

// ================================================================ finish
__global__ void __launch_bounds__(NT)
finish_v7_kernel(const __grid_constant__ SsadBatchDesc d, const __grid_constant__ Geo7 g, ImgState* __restrict__ state,
                 int64_t* __restrict__ y_out, SsadDebug* __restrict__ debug) {
    __shared__ Plan D;
    __shared__ Tables TF;
    const int tid = threadIdx.x, b = blockIdx.x;
    const int npix = g.H * g.W;
    ImgState* const S = state + b;
    pdl_launch_dependents();
    pdl_wait();
    copy_words(&D, &S->D, tid);
    __syncthreads();
    const int cpos = D.cpos, nops = D.nops;
    int cmean = -1;
    if (cpos >= 0) {
        // int(mean + 0.5) with mean = sum / count in float64 (ImageEnhance.py Contrast) == integer rounding
        const uint32_t lsum = S->acc[3] + S->acc[7];
        cmean = (int)((2ull * lsum + (uint32_t)npix) / (2ull * (uint32_t)npix));
    }
    // what pass 2 still has to do: Contrast and the ops after it on pre(x), or the whole chain on x
    build_prog(D.fin, TF, D.ops, D.fac, D.s_in_pre ? cpos : 0, nops, cmean, tid);
    if (tid == 0) D.cmean = cmean;
    __syncthreads();
    copy_words(&S->D, &D, tid);
    copy_words(&S->TF, &TF, tid);
    if (tid == 0) {
        const int y = D.y;
        if (y_out) y_out[b] = y;
        if (debug) {
            SsadDebug o;
            memset(&o, 0, sizeof(o));
            o.n_coords = D.n_coords;
            o.label_out = y;
            o.contrast_mean = cmean;
            if (y != 0) {
                o.sim = D.sim;
                o.cosine = D.cosine;
                for (int k = 0; k < D.npaste; ++k) { o.paste[k][0] = D.paste[k][0]; o.paste[k][1] = D.paste[k][1]; }
                for (int k = 0; k < 3; ++k) o.avg_rgb[k] = D.avg_rgb[k];
            }
            debug[b] = o;
        }
    }
}

// ================================================================ pass 2
__global__ void __launch_bounds__(NT, SSAD_V7_P2_MINB)
pass2_v7_kernel(const __grid_constant__ SsadBatchDesc d, const __grid_constant__ Geo7 g, const uint8_t* __restrict__ src,
                const ImgState* __restrict__ state, const Hdr* __restrict__ hdr, const uint8_t* xbuf,
                const void* __restrict__ norm_lut, uint8_t* __restrict__ out_u8, void* __restrict__ out_norm) {
    __shared__ Tables TF;
    __shared__ __align__(16) uint8_t lut_s[3072];
    const int tid = threadIdx.x;
    const uint32_t j = g.nslab_shift >= 0 ? blockIdx.x >> g.nslab_shift : blockIdx.x / (unsigned)g.nslab;
    const int slab = (int)(blockIdx.x - j * (unsigned)g.nslab);
    pdl_launch_dependents();
    pdl_wait();
    const int4 h0 = __ldg(reinterpret_cast<const int4*>(hdr + j));                  // b, kind, flags, cx0
    const int4 h1 = __ldg(reinterpret_cast<const int4*>(hdr + j) + 1);              // cy0, fs, source image
    const int b = h0.x, kind = h0.y;
    if (b < 0) return;
    const int H = g.H, W = g.W, npix = H * W;
    const int r0 = slab * g.rows_slab, r1 = min(H, r0 + g.rows_slab);
    const bool pasted = r0 < (int)((uint32_t)h1.w >> 16) && r1 > (int)((uint32_t)h1.w & 0xFFFFu);
    // pass 2 starts from the x buffer where pass 1 left something there: warped x, pre(x), or a slab with pastes
    const bool stored = (h0.z & (HF_WARP | HF_PRE)) != 0 || ((h0.z & HF_DEFECT) != 0 && pasted);
    const int nunits = (r1 - r0) * g.upr;
    const uint8_t* const in2 = stored ? xbuf + (size_t)b * npix * 3
                                      : src + (size_t)h1.z * d.src_stride + (size_t)h1.x * d.src_pitch + (size_t)h0.w * 3;
    const int in2_pitch = stored ? W * 3 : (int)d.src_pitch;
    // the first loads of the stream leave before the tables arrive
    FastCol F;
    Raw nx;
    if (g.fast) {
        F = fast_col(in2 + (size_t)r0 * in2_pitch, in2_pitch, g, tid);
        ld_raw(F.ip, F.mis != 0, false, nx);
    }
    if (kind != LK_COPY) copy_words(&TF, &state[b].TF, tid);
    if (out_norm != nullptr) {
        const int nw = (d.out_norm_dtype == SSAD_NORM_F32) ? 768 : 384;
        for (int i = tid; i < nw; i += NT)
            reinterpret_cast<uint32_t*>(lut_s)[i] = reinterpret_cast<const uint32_t*>(norm_lut)[i];
    }
    __syncthreads();
    ProgRegs R;
    R.fs = __int_as_float(h1.y); R.ta = smem_u32(TF.a); R.tab = smem_u32(TF.ab); R.tbc = smem_u32(TF.b) + (uint32_t)TB_OFF - kMagicBits;
    uint8_t* const obase = out_u8 ? out_u8 + (size_t)b * d.out_u8_stride : nullptr;
    void* onorm = nullptr;
    if (out_norm) onorm = (d.out_norm_dtype == SSAD_NORM_F32) ? (void*)(reinterpret_cast<float*>(out_norm) + (size_t)b * 3 * npix)
                                                               : (void*)(reinterpret_cast<uint16_t*>(out_norm) + (size_t)b * 3 * npix);
    if (g.fast) {
        #define SSAD_CALL(LK) pass2_fast<LK>(F, nx, r0, W, npix, R, obase, onorm, d.out_norm_dtype, lut_s)
        SSAD_V7_BY_KIND(kind, SSAD_CALL)
        #undef SSAD_CALL
    } else {
        const Slab SL = make_slab(r0, nunits, g.upr, tid);
        #define SSAD_CALL(LK) pass2_gen<LK>(SL, in2, in2_pitch, R, W, npix, obase, onorm, d.out_norm_dtype, lut_s)
        SSAD_V7_BY_KIND(kind, SSAD_CALL)
        #undef SSAD_CALL
    }
}

// Tuning knobs from the environment (experiments, tests), read once at the first call (ssad_reload_env re-reads them)
struct PipeEnv {
    int upt, nofast, nc, lanes, chunk_mb, split, nopdl;
};
static PipeEnv g_env;
static std::once_flag g_env_once;
static void load_env() {
    auto geti = [](const char* name, int lo, int hi) { const char* f = getenv(name); if (!f) return 0; int v = atoi(f); return (v >= lo && v <= hi) ? v : 0; };
    g_env.upt = geti("SSAD_V7_UPT", 1, 64);
    g_env.nofast = getenv("SSAD_V7_NOFAST") ? 1 : 0;
    g_env.nc = geti("SSAD_V7_NC", 1, 32);
    g_env.lanes = geti("SSAD_V7_LANES", 1, 8);
    g_env.chunk_mb = geti("SSAD_V7_CHUNK_MB", 1, 4096);
    g_env.split = geti("SSAD_V7_SPLIT", 1, 64);
    g_env.nopdl = getenv("SSAD_V7_NOPDL") ? 1 : 0;
}
static const PipeEnv& env() {
    std::call_once(g_env_once, load_env);
    return g_env;
}
extern "C" void ssad_reload_env(void) {
    (void)env();
    load_env();
}

static bool make_geo7(const SsadBatchDesc& d, Geo7& g) {
    const PipeEnv& E = env();
    g.H = d.height; g.W = d.width; g.batch = d.batch;
    if (g.W % 8 != 0 || g.W < 8) return false;
    if (!(d.sat_factor_max <= kSatMax)) return false;       // the table after Saturation covers factors up to kSatMax
    g.upr = g.W / 8;
    const int upt = E.upt > 0 ? E.upt : UPT;                // units per thread and slab
    g.rows_slab = (NT * upt + g.upr - 1) / g.upr;
    if (g.rows_slab < 1) g.rows_slab = 1;
    if (g.rows_slab > g.H) g.rows_slab = g.H;
    g.nslab = (g.H + g.rows_slab - 1) / g.rows_slab;
    g.fast = (NT % g.upr == 0 && upt == UPT && g.rows_slab * g.upr == NT * UPT && g.H % g.rows_slab == 0) ? 1 : 0;
    if (E.nofast) g.fast = 0;
    g.nslab_shift = -1;
    for (int k = 0; k < 16; ++k) if (g.nslab == (1 << k)) g.nslab_shift = k;
    g.inv_upr = (65536u + (uint32_t)g.upr - 1u) / (uint32_t)g.upr;
    // canvas CTAs per image, per chunk; measured: fewer, longer canvas CTAs win once the grid fills the GPU, larger
    // images (taller canvases) want more
    g.ncanvas = (d.batch >= 1024 ? 1 : 2) * (d.height >= 1024 ? 2 : 1);
    if (E.nc > 0) g.ncanvas = E.nc;
    while ((d.max_patch_h + g.ncanvas - 1) / g.ncanvas > kCanvasRows) ++g.ncanvas;
    g.cov_pitch = (uint32_t)(d.max_patch_w + 31) / 32u;
    g.cov_rows = (uint32_t)d.max_patch_h;
    g.cov_stride = (size_t)g.cov_pitch * g.cov_rows;
    const int cw = d.max_scar_w > 0 ? d.max_scar_w : d.max_patch_w, ch = d.max_scar_h > 0 ? d.max_scar_h : d.max_patch_h;
    g.canvas_pitch = (uint32_t)cw;                          // the colour canvas holds rotated scars only
    g.canvas_stride = (size_t)cw * (size_t)ch;
    auto up = [](size_t v) { return (v + 255) / 256 * 256; };
    g.o_canvas = up((size_t)d.batch * d.height * d.width * 3);
    g.o_state = g.o_canvas + up((size_t)d.batch * g.canvas_stride * 4);
    g.o_cov = g.o_state + up((size_t)d.batch * sizeof(ImgState));
    g.o_sched = g.o_cov + up((size_t)d.batch * g.cov_stride * 4);
    g.o_hdr = g.o_sched + up(sizeof(Sched));
    g.ws_total = g.o_hdr + up((size_t)2 * d.batch * sizeof(Hdr));
    g.smem_plan = (((uint32_t)g.H + 1u) * 4u + 15u) & ~15u;
    g.src_indexed = d.src_indexed ? 1 : 0;
    return true;
}

}  // namespace v7
}  // namespace ssad

// A call is cut into chunks that run on `lanes` streams (chunk c on lane c % lanes, one workspace region per lane, so
// the x buffer of a lane is rewritten while it is still in the L2 and pass 2 reads it from there).  Lanes: 2 for plain
// launches (every chunk is 6 launches: more would make the host the bottleneck), 4 when the call is being captured
// into a CUDA graph (one cudaGraphLaunch replays everything).  Chunks: one per lane (more only on request: SSAD_V7_CHUNK_MB
// caps a chunk's x buffer, SSAD_V7_SPLIT pins the number; both measured slower than one large chunk per lane).
struct V7Split { int nchunk, csize, nlanes; };
static V7Split v7_split(const SsadBatchDesc& d, bool capturing) {
    const ssad::v7::PipeEnv& E = ssad::v7::env();
    int lanes = capturing ? 4 : 2;
    if (E.lanes > 0) lanes = E.lanes;
    const size_t mb = (size_t)E.chunk_mb;         // measured: smaller chunks lose more to the latency chains than the L2 gives back
    int n = 1;
    if (d.batch >= 256) {
        const size_t total = (size_t)d.height * d.width * 3 * (size_t)d.batch, lim = mb << 20;
        const size_t by_size = lim ? (total + lim - 1) / lim : 0;
        n = lanes;
        if (by_size > (size_t)n) n = by_size > 64 ? 64 : (int)by_size;
    }
    if (E.split > 0) n = E.split;
    if (n > d.batch) n = d.batch;
    if (n < 1) n = 1;
    V7Split sp;
    sp.csize = (d.batch + n - 1) / n;
    sp.nchunk = (d.batch + sp.csize - 1) / sp.csize;        // no empty chunk
    sp.nlanes = lanes < sp.nchunk ? lanes : sp.nchunk;
    return sp;
}
static size_t v7_chunk_workspace(const SsadBatchDesc& d, int csize) {
    ssad::v7::Geo7 g;
    SsadBatchDesc dc = d;
    dc.batch = csize;
    if (!ssad::v7::make_geo7(dc, g)) return 0;
    return g.ws_total;
}

size_t ssad_pipeline_workspace(const SsadBatchDesc& d) {           // enough for a plain and for a captured call
    const V7Split s0 = v7_split(d, false), s1 = v7_split(d, true);
    const size_t a = v7_chunk_workspace(d, s0.csize) * (size_t)s0.nlanes, b = v7_chunk_workspace(d, s1.csize) * (size_t)s1.nlanes;
    return a > b ? a : b;
}

// kernels one ssad_generate call launches (bench.py's gpu_launches)
int ssad_pipeline_launches(const SsadBatchDesc& d, bool capturing) { return 6 * v7_split(d, capturing).nchunk; }

static int launch_v7_chunk(const SsadBatchDesc& d, const SsadImageParams* params, const uint8_t* src, const uint8_t* donor_bank,
                           const uint32_t* mask_bits, const void* norm_lut, uint8_t* out_u8, void* out_norm, int64_t* y_out,
                           SsadDebug* debug, void* workspace, size_t workspace_bytes, cudaStream_t stream, bool use_pdl);

// A context: the helper streams and fork / join events of one device (ssad_context_create).
struct SsadContext {
    int device;
    cudaStream_t s[8];
    cudaEvent_t fork, join[8];
};
void ssad_pipeline_context_delete(SsadContext* c);
SsadContext* ssad_pipeline_context_new(int device) {
    SsadContext* c = new SsadContext();
    memset(c, 0, sizeof(*c));
    c->device = device;
    bool ok = cudaEventCreateWithFlags(&c->fork, cudaEventDisableTiming) == cudaSuccess;
    for (int i = 1; i < 8 && ok; ++i)
        ok = cudaStreamCreateWithFlags(&c->s[i], cudaStreamNonBlocking) == cudaSuccess &&
             cudaEventCreateWithFlags(&c->join[i], cudaEventDisableTiming) == cudaSuccess;
    if (!ok) { ssad_pipeline_context_delete(c); return nullptr; }
    return c;
}
void ssad_pipeline_context_delete(SsadContext* c) {
    if (!c) return;
    for (int i = 1; i < 8; ++i) {
        if (c->s[i]) cudaStreamDestroy(c->s[i]);
        if (c->join[i]) cudaEventDestroy(c->join[i]);
    }
    if (c->fork) cudaEventDestroy(c->fork);
    delete c;
}
int ssad_pipeline_context_device(const SsadContext* c) { return c->device; }

// returns SSAD_OK when launched, SSAD_EUNSUPPORTED for a request this library cannot serve
int ssad_pipeline_launch(SsadContext* HS, const SsadBatchDesc& d, const SsadImageParams* params, const uint8_t* src,
                         const uint8_t* donor_bank, const uint32_t* mask_bits, const void* norm_lut, uint8_t* out_u8, void* out_norm,
                         int64_t* y_out, SsadDebug* debug, void* workspace, size_t workspace_bytes, cudaStream_t stream) {
    cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
    if (cudaStreamIsCapturing(stream, &cap) != cudaSuccess) { cap = cudaStreamCaptureStatusNone; (void)cudaGetLastError(); }
    const bool capturing = cap == cudaStreamCaptureStatusActive;
    // programmatic dependent launch pays on plain streams; inside a graph the programmatic edges cost more than they hide
    const bool use_pdl = !capturing && !ssad::v7::env().nopdl;
    const V7Split sp = v7_split(d, capturing);
    const int nchunk = sp.nchunk, csize = sp.csize, nlanes = sp.nlanes;
    const size_t need = ssad_pipeline_workspace(d);
    if (need == 0) return SSAD_EUNSUPPORTED;                 // width not a multiple of 8, or a Saturation factor above 2
    if (((long long)d.height * d.width) % 8 != 0) return SSAD_EUNSUPPORTED;
    if ((reinterpret_cast<uintptr_t>(src) & 15) || (d.src_pitch & 7) || (d.src_stride & 7)) return SSAD_EUNSUPPORTED;
    if (nchunk > 1 && (d.src_stride & 15)) return SSAD_EUNSUPPORTED;      // every chunk's first image must stay 16-byte aligned
    if (!workspace || workspace_bytes < need) return SSAD_EWORKSPACE;
    if (reinterpret_cast<uintptr_t>(workspace) & 255) return SSAD_EALIGN;
    const size_t ws_chunk = v7_chunk_workspace(d, csize);
    cudaError_t e = cudaSuccess;
    if (nlanes > 1) {
        e = cudaEventRecord(HS->fork, stream);
        for (int l = 1; l < nlanes && e == cudaSuccess; ++l) e = cudaStreamWaitEvent(HS->s[l], HS->fork, 0);
        if (e != cudaSuccess) return ssad_set_cuda_error(e);
    }
    const size_t npix = (size_t)d.height * d.width;
    const size_t norm_elem = (d.out_norm_dtype == SSAD_NORM_F32) ? 4 : 2;
    int rc = SSAD_OK;
    for (int c = 0; c < nchunk; ++c) {
        const int b0 = c * csize, nb = (d.batch - b0 < csize) ? d.batch - b0 : csize;
        if (nb <= 0) break;
        SsadBatchDesc dc = d;
        dc.batch = nb;
        if (dc.line_xy) dc.line_xy += (size_t)b0 * d.line_max * 2;
        const uint32_t* mb = mask_bits;
        if (mb && d.mask_mode == SSAD_MASK_PER_IMAGE && !d.src_indexed) mb += (size_t)b0 * d.mask_stride_words;
        const int lane = c % nlanes;             // chunks of a lane run one after the other and share its workspace
        cudaStream_t st = (lane == 0) ? stream : HS->s[lane];
        rc = launch_v7_chunk(dc, params + b0, d.src_indexed ? src : src + (size_t)b0 * d.src_stride, donor_bank, mb, norm_lut,
                             out_u8 ? out_u8 + (size_t)b0 * d.out_u8_stride : nullptr,
                             out_norm ? static_cast<uint8_t*>(out_norm) + (size_t)b0 * 3 * npix * norm_elem : nullptr,
                             y_out ? y_out + b0 : nullptr, debug ? debug + b0 : nullptr,
                             static_cast<uint8_t*>(workspace) + (size_t)lane * ws_chunk, ws_chunk, st, use_pdl);
        if (rc != SSAD_OK) break;
    }
    // join (also after a failed launch: the helper streams must not stay forked inside a capture)
    for (int l = 1; l < nlanes; ++l) {
        cudaError_t e2 = cudaEventRecord(HS->join[l], HS->s[l]);
        if (e2 == cudaSuccess) e2 = cudaStreamWaitEvent(stream, HS->join[l], 0);
        if (e2 != cudaSuccess && rc == SSAD_OK) rc = ssad_set_cuda_error(e2);
    }
    return rc;
}

static int launch_v7_chunk(const SsadBatchDesc& d, const SsadImageParams* params, const uint8_t* src, const uint8_t* donor_bank,
                           const uint32_t* mask_bits, const void* norm_lut, uint8_t* out_u8, void* out_norm, int64_t* y_out,
                           SsadDebug* debug, void* workspace, size_t workspace_bytes, cudaStream_t stream, bool use_pdl) {
    using namespace ssad::v7;
    Geo7 g;
    if (!make_geo7(d, g)) return SSAD_EUNSUPPORTED;
    if ((reinterpret_cast<uintptr_t>(src) & 15) || (d.src_pitch & 7) || (d.src_stride & 7)) return SSAD_EUNSUPPORTED;
    if (g.smem_plan > 40u * 1024u) return SSAD_EUNSUPPORTED;
    if ((long long)d.batch * g.nslab > 0x7FFFFFFFll) return SSAD_EUNSUPPORTED;
    if (!workspace || workspace_bytes < g.ws_total) return SSAD_EWORKSPACE;
    if (reinterpret_cast<uintptr_t>(workspace) & 255) return SSAD_EALIGN;
    uint8_t* const ws = static_cast<uint8_t*>(workspace);
    uint8_t* const xbuf = ws;
    uint32_t* const canvas = reinterpret_cast<uint32_t*>(ws + g.o_canvas);
    ImgState* const state = reinterpret_cast<ImgState*>(ws + g.o_state);
    uint32_t* const cov = reinterpret_cast<uint32_t*>(ws + g.o_cov);
    Sched* const sched = reinterpret_cast<Sched*>(ws + g.o_sched);
    Hdr* const hdr = reinterpret_cast<Hdr*>(ws + g.o_hdr);
    const unsigned B = (unsigned)d.batch;
    cudaError_t e = cudaMemsetAsync(sched, 0, sizeof(Sched), stream);
    if (e != cudaSuccess) return ssad_set_cuda_error(e);
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    auto cfg_of = [&](unsigned grid, size_t smem, bool pdl) {
        cudaLaunchConfig_t c;
        memset(&c, 0, sizeof(c));
        c.gridDim = dim3(grid, 1, 1); c.blockDim = dim3(NT, 1, 1); c.dynamicSmemBytes = smem; c.stream = stream;
        c.attrs = attr; c.numAttrs = (pdl && use_pdl) ? 1 : 0;
        return c;
    };
    const Hdr* const hdr2 = hdr + d.batch;
    const uint8_t* const xbuf_c = xbuf;
    const ImgState* const state_c = state;
    const Sched* const sched_c = sched;
    const uint32_t* const cov_c = cov;
    cudaLaunchConfig_t c0 = cfg_of(B, g.smem_plan, false);          // follows the memset: a plain launch
    e = cudaLaunchKernelEx(&c0, plan_v7_kernel, d, g, params, mask_bits, state, cov, sched);
    cudaLaunchConfig_t c1 = cfg_of((B + NT - 1) / NT, 0, true);
    if (e == cudaSuccess) e = cudaLaunchKernelEx(&c1, lists_v7_kernel, g, params, state_c, sched_c, hdr);
    cudaLaunchConfig_t c2 = cfg_of(B * (unsigned)g.nslab, 0, true);
    if (e == cudaSuccess) e = cudaLaunchKernelEx(&c2, pass1_v7_kernel, d, g, params, src, donor_bank, state, (const Hdr*)hdr, xbuf);
    cudaLaunchConfig_t c3 = cfg_of(B * (unsigned)g.ncanvas, 0, true);
    if (e == cudaSuccess) e = cudaLaunchKernelEx(&c3, canvas_v7_kernel, d, g, params, src, donor_bank, state, cov_c, canvas, xbuf);
    cudaLaunchConfig_t c4 = cfg_of(B, 0, true);
    if (e == cudaSuccess) e = cudaLaunchKernelEx(&c4, finish_v7_kernel, d, g, state, y_out, debug);
    cudaLaunchConfig_t c5 = cfg_of(B * (unsigned)g.nslab, 0, true);
    if (e == cudaSuccess) e = cudaLaunchKernelEx(&c5, pass2_v7_kernel, d, g, src, state_c, hdr2, xbuf_c, norm_lut, out_u8, out_norm);
    if (e != cudaSuccess) return ssad_set_cuda_error(e);
    e = cudaGetLastError();
    if (e != cudaSuccess) return ssad_set_cuda_error(e);
    return SSAD_OK;
}
