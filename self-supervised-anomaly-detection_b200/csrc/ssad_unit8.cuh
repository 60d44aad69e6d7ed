// 8-pixel register unit of the jitter chain + the similarity decision, shared by the generator kernels.
//
// A unit is 8 consecutive pixels of one row (24 bytes = 6 words) held as 12 packed f32x2 values in "magic" form
// (2^23 + v).  ColorJitter's Saturation (torchvision _functional_pil.py:71-97 -> ImageEnhance.Color -> Blend.c) runs on
// them in ssad_pipeline_common.cuh with every product and sum rounded on its own (bit-identical to Pillow's float32
// code); Brightness / Contrast are 256-entry tables; RGB -> L is Convert.c's integer formula.
#pragma once
#include "ssad_device.cuh"
#include "ssad_sm100.cuh"

namespace ssad {
namespace u8x {

// ---------------------------------------------------------------- 8-pixel unit, packed pairs
// v[c][k] = (pixel 2k, pixel 2k+1) of channel c in magic form (2^23 + value)
struct Unit8 {
    f32x2 v[3][4];
};

__device__ __forceinline__ void unit_unpack(Unit8& u, const uint2& a, const uint2& b, const uint2& c) {
    const uint32_t w[6] = {a.x, a.y, b.x, b.y, c.x, c.y};
    #pragma unroll
    for (int k = 0; k < 4; ++k) {
        #pragma unroll
        for (int ch = 0; ch < 3; ++ch) {
            const int b0 = 6 * k + ch, b1 = 6 * k + 3 + ch;
            u.v[ch][k] = pack2(magic_from_byte(w[b0 >> 2], b0 & 3), magic_from_byte(w[b1 >> 2], b1 & 3));
        }
    }
}
__device__ __forceinline__ void unit_scalars(const Unit8& u, float (&m)[24]) {
    #pragma unroll
    for (int k = 0; k < 4; ++k) {
        #pragma unroll
        for (int ch = 0; ch < 3; ++ch) unpack2(u.v[ch][k], m[6 * k + ch], m[6 * k + 3 + ch]);
    }
}
__device__ __forceinline__ void unit_pack(const Unit8& u, uint2& a, uint2& b, uint2& c) {
    float m[24];
    unit_scalars(u, m);
    uint32_t w[6];
    #pragma unroll
    for (int i = 0; i < 6; ++i) w[i] = magic_pack4(m[4 * i], m[4 * i + 1], m[4 * i + 2], m[4 * i + 3]);
    a = make_uint2(w[0], w[1]);
    b = make_uint2(w[2], w[3]);
    c = make_uint2(w[4], w[5]);
}

__device__ __forceinline__ uint32_t luma_bits(uint32_t r, uint32_t g, uint32_t b) {
    return 19595u * r + 38470u * g + 7471u * b + 0x8000u;     // on magic bit patterns: the 0x4B000000 offsets cancel mod 2^32
}
// check_color_similarity(...) > 0.99 from exact integer channel sums (dataset_generator.py:147-159): the cheap
// cosine <s1,s2>/sqrt(<s1,s1><s2,s2>) decides unless it is within 1e-9 of the threshold; then the reference's own
// operation order (mean / 255, normalise, dot) is evaluated.
__device__ __forceinline__ bool similarity_branch(const uint32_t* s1, double n1, const uint32_t* s2, double n2, double& cosv) {
    const double a0 = s1[0], a1 = s1[1], a2 = s1[2], b0 = s2[0], b1 = s2[1], b2 = s2[2];
    const double dot = a0 * b0 + a1 * b1 + a2 * b2;
    const double nn = (a0 * a0 + a1 * a1 + a2 * a2) * (b0 * b0 + b1 * b1 + b2 * b2);
    double c = (nn > 0.0) ? dot * rsqrt(nn) : 0.0;
    if (fabs(c - 0.99) < 1e-9 || nn == 0.0) c = cosine_of_means(s1, n1, s2, n2);
    cosv = c;
    return c > 0.99;
}

}  // namespace u8x
}  // namespace ssad
