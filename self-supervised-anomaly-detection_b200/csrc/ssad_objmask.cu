// obj_mask (reference dataset_generator.py:27-39) on device -- placeholder until the multi-pass
// canny / morphology / connected-components pipeline lands.
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/ssad.h"

extern "C" size_t ssad_obj_mask_workspace(int32_t batch, int32_t height, int32_t width) {
    (void)batch; (void)height; (void)width;
    return 0;
}

extern "C" int ssad_obj_mask(int32_t batch, int32_t height, int32_t width, const uint8_t* src, int64_t src_pitch,
                             int64_t src_stride, uint32_t* mask_bits, int32_t mask_pitch_words, int64_t mask_stride_words,
                             void* workspace, size_t workspace_bytes, void* stream) {
    (void)batch; (void)height; (void)width; (void)src; (void)src_pitch; (void)src_stride; (void)mask_bits;
    (void)mask_pitch_words; (void)mask_stride_words; (void)workspace; (void)workspace_bytes; (void)stream;
    return SSAD_EUNSUPPORTED;
}
