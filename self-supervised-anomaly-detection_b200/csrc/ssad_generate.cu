// Host side of the C ABI for ssad_generate: argument validation, the per-device contexts (helper streams of the kernel
// pipeline), error texts.  The kernels live in ssad_generate_pipeline.cu -- the one generator path of this library:
// a request it cannot serve returns SSAD_EUNSUPPORTED, there is no other kernel to fall back to.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include <mutex>

#include "../../include/ssad.h"

int ssad_pipeline_launch(SsadContext* ctx, const SsadBatchDesc& d, const SsadImageParams* params, const uint8_t* src,
                         const uint8_t* donor_bank, const uint32_t* mask_bits, const void* norm_lut, uint8_t* out_u8, void* out_norm,
                         int64_t* y_out, SsadDebug* debug, void* workspace, size_t workspace_bytes, cudaStream_t stream);
size_t ssad_pipeline_workspace(const SsadBatchDesc& d);
int ssad_pipeline_launches(const SsadBatchDesc& d, bool capturing);
SsadContext* ssad_pipeline_context_new(int device);
void ssad_pipeline_context_delete(SsadContext* ctx);
int ssad_pipeline_context_device(const SsadContext* ctx);

static thread_local char g_cuda_err[256] = "";

int ssad_set_cuda_error(cudaError_t e) {
    strncpy(g_cuda_err, cudaGetErrorString(e), sizeof(g_cuda_err) - 1);
    return SSAD_ELAUNCH;
}

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
        case SSAD_ESTATIC: return "request needs more scratch than the static device array holds (SSAD_OBJ_MASK_STATIC_BYTES = 16 MiB): "
                                  "use ssad_obj_mask with a caller workspace, or a smaller batch";
        default: return "unknown error";
    }
}

// ------------------------------------------------------------------ contexts
// A context owns the helper streams / events one ssad_generate call forks onto; calls that share a context must not
// overlap on the host.  ssad_generate uses one default context per device (created on first use, under a lock);
// host threads that drive the same device concurrently create their own with ssad_context_create.
extern "C" int ssad_context_create(SsadContext** out) {
    if (!out) return SSAD_EINVAL;
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return ssad_set_cuda_error(e);
    SsadContext* c = ssad_pipeline_context_new(dev);
    if (!c) return SSAD_ELAUNCH;
    *out = c;
    return SSAD_OK;
}

extern "C" void ssad_context_destroy(SsadContext* ctx) {
    if (ctx) ssad_pipeline_context_delete(ctx);
}

static SsadContext* default_context() {
    static std::mutex mu;
    static SsadContext* per_dev[64];
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return nullptr;
    std::lock_guard<std::mutex> lock(mu);
    if (!per_dev[dev]) per_dev[dev] = ssad_pipeline_context_new(dev);
    return per_dev[dev];
}

extern "C" size_t ssad_generate_workspace(const SsadBatchDesc* desc) {
    if (!desc || desc->batch <= 0) return 0;
    return ssad_pipeline_workspace(*desc);
}

extern "C" int ssad_generate_launches(const SsadBatchDesc* desc, int captured) {
    if (!desc || desc->batch <= 0) return 0;
    return ssad_pipeline_launches(*desc, captured != 0);
}

extern "C" int ssad_generate_ctx(SsadContext* ctx, const SsadBatchDesc* desc, const SsadImageParams* params, const uint8_t* src,
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
    if (d.src_indexed && (d.src_count <= 0)) return SSAD_EINVAL;
    if (d.donor_count < 0 || (d.donor_count > 0 && !donor_bank)) return SSAD_EINVAL;
    if (d.mask_mode != SSAD_MASK_ALL && (d.mask_height < d.src_height || d.mask_pitch_words * 32 < d.src_width)) return SSAD_EINVAL;
    if ((reinterpret_cast<uintptr_t>(params) & 15) || (reinterpret_cast<uintptr_t>(out_u8) & 15) ||
        (reinterpret_cast<uintptr_t>(out_norm) & 15))
        return SSAD_EALIGN;
    if (out_u8 && (d.out_u8_pitch != 3ll * d.width || d.out_u8_stride != 3ll * d.width * d.height)) return SSAD_EUNSUPPORTED;
    if (!ctx) ctx = default_context();
    if (!ctx) return SSAD_ELAUNCH;
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess || dev != ssad_pipeline_context_device(ctx)) return SSAD_EINVAL;   // a context belongs to its device
    return ssad_pipeline_launch(ctx, d, params, src, donor_bank, mask_bits, norm_lut, out_u8, out_norm, y_out, debug, workspace,
                                workspace_bytes, static_cast<cudaStream_t>(stream));
}

extern "C" int ssad_generate(const SsadBatchDesc* desc, const SsadImageParams* params, const uint8_t* src,
                             const uint8_t* donor_bank, const uint32_t* mask_bits, const void* norm_lut,
                             uint8_t* out_u8, void* out_norm, int64_t* y_out, SsadDebug* debug,
                             void* workspace, size_t workspace_bytes, void* stream) {
    return ssad_generate_ctx(nullptr, desc, params, src, donor_bank, mask_bits, norm_lut, out_u8, out_norm, y_out, debug, workspace,
                             workspace_bytes, stream);
}
