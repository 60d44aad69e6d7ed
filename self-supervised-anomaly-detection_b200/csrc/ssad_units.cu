// Unit entry points of the C ABI: each reference routine of the path on its own, for parity tests
// (rect2poly / ImageDraw.polygon, Image.transform AFFINE NEAREST, ColorJitter, coords_map[k], ToTensor).
#include <cuda_runtime.h>
#include <stdint.h>

#include "ssad_device.cuh"

int ssad_set_cuda_error(cudaError_t e);

namespace ssad {

// ---- rect2poly(patch, regular=False, sides=8) -> alpha of the RGBA mask (dataset_generator.py:63-100)
__global__ void polygon_mask_kernel(int n, const int16_t* __restrict__ poly_xy, const uint8_t* __restrict__ poly_n,
                                    const int16_t* __restrict__ size, uint8_t* __restrict__ out, int max_w, int max_h) {
    extern __shared__ uint32_t rows[];          // blockDim.x rows of pitch words
    const int i = blockIdx.x;
    if (i >= n) return;
    const int pw = size[2 * i], ph = size[2 * i + 1];
    const int pitch = (max_w + 31) / 32;
    __shared__ int16_t xy[8][2];
    if (threadIdx.x < 16) (&xy[0][0])[threadIdx.x] = poly_xy[16 * i + threadIdx.x];
    __syncthreads();
    Polygon G;
    polygon_setup(G, xy, poly_n[i], ph);
    uint32_t* row = rows + (size_t)threadIdx.x * pitch;
    uint8_t* o = out + (size_t)i * max_w * max_h;
    for (int r = threadIdx.x; r < max_h; r += blockDim.x) {
        for (int j = 0; j < pitch; ++j) row[j] = 0;
        if (r < ph) polygon_fill_row(G, r, pw, row);
        for (int x = 0; x < max_w; ++x) o[(size_t)r * max_w + x] = ((row[x >> 5] >> (x & 31)) & 1u) ? 255 : 0;
    }
}

// ---- Image.transform(size, AFFINE, coeffs, NEAREST) via affine_fixed (Geometry.c)
struct Fix6 { int32_t a[6]; };
__global__ void affine_nearest_kernel(const uint8_t* __restrict__ src, int sw, int sh, int ch, Fix6 f,
                                      uint8_t* __restrict__ out, int ow, int oh) {
    int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= ow * oh) return;
    int y = p / ow, x = p - y * ow;
    int xin, yin;
    affine_fixed_map(f.a, x, y, xin, yin);
    bool ok = xin >= 0 && xin < sw && yin >= 0 && yin < sh;
    for (int c = 0; c < ch; ++c) out[(size_t)p * ch + c] = ok ? src[((size_t)yin * sw + xin) * ch + c] : 0;
}

// ---- ColorJitter.forward (torchvision transforms.py:1259-1290) on PIL semantics, one CTA per image
__global__ void __launch_bounds__(256)
color_jitter_kernel(int height, int width, const uint8_t* __restrict__ src, uint8_t* __restrict__ out,
                    const uint8_t* __restrict__ order, const float* __restrict__ factors, const uint8_t* __restrict__ enable) {
    __shared__ uint32_t red[64];
    const int b = blockIdx.x, tid = threadIdx.x;
    const int npix = height * width;
    const uint8_t* s = src + (size_t)b * npix * 3;
    uint8_t* o = out + (size_t)b * npix * 3;
    for (int i = tid; i < npix * 3; i += blockDim.x) o[i] = s[i];
    __syncthreads();
    for (int st = 0; st < 4; ++st) {
        const int fn = order[4 * b + st];
        if (fn >= 3 || !((enable[b] >> fn) & 1)) continue;
        const float f = factors[4 * b + fn];
        const bool clip = blend_needs_clip(f);
        int mean = 0;
        if (fn == 1) {
            uint32_t ls = 0;
            for (int p = tid; p < npix; p += blockDim.x) ls += luma_u8(o[3 * p], o[3 * p + 1], o[3 * p + 2]);
            ls = block_sum_u32<256>(ls, red);
            mean = (int)((2ull * ls + (uint32_t)npix) / (2ull * (uint32_t)npix));
        }
        for (int p = tid; p < npix; p += blockDim.x) {
            uint32_t L = luma_u8(o[3 * p], o[3 * p + 1], o[3 * p + 2]);
            for (int c = 0; c < 3; ++c) {
                float mB = magic_from_u8(o[3 * p + c]);
                uint32_t A = (fn == 0) ? 0u : (fn == 1 ? (uint32_t)mean : L);
                float r = clip ? magic_blend<true>(mB, magic_from_u8(A), (float)A, f)
                               : magic_blend<false>(mB, magic_from_u8(A), (float)A, f);
                o[3 * p + c] = (uint8_t)magic_to_u8(r);
            }
        }
        __syncthreads();
    }
}

// ---- coords_map[k] inside a window of a bit-packed mask (datasets.py:263-264)
__global__ void __launch_bounds__(256)
select_kth_kernel(const uint32_t* __restrict__ mask, int pitch, int wx, int wy, int ww, int wh, int n,
                  const uint32_t* __restrict__ ks, int32_t* __restrict__ out_xy, int32_t* __restrict__ count_out) {
    extern __shared__ uint32_t pre[];           // wh + 1
    const int tid = threadIdx.x;
    for (int r = tid; r < wh; r += blockDim.x) pre[r + 1] = row_window_popc(mask + (size_t)(wy + r) * pitch, wx, ww);
    __syncthreads();
    if (tid == 0) {
        pre[0] = 0;
        for (int r = 0; r < wh; ++r) pre[r + 1] += pre[r];
        if (count_out) *count_out = (int32_t)pre[wh];
    }
    __syncthreads();
    for (int i = tid; i < n; i += blockDim.x) {
        uint32_t k = ks[i];
        int x = -1, y = -1;
        if (k < pre[wh]) {
            int lo = 0, hi = wh;
            while (hi - lo > 1) { int mid = (lo + hi) >> 1; if (pre[mid] <= k) lo = mid; else hi = mid; }
            y = lo;
            x = row_window_select(mask + (size_t)(wy + y) * pitch, wx, ww, k - pre[lo]);
        }
        out_xy[2 * i] = x; out_xy[2 * i + 1] = y;
    }
}

// ---- ToTensor(original) (datasets.py:394): HWC u8 -> CHW f32, value / 255 (IEEE division as torch)
__global__ void to_tensor_kernel(int height, int width, const uint8_t* __restrict__ src, int64_t pitch, int64_t stride,
                                 float* __restrict__ out) {
    const int b = blockIdx.y;
    const int npix = height * width;
    int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= npix) return;
    int y = p / width, x = p - y * width;
    const uint8_t* q = src + (size_t)b * stride + (size_t)y * pitch + (size_t)x * 3;
    float* o = out + (size_t)b * 3 * npix;
    o[p] = __fdiv_rn((float)q[0], 255.0f);
    o[npix + p] = __fdiv_rn((float)q[1], 255.0f);
    o[2 * npix + p] = __fdiv_rn((float)q[2], 255.0f);
}

// ---- upload of only what a patch-mode call reads of a host image (ssad_upload_windows) -------------------------------
// src is pinned host memory the device can address; dst the device copy with the same layout.  Per image: the working
// window (crop_off, win_w x win_h) and, for labels 1 / 2 whose patch comes from the image itself, the cut rectangle
// (donor_off + rect).  One warp per row segment, 16-byte blocks that cover it (source and destination share their
// alignment; the few extra bytes either side are the image's own pixels).
__global__ void __launch_bounds__(256)
upload_windows_kernel(int n, const SsadImageParams* __restrict__ params, const uint8_t* __restrict__ src, uint8_t* __restrict__ dst,
                      long long pitch, long long stride, int src_h, int src_w, int win_h, int win_w, int rows) {
    // a small persistent grid (the launch sizes it): the kernel waits on PCIe, not on SMs, and must leave them to the
    // generation kernels of the previous piece of the batch that run beside it
    const int lane = threadIdx.x & 31;
    const long long nwarp = (long long)gridDim.x * 8, ntask = (long long)n * rows;
    for (long long t = (long long)blockIdx.x * 8 + (threadIdx.x >> 5); t < ntask; t += nwarp) {
        const int b = (int)(t / rows), r = (int)(t - (long long)b * rows);
        const SsadImageParams* const p = params + b;
        const bool a = r < win_h;
        int x, y, w;
        if (a) {
            x = p->crop_off[0]; y = p->crop_off[1] + r; w = win_w;
        } else {
            const bool cut = (p->label == 1 || p->label == 2) && p->donor_idx < 0;
            if (!cut || r - win_h >= p->rect[3]) continue;
            x = p->donor_off[0] + p->rect[0]; y = p->donor_off[1] + p->rect[1] + (r - win_h); w = p->rect[2];
        }
        if (y < 0 || y >= src_h) continue;
        const int x0 = max(x, 0), x1 = min(x + w, src_w);
        if (x1 <= x0) continue;
        const long long img = (long long)b * stride;
        const long long lo = ((long long)y * pitch + 3ll * x0) & ~15ll, hi = min(((long long)y * pitch + 3ll * x1 + 15) & ~15ll, stride);
        const int4* s4 = reinterpret_cast<const int4*>(src + img + lo);
        int4* d4 = reinterpret_cast<int4*>(dst + img + lo);
        const int nb = (int)((hi - lo) >> 4);
        for (int k = lane; k < nb; k += 32) d4[k] = s4[k];
    }
}

}  // namespace ssad

#define SSAD_CHECK_LAUNCH()                                   \
    do {                                                      \
        cudaError_t e_ = cudaGetLastError();                  \
        if (e_ != cudaSuccess) return ssad_set_cuda_error(e_); \
    } while (0)

namespace ssad {
// ---- Image.paste(patch, (left, top), mask) (dataset_generator.py:268-275; Pillow Paste.c for alpha in {0, 255})
__global__ void __launch_bounds__(256)
paste_kernel(uint8_t* __restrict__ img, int W, int H, const uint8_t* __restrict__ patch, int pw, int ph, int ch,
             const uint8_t* __restrict__ mask, int left, int top) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= pw * ph) return;
    const int v = i / pw, u = i - v * pw;
    const int X = left + u, Y = top + v;
    if (X < 0 || Y < 0 || X >= W || Y >= H) return;            // the box is clipped to the image
    const uint32_t a = mask ? mask[i] : (ch == 4 ? patch[(size_t)i * 4 + 3] : 255u);
    if (a == 0) return;
    uint8_t* q = img + 3 * ((size_t)Y * W + X);
    const uint8_t* s = patch + (size_t)i * ch;
    if (a == 255) { q[0] = s[0]; q[1] = s[1]; q[2] = s[2]; return; }      // (on the dataset's path alpha is 0 or 255)
    // Paste.c BLEND8 for a soft mask: DIV255(out * (255 - a) + in * a), DIV255(t) = ((t + 128) >> 8) + (t + 128) >> 8
    #pragma unroll
    for (int c = 0; c < 3; ++c) {
        const uint32_t t = (uint32_t)q[c] * (255u - a) + (uint32_t)s[c] * a + 128u;
        q[c] = (uint8_t)(((t >> 8) + t) >> 8);
    }
}

// ---- ImageDraw.line(points, fill, width) (datasets.py:383-388): the pixels set, one CTA per poly-line
__global__ void __launch_bounds__(256)
line_mask_kernel(const int16_t* __restrict__ xy, const uint8_t* __restrict__ npts, int max_pts, int width, uint8_t* __restrict__ out,
                 int W, int H) {
    extern __shared__ uint32_t cov[];            // H rows of ceil(W / 32) words (box coordinates: row stride = box pitch)
    __shared__ Polygon quads[31];
    __shared__ LineBox B;
    const int i = blockIdx.x, tid = threadIdx.x;
    const int16_t* p = xy + (size_t)i * max_pts * 2;
    const int n = min((int)npts[i], min(max_pts, 32));
    if (tid == 0) B = line_box(p, n, width, W, H);
    for (int k = tid; k < H * ((W + 31) / 32); k += 256) cov[k] = 0u;
    __syncthreads();
    const int pitch = (B.w + 31) / 32;
    line_rasterize(p, n, width, W, H, B, cov, max(pitch, 1), quads, tid, 256);
    __syncthreads();
    uint8_t* o = out + (size_t)i * W * H;
    for (int k = tid; k < W * H; k += 256) {
        const int y = k / W, x = k - y * W;
        const int u = x - B.x0, v = y - B.y0;
        const bool on = B.w > 0 && (unsigned)u < (unsigned)B.w && (unsigned)v < (unsigned)B.h && ((cov[v * pitch + (u >> 5)] >> (u & 31)) & 1u);
        o[k] = on ? 255 : 0;
    }
}

// ---- check_color_similarity (dataset_generator.py:147-159): cosine of the mean colours of two images
__global__ void __launch_bounds__(256)
color_similarity_kernel(const uint8_t* __restrict__ a, int na, const uint8_t* __restrict__ b, int nb, double* __restrict__ out) {
    __shared__ unsigned long long acc[6];
    const int tid = threadIdx.x;
    if (tid < 6) acc[tid] = 0ull;
    __syncthreads();
    unsigned long long s[6] = {0, 0, 0, 0, 0, 0};
    for (int i = tid; i < na; i += 256) { s[0] += a[3 * (size_t)i]; s[1] += a[3 * (size_t)i + 1]; s[2] += a[3 * (size_t)i + 2]; }
    for (int i = tid; i < nb; i += 256) { s[3] += b[3 * (size_t)i]; s[4] += b[3 * (size_t)i + 1]; s[5] += b[3 * (size_t)i + 2]; }
    #pragma unroll
    for (int k = 0; k < 6; ++k) atomicAdd(&acc[k], s[k]);
    __syncthreads();
    if (tid == 0) {
        // np.mean(axis=(0,1)) / 255 per channel, then sklearn cosine_similarity: float64 throughout
        double u[3], v[3];
        for (int k = 0; k < 3; ++k) { u[k] = ((double)acc[k] / (double)na) / 255.0; v[k] = ((double)acc[3 + k] / (double)nb) / 255.0; }
        const double nu = sqrt(u[0] * u[0] + u[1] * u[1] + u[2] * u[2]), nv = sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
        double c = 0.0;
        if (nu > 0.0 && nv > 0.0) {
            const double un[3] = {u[0] / nu, u[1] / nu, u[2] / nu}, vn[3] = {v[0] / nv, v[1] / nv, v[2] / nv};
            c = un[0] * vn[0] + un[1] * vn[1] + un[2] * vn[2];
        }
        out[0] = c;
    }
}
}  // namespace ssad

extern "C" int ssad_polygon_mask(int32_t n, const int16_t* poly_xy, const uint8_t* poly_n, const int16_t* size,
                                 uint8_t* out, int32_t max_w, int32_t max_h, void* stream) {
    if (n < 0 || !poly_xy || !poly_n || !size || !out || max_w <= 0 || max_h <= 0) return SSAD_EINVAL;
    if (n == 0) return SSAD_OK;
    const int threads = 128;
    size_t smem = (size_t)threads * ((max_w + 31) / 32) * 4;
    if (smem > 48 * 1024) return SSAD_EUNSUPPORTED;
    ssad::polygon_mask_kernel<<<n, threads, smem, (cudaStream_t)stream>>>(n, poly_xy, poly_n, size, out, max_w, max_h);
    SSAD_CHECK_LAUNCH();
    return SSAD_OK;
}

extern "C" int ssad_affine_nearest(const uint8_t* src, int32_t src_w, int32_t src_h, int32_t channels,
                                   const int32_t* fix6_host, uint8_t* out, int32_t out_w, int32_t out_h, void* stream) {
    if (!src || !out || !fix6_host || src_w <= 0 || src_h <= 0 || out_w <= 0 || out_h <= 0) return SSAD_EINVAL;
    if (channels != 3 && channels != 4) return SSAD_EINVAL;
    ssad::Fix6 f;
    for (int i = 0; i < 6; ++i) f.a[i] = fix6_host[i];
    int n = out_w * out_h;
    ssad::affine_nearest_kernel<<<(n + 255) / 256, 256, 0, (cudaStream_t)stream>>>(src, src_w, src_h, channels, f, out, out_w, out_h);
    SSAD_CHECK_LAUNCH();
    return SSAD_OK;
}

extern "C" int ssad_color_jitter(int32_t n, int32_t height, int32_t width, const uint8_t* src, uint8_t* out,
                                 const uint8_t* order, const float* factors, const uint8_t* enable, void* stream) {
    if (n < 0 || height <= 0 || width <= 0 || !src || !out || !order || !factors || !enable) return SSAD_EINVAL;
    if (n == 0) return SSAD_OK;
    ssad::color_jitter_kernel<<<n, 256, 0, (cudaStream_t)stream>>>(height, width, src, out, order, factors, enable);
    SSAD_CHECK_LAUNCH();
    return SSAD_OK;
}

extern "C" int ssad_select_kth(const uint32_t* mask_bits, int32_t mask_pitch_words, int32_t win_x, int32_t win_y,
                               int32_t win_w, int32_t win_h, int32_t n, const uint32_t* k, int32_t* out_xy,
                               int32_t* count_out, void* stream) {
    if (!mask_bits || mask_pitch_words <= 0 || win_x < 0 || win_y < 0 || win_w <= 0 || win_h <= 0 || n < 0) return SSAD_EINVAL;
    if (n > 0 && (!k || !out_xy)) return SSAD_EINVAL;
    size_t smem = (size_t)(win_h + 1) * 4;
    if (smem > 48 * 1024) return SSAD_EUNSUPPORTED;
    ssad::select_kth_kernel<<<1, 256, smem, (cudaStream_t)stream>>>(mask_bits, mask_pitch_words, win_x, win_y, win_w, win_h,
                                                                      n, k, out_xy, count_out);
    SSAD_CHECK_LAUNCH();
    return SSAD_OK;
}

extern "C" int ssad_to_tensor(int32_t n, int32_t height, int32_t width, const uint8_t* src, int64_t src_pitch,
                              int64_t src_stride, float* out, void* stream) {
    if (n < 0 || height <= 0 || width <= 0 || !src || !out || src_pitch < 3ll * width) return SSAD_EINVAL;
    if (n == 0) return SSAD_OK;
    dim3 grid((height * width + 255) / 256, n);
    ssad::to_tensor_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(height, width, src, src_pitch, src_stride, out);
    SSAD_CHECK_LAUNCH();
    return SSAD_OK;
}

extern "C" int ssad_paste(uint8_t* image, int32_t width, int32_t height, const uint8_t* patch, int32_t pw, int32_t ph,
                          int32_t channels, const uint8_t* mask, int32_t left, int32_t top, void* stream) {
    if (!image || !patch || width <= 0 || height <= 0 || pw <= 0 || ph <= 0 || (channels != 3 && channels != 4)) return SSAD_EINVAL;
    const int n = pw * ph;
    ssad::paste_kernel<<<(n + 255) / 256, 256, 0, static_cast<cudaStream_t>(stream)>>>(image, width, height, patch, pw, ph, channels,
                                                                                   mask, left, top);
    SSAD_CHECK_LAUNCH();
    return SSAD_OK;
}

extern "C" int ssad_color_similarity(const uint8_t* image, int32_t n_image, const uint8_t* patch, int32_t n_patch, double* out,
                                     void* stream) {
    if (!image || !patch || !out || n_image <= 0 || n_patch <= 0) return SSAD_EINVAL;
    ssad::color_similarity_kernel<<<1, 256, 0, static_cast<cudaStream_t>(stream)>>>(image, n_image, patch, n_patch, out);
    SSAD_CHECK_LAUNCH();
    return SSAD_OK;
}

extern "C" int ssad_line_mask(int32_t n, const int16_t* xy, const uint8_t* npts, int32_t max_pts, int32_t line_width, uint8_t* out,
                              int32_t width, int32_t height, void* stream) {
    if (n < 0 || !xy || !npts || !out || max_pts < 2 || width <= 0 || height <= 0 || line_width < 0) return SSAD_EINVAL;
    if (n == 0) return SSAD_OK;
    const size_t smem = (size_t)height * ((width + 31) / 32) * 4;
    if (smem > 40 * 1024 || width > 1024) return SSAD_EUNSUPPORTED;
    ssad::line_mask_kernel<<<n, 256, smem, static_cast<cudaStream_t>(stream)>>>(xy, npts, max_pts, line_width, out, width, height);
    SSAD_CHECK_LAUNCH();
    return SSAD_OK;
}

extern "C" int ssad_upload_windows(int32_t n, const SsadImageParams* params, const uint8_t* host_src, uint8_t* dev_src, int64_t src_pitch,
                                   int64_t src_stride, int32_t src_height, int32_t src_width, int32_t win_height, int32_t win_width,
                                   int32_t max_patch_h, void* stream) {
    if (n < 0 || !params || !host_src || !dev_src || src_height <= 0 || src_width <= 0 || win_height <= 0 || win_width <= 0 ||
        max_patch_h < 0 || src_pitch < 3ll * src_width || src_stride < src_pitch * src_height) return SSAD_EINVAL;
    if (n == 0) return SSAD_OK;
    if ((reinterpret_cast<uintptr_t>(host_src) & 15) || (reinterpret_cast<uintptr_t>(dev_src) & 15) || (src_stride & 15)) return SSAD_EALIGN;
    const int rows = win_height + max_patch_h;
    static int sm_count[64];
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return ssad_set_cuda_error(e);
    if (dev < 0 || dev >= 64) return SSAD_EUNSUPPORTED;
    if (sm_count[dev] == 0) {
        int k = 0;
        e = cudaDeviceGetAttribute(&k, cudaDevAttrMultiProcessorCount, dev);
        if (e != cudaSuccess) return ssad_set_cuda_error(e);
        sm_count[dev] = k > 0 ? k : 1;
    }
    const long long ntask = (long long)n * rows;
    long long grid = (ntask + 7) / 8;
    if (grid > 2ll * sm_count[dev]) grid = 2ll * sm_count[dev];        // two CTAs per SM keep ~2 MB of reads in flight
    ssad::upload_windows_kernel<<<(unsigned)grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(n, params, host_src, dev_src, src_pitch,
                                                                                             src_stride, src_height, src_width,
                                                                                             win_height, win_width, rows);
    SSAD_CHECK_LAUNCH();
    return SSAD_OK;
}
