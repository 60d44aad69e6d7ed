// Standalone probe of the sm_100a primitives the cluster kernel relies on (run on the GPU box):
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o sm100_probe tools/sm100_probe.cu && ./sm100_probe
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include <cuda.h>
#include <cuda_runtime.h>
#include "../self-supervised-anomaly-detection_b200/csrc/ssad_sm100.cuh"
using namespace ssad;

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("  CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); return 1; } } while (0)

__global__ void k_f32x2(const float* a, float* o, int mode) {
    int i = threadIdx.x;
    f32x2 x = pack2(a[2 * i], a[2 * i + 1]), y = pack2(a[64 + 2 * i], a[64 + 2 * i + 1]), z;
    if (mode == 0) z = add2_rn(x, y);
    else if (mode == 1) z = add2_rz(x, y);
    else if (mode == 2) z = mul2_rn(x, y);
    else asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(z) : "l"(x), "l"(y));
    float lo, hi; unpack2(z, lo, hi);
    o[2 * i] = lo; o[2 * i + 1] = hi;
}

__global__ void __cluster_dims__(2, 1, 1) k_cluster(uint32_t* out) {
    __shared__ uint32_t slot[4];
    uint32_t rank = cluster_ctarank();
    if (threadIdx.x == 0) slot[0] = 100 + rank + 10 * cluster_id_x();
    cluster_arrive(); cluster_wait();
    if (threadIdx.x == 0) out[blockIdx.x] = dsmem_ld_u32(&slot[0], rank ^ 1u) + 1000 * cluster_nclusters_x();
    cluster_arrive(); cluster_wait();
}

__global__ void k_mbar_only(uint32_t* out) {
    __shared__ __align__(8) uint64_t bar;
    if (threadIdx.x == 0) { mbar_init(&bar, 1); mbar_fence_init(); }
    __syncthreads();
    if (threadIdx.x == 0) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(&bar)) : "memory"); }
    mbar_wait(&bar, 0);
    if (threadIdx.x == 0) out[0] = 7;
}

__global__ void __cluster_dims__(2, 1, 1) k_tma(const __grid_constant__ CUtensorMap tmap, uint8_t* out, int box_w, int box_h, int x, int y, int z, int step) {
    extern __shared__ __align__(128) uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((128u - (smem_u32(smem_raw) & 127u)) & 127u);
    __shared__ __align__(8) uint64_t bar;
    if (threadIdx.x == 0) {
        mbar_init(&bar, 1); mbar_fence_init();
        if (step >= 1) tma_prefetch_desc(&tmap);
    }
    __syncthreads();
    if (threadIdx.x == 0 && step >= 2) {
        fence_proxy_async();
        mbar_expect_tx(&bar, box_w * box_h);
        if (step >= 3) tma_load_3d(smem, &tmap, &bar, x, y + (int)cluster_ctarank() * box_h, z);
    }
    if (step >= 3) mbar_wait(&bar, 0);
    __syncthreads();
    for (int i = threadIdx.x; i < box_w * box_h; i += blockDim.x) out[(size_t)blockIdx.x * box_w * box_h + i] = (step >= 3) ? smem[i] : 1;
    cluster_arrive(); cluster_wait();
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int main() {
    float h[128], *da, *dо;
    for (int i = 0; i < 128; ++i) h[i] = 1.0f + 0.37f * i;
    CK(cudaMalloc(&da, sizeof(h))); CK(cudaMalloc(&dо, sizeof(h)));
    CK(cudaMemcpy(da, h, sizeof(h), cudaMemcpyHostToDevice));
    const char* names[4] = {"add.rn.f32x2", "add.rz.f32x2", "mul.rn.f32x2", "sub.rn.f32x2"};
    for (int m = 0; m < 4; ++m) {
        k_f32x2<<<1, 32>>>(da, dо, m);
        cudaError_t e = cudaDeviceSynchronize();
        printf("%s: %s\n", names[m], cudaGetErrorString(e));
        if (e != cudaSuccess) return 1;
    }
    uint32_t* du; CK(cudaMalloc(&du, 64));
    k_cluster<<<4, 64>>>(du);
    cudaError_t e = cudaDeviceSynchronize();
    uint32_t hu[4] = {0, 0, 0, 0};
    if (e == cudaSuccess) cudaMemcpy(hu, du, 16, cudaMemcpyDeviceToHost);
    printf("cluster barrier + dsmem: %s  [%u %u %u %u]\n", cudaGetErrorString(e), hu[0], hu[1], hu[2], hu[3]);
    if (e != cudaSuccess) return 1;
    k_mbar_only<<<1, 64>>>(du);
    e = cudaDeviceSynchronize();
    printf("mbarrier init/arrive/wait: %s\n", cudaGetErrorString(e));
    if (e != cudaSuccess) return 1;

    // TMA: [B=3][H=40][W*3=480] u8, box 192 x 8 at arbitrary x
    const int B = 3, H = 40, RB = 480;
    std::vector<uint8_t> img((size_t)B * H * RB);
    for (size_t i = 0; i < img.size(); ++i) img[i] = (uint8_t)((i * 7 + i / RB) & 0xFF);
    uint8_t *dsrc, *dout;
    CK(cudaMalloc(&dsrc, img.size())); CK(cudaMemcpy(dsrc, img.data(), img.size(), cudaMemcpyHostToDevice));
    const int bw = 192, bh = 8;
    CK(cudaMalloc(&dout, 2 * bw * bh));
    void* p = nullptr; cudaDriverEntryPointQueryResult q;
    CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q));
    if (!p) { printf("no cuTensorMapEncodeTiled\n"); return 1; }
    CUtensorMap tmap;
    cuuint64_t gdim[3] = {RB, H, B}; cuuint64_t gstr[2] = {RB, (cuuint64_t)H * RB};
    cuuint32_t box[3] = {bw, bh, 1}, es[3] = {1, 1, 1};
    CUresult cr = ((EncodeTiledFn)p)(&tmap, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, dsrc, gdim, gstr, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                     CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("encode: %d\n", (int)cr);
    if (cr != CUDA_SUCCESS) return 1;
    for (int step = 0; step <= 3; ++step) {
        const int x = 37, y = 5, z = 2;
        k_tma<<<2, 128, bw * bh + 128>>>(tmap, dout, bw, bh, x, y, z, step);
        e = cudaDeviceSynchronize();
        printf("tma step %d: %s\n", step, cudaGetErrorString(e));
        if (e != cudaSuccess) return 1;
        if (step == 3) {
            std::vector<uint8_t> o(2 * bw * bh);
            cudaMemcpy(o.data(), dout, o.size(), cudaMemcpyDeviceToHost);
            int bad = 0;
            for (int r = 0; r < 2; ++r) for (int yy = 0; yy < bh; ++yy) for (int xx = 0; xx < bw; ++xx)
                if (o[(size_t)r * bw * bh + yy * bw + xx] != img[((size_t)z * H + y + r * bh + yy) * RB + x + xx]) ++bad;
            printf("tma data mismatches: %d\n", bad);
        }
    }
    printf("probe done\n");
    return 0;
}
