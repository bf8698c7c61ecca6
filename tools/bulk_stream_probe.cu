// Probe: how fast can ONE thread per SM stream global memory into a shared-memory ring, (a) with 1-D cp.async.bulk copies,
// (b) with 2-D tensor-map TMA loads of the same 16 KB?  No consumer: the thread itself waits for slot i - depth to land before
// it requests slot i.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/bulk_stream_probe tools/bulk_stream_probe.cu -lcuda
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

constexpr uint32_t kSlot = 16384;

__device__ __forceinline__ uint32_t s_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t done = 0;
    while (!done)
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(done) : "r"(bar), "r"(parity) : "memory");
}

template <int MODE>   // 0: 1-D bulk, 1: 2-D tensor map (box 256 x 32 halves)
__global__ void __launch_bounds__(64, 1) probe(const __grid_constant__ CUtensorMap map, const uint8_t* src, size_t bytes_per_cta, int depth,
                                               int split) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ __align__(8) uint64_t full[16];
    if (threadIdx.x == 0) {
        for (int i = 0; i < depth; ++i) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(s_u32(&full[i])));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x != 0) return;
    const uint32_t n = (uint32_t)(bytes_per_cta / kSlot);
    const uint8_t* base = src + (size_t)blockIdx.x * bytes_per_cta;
    for (uint32_t i = 0; i < n + depth; ++i) {
        if (i >= (uint32_t)depth) mbar_wait(s_u32(&full[(i - depth) % depth]), ((i - depth) / depth) & 1u);
        if (i < n) {
            const uint32_t idx = i % depth, bar = s_u32(&full[idx]), dst = s_u32(smem) + idx * kSlot;
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(kSlot) : "memory");
            if (MODE == 0) {
                const uint32_t part = kSlot / split;
                for (int s = 0; s < split; ++s)
                    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                                 ::"r"(dst + s * part), "l"(base + (size_t)i * kSlot + s * part), "r"(part), "r"(bar) : "memory");
            } else {
                const int row = (int)(((size_t)blockIdx.x * bytes_per_cta + (size_t)i * kSlot) / 512);
                asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                             ::"r"(dst), "l"(&map), "r"(bar), "r"(0), "r"(row) : "memory");
            }
        }
    }
}

int main(int argc, char** argv) {
    const size_t per_cta = 8u << 20;   // 8 MB per CTA: 1.2 GB in all, nothing from L2
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    const size_t total = per_cta * sms;
    uint8_t* buf;
    if (cudaMalloc(&buf, total) != cudaSuccess) { printf("alloc failed\n"); return 1; }
    cudaMemset(buf, 1, total);
    CUtensorMap map;
    cuInit(0);
    cuuint64_t dims[2] = {256, total / 512};
    cuuint64_t strides[1] = {512};
    cuuint32_t box[2] = {256, 32}, estr[2] = {1, 1};
    CUresult r = cuTensorMapEncodeTiled(&map, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, buf, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                        CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { printf("tensor map failed %d\n", (int)r); return 1; }
    cudaFuncSetAttribute(probe<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    cudaFuncSetAttribute(probe<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    const int depths[] = {2, 4, 8, 11};
    for (int mode = 0; mode < 2; ++mode)
        for (int split : {1, 4})
            for (int depth : depths) {
                if (mode == 1 && split != 1) continue;
                for (int rep = 0; rep < 2; ++rep) {
                    cudaEventRecord(e0);
                    if (mode == 0) probe<0><<<sms, 64, depth * kSlot>>>(map, buf, per_cta, depth, split);
                    else probe<1><<<sms, 64, depth * kSlot>>>(map, buf, per_cta, depth, split);
                    cudaEventRecord(e1);
                    cudaError_t err = cudaDeviceSynchronize();
                    if (err != cudaSuccess) { printf("kernel failed: %s\n", cudaGetErrorString(err)); return 1; }
                    float ms;
                    cudaEventElapsedTime(&ms, e0, e1);
                    if (rep == 1)
                        printf("%s split %d depth %2d (%3d KB in flight per SM): %.3f ms  %.0f GB/s\n", mode == 0 ? "bulk-1d  " : "tensor-2d", split, depth,
                               depth * 16, ms, total / ms * 1e-6);
                }
            }
    return 0;
}
