// Probe 2: the producer / consumer ring of decode_mega.cu in isolation.  One producer thread streams 16 KB slots with 1-D bulk
// copies; NW consumer warps wait for every slot (full barrier), optionally touch it, and arrive on its empty barrier (count NW),
// which the producer waits for before it reuses the slot.  Variants: polling instruction, back-off, who polls.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/ring_probe tools/ring_probe.cu
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

constexpr uint32_t kSlot = 16384;
__device__ __forceinline__ uint32_t s_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
template <int POLL>  // 0 try_wait, 1 test_wait, 2 test_wait + nanosleep
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t done = 0;
    while (!done) {
        if (POLL == 0)
            asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                         : "=r"(done) : "r"(bar), "r"(parity) : "memory");
        else
            asm volatile("{\n\t.reg .pred p;\n\tmbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                         : "=r"(done) : "r"(bar), "r"(parity) : "memory");
        if (POLL == 2 && !done) __nanosleep(40);
    }
}
// WHO: 0 = every lane of every consumer warp polls; 1 = lane 0 polls, then __syncwarp; 2 = warp 0 polls, bar.sync releases the rest
template <int POLL, int WHO, int TOUCH>
__global__ void __launch_bounds__(544, 1) ring(const uint8_t* src, size_t bytes_per_cta, int depth, int nw, float* sink) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ __align__(8) uint64_t full[16], empty[16];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    if (tid == 0) {
        for (int i = 0; i < depth; ++i) {
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(s_u32(&full[i])));
            asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(s_u32(&empty[i])), "r"(WHO == 2 ? 1 : nw));
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    const uint32_t n = (uint32_t)(bytes_per_cta / kSlot);
    if (warp == 16) {
        if (lane != 0) return;
        const uint8_t* base = src + (size_t)blockIdx.x * bytes_per_cta;
        uint32_t idx = 0, par = 1;
        for (uint32_t i = 0; i < n; ++i) {
            mbar_wait<POLL>(s_u32(&empty[idx]), par);
            const uint32_t bar = s_u32(&full[idx]);
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(kSlot) : "memory");
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                         ::"r"(s_u32(smem) + idx * kSlot), "l"(base + (size_t)i * kSlot), "r"(kSlot), "r"(bar) : "memory");
            if (++idx == (uint32_t)depth) { idx = 0; par ^= 1; }
        }
        return;
    }
    if (warp >= nw) return;
    uint32_t idx = 0, par = 0;
    float acc = 0.f;
    for (uint32_t i = 0; i < n; ++i) {
        if (WHO == 0) mbar_wait<POLL>(s_u32(&full[idx]), par);
        else if (WHO == 1) { if (lane == 0) mbar_wait<POLL>(s_u32(&full[idx]), par); __syncwarp(); }
        else { if (warp == 0) mbar_wait<POLL>(s_u32(&full[idx]), par); asm volatile("bar.sync 1, %0;" ::"r"(nw * 32) : "memory"); }
        if (TOUCH) {
            const uint4 v = *reinterpret_cast<const uint4*>(smem + idx * kSlot + (size_t)tid * 16);
            acc += __uint_as_float(v.x) + __uint_as_float(v.y) + __uint_as_float(v.z) + __uint_as_float(v.w);
        }
        if (WHO == 2) {
            asm volatile("bar.sync 2, %0;" ::"r"(nw * 32) : "memory");
            if (tid == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(s_u32(&empty[idx])) : "memory");
        } else {
            __syncwarp();
            if (lane == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(s_u32(&empty[idx])) : "memory");
        }
        if (++idx == (uint32_t)depth) { idx = 0; par ^= 1; }
    }
    if (acc == 123.456f) sink[0] = acc;
}

template <int POLL, int WHO, int TOUCH>
static void run(const char* name, const uint8_t* buf, size_t per_cta, int sms, float* sink) {
    auto k = ring<POLL, WHO, TOUCH>;
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    for (int depth : {4, 11})
        for (int nw : {1, 16}) {
            float ms = 0;
            for (int rep = 0; rep < 2; ++rep) {
                cudaEventRecord(e0);
                k<<<sms, 544, depth * kSlot>>>(buf, per_cta, depth, nw, sink);
                cudaEventRecord(e1);
                if (cudaDeviceSynchronize() != cudaSuccess) { printf("%s failed: %s\n", name, cudaGetErrorString(cudaGetLastError())); return; }
                cudaEventElapsedTime(&ms, e0, e1);
            }
            printf("%-44s depth %2d consumers %2d warps: %.3f ms  %5.0f GB/s\n", name, depth, nw, ms, per_cta * sms / ms * 1e-6);
        }
}

int main() {
    const size_t per_cta = 8u << 20;
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    uint8_t* buf;
    float* sink;
    if (cudaMalloc(&buf, per_cta * sms) != cudaSuccess) return 1;
    cudaMalloc(&sink, 16);
    cudaMemset(buf, 0, per_cta * sms);
    run<0, 0, 0>("try_wait, all lanes poll", buf, per_cta, sms, sink);
    run<1, 0, 0>("test_wait, all lanes poll", buf, per_cta, sms, sink);
    run<2, 0, 0>("test_wait + nanosleep, all lanes poll", buf, per_cta, sms, sink);
    run<0, 1, 0>("try_wait, lane 0 polls", buf, per_cta, sms, sink);
    run<0, 2, 0>("try_wait, warp 0 polls + bar.sync", buf, per_cta, sms, sink);
    run<0, 0, 1>("try_wait, all lanes poll, touch data", buf, per_cta, sms, sink);
    run<0, 1, 1>("try_wait, lane 0 polls, touch data", buf, per_cta, sms, sink);
    return 0;
}
