// The hop between shards: replaces Communicator.transfer_data / receive_data
// (utils/node_worker.py:44-67: torch.save -> file -> ZMQ PUSH | PULL -> file -> torch.load) by
// 128-bit stores from a kernel straight into the next GPU's mailbox over NVLink, followed by a
// system-scope release of a sequence flag; the receiver acquires the flag on its own GPU.
#include "common.cuh"
#include "internal.h"

namespace b200 {

__device__ __forceinline__ void st_release_sys(uint64_t* p, uint64_t v) {
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ uint64_t ld_acquire_sys(const uint64_t* p) {
    uint64_t v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}

// grid of up to 64 CTAs; the last CTA to finish publishes the flag (counter lives on the sender)
__global__ void __launch_bounds__(1024) hop_send_kernel(const uint4* __restrict__ src, int64_t n16, uint4* peer_slot,
                                                        uint64_t* peer_flag, uint64_t seq, unsigned int* counter) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += (int64_t)gridDim.x * blockDim.x)
        peer_slot[i] = src[i];
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
        bool last = true;
        if (gridDim.x > 1) {  // a single CTA (payloads up to 64 KiB, flag-only sends) needs no arrival counter
            const unsigned int done = atomicAdd(counter, 1u) + 1u;
            last = (done == gridDim.x);
            if (last) {
                *counter = 0u;
                __threadfence_system();
            }
        }
        if (last) st_release_sys(peer_flag, seq + 1);  // cumulative: ordered after every store fenced before the barrier
    }
}

__global__ void __launch_bounds__(1024) hop_recv_kernel(uint4* __restrict__ dst, int64_t n16, const uint4* local_slot,
                                                        const uint64_t* local_flag, uint64_t seq, long long max_wait_cycles,
                                                        int* err) {
    __shared__ int ok;
    if (threadIdx.x == 0) {
        const long long t0 = clock64();
        int good = 1;
        while (ld_acquire_sys(local_flag) < seq + 1) {
            if (clock64() - t0 > max_wait_cycles) { good = 0; break; }
            __nanosleep(20);
        }
        ok = good;
        if (!good) *err = 1;
    }
    __syncthreads();
    if (!ok) return;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += (int64_t)gridDim.x * blockDim.x)
        dst[i] = local_slot[i];
}

// Message = 256-byte header (built on the host, passed by value as a kernel parameter: no staging copy) + payload.
// The header lands at the start of the peer's slot, the payload right behind it, then the flag is released.
__global__ void __launch_bounds__(1024) hop_send_msg_kernel(const HopHeader hdr, const uint4* __restrict__ src, int64_t n16,
                                                            uint4* peer_slot, uint64_t* peer_flag, uint64_t seq,
                                                            unsigned int* counter) {
    if (blockIdx.x == 0 && threadIdx.x < kHopHeaderBytes / 16) peer_slot[threadIdx.x] = reinterpret_cast<const uint4*>(hdr.b)[threadIdx.x];
    uint4* dst = peer_slot + kHopHeaderBytes / 16;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += (int64_t)gridDim.x * blockDim.x) dst[i] = src[i];
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
        bool last = true;
        if (gridDim.x > 1) {
            const unsigned int done = atomicAdd(counter, 1u) + 1u;
            last = (done == gridDim.x);
            if (last) {
                *counter = 0u;
                __threadfence_system();
            }
        }
        if (last) st_release_sys(peer_flag, seq + 1);
    }
}

static int hop_grid(int64_t n16) {
    int64_t g = (n16 + 1024 * 4 - 1) / (1024 * 4);
    if (g < 1) g = 1;
    if (g > 64) g = 64;
    return (int)g;
}

void launch_hop_send(const void* payload, int64_t bytes, void* peer_slot, uint64_t* peer_flag, uint64_t seq,
                     unsigned int* counter, cudaStream_t st) {
    const int64_t n16 = (bytes + 15) / 16;
    hop_send_kernel<<<hop_grid(payload ? n16 : 0), payload ? 1024 : 32, 0, st>>>(reinterpret_cast<const uint4*>(payload), payload ? n16 : 0,
                                                    reinterpret_cast<uint4*>(peer_slot), peer_flag, seq, counter);
}
void launch_hop_send_msg(const HopHeader& hdr, const void* payload, int64_t bytes, void* peer_slot, uint64_t* peer_flag,
                         uint64_t seq, unsigned int* counter, cudaStream_t st) {
    const int64_t n16 = (bytes + 15) / 16;
    hop_send_msg_kernel<<<hop_grid(n16), n16 > 0 ? 1024 : 32, 0, st>>>(hdr, reinterpret_cast<const uint4*>(payload), n16,
                                                                       reinterpret_cast<uint4*>(peer_slot), peer_flag, seq, counter);
}
void launch_hop_recv(void* out, int64_t bytes, const void* local_slot, const uint64_t* local_flag, uint64_t seq,
                     int64_t max_wait_cycles, int* err, cudaStream_t st) {
    const int64_t n16 = (bytes + 15) / 16;
    hop_recv_kernel<<<hop_grid(n16), n16 > 0 ? 1024 : 32, 0, st>>>(reinterpret_cast<uint4*>(out), n16,
                                                    reinterpret_cast<const uint4*>(local_slot), local_flag, seq,
                                                    (long long)max_wait_cycles, err);
}

}  // namespace b200
