// EXPERIMENT, off by default (B200SHARD_MEGA=1 turns it on; parity-green, slower than the per-kernel path — numbers below).
//
// One persistent kernel for all decoder layers of a single-token, single-row decode step (the BASELINE configs[1] hot loop:
// NodeWorker.pass_through_shard for one new token, utils/node_worker.py:252-258 -> LlamaShardPart.forward,
// utils/shard_loader.py:71-75 -> LlamaDecoderLayer, modeling_llama.py:299-332).
//
// Idea: launched as five kernels per layer, every kernel boundary drains the memory pipeline (wait for the predecessor,
// stage the activations, first DRAM round trip, ragged tail: about 2.5 us each, and the latency-bound attention kernel
// streams almost nothing for 6 us) — the step sits at 0.79-0.84 of HBM bandwidth although each GEMV alone reaches 0.90.
// Here one CTA per SM stays resident for the whole call and the weight stream does not stop at phase boundaries:
//   * warp 16 of every CTA is a producer: one thread walks this CTA's share of every projection of every layer in order
//     and copies it, 16 KB at a time, into a shared-memory ring (cp.async.bulk global -> shared, completion on an mbarrier).
//     It depends on nothing but free ring slots, so it runs ahead across phase and layer boundaries: while the consumers
//     sit in a grid barrier, stage activations or walk the attention chain, the ring (176 KB per SM, 26 MB per GPU = 4 us
//     of HBM time) keeps filling with the next projection's rows.
//   * warps 0..15 are consumers: together they take the oldest slot (thread t owns the 16-byte chunks t, t + 512, ... of
//     its rows), form the dot products against the activation vector staged in shared memory (fp32), and free the slot.
//   * phases of a layer — [RMSNorm + QKV] | attention (RoPE, paged K/V append, split-K softmax, merge) | [o_proj +
//     residual] | [RMSNorm + gate/up] | [SwiGLU + down_proj + residual] — are separated by a grid barrier (one release
//     reduction per CTA on a 64-bit counter that only grows, bounded spin); everything that crosses CTAs goes through L2.
// Rows are dealt to CTAs in contiguous ranges (a CTA's share of a projection is one contiguous byte range of the weight
// matrix), so RoPE pairs and gate/up pairs may straddle CTAs: the projections store raw fp32 results and the pairing is
// applied by the consumer of the NEXT phase while it stages its activations.
// Launched cooperatively (all CTAs co-resident or the launch fails), grid = number of SMs.
//
// Measured (B200, Llama-2-7B layer, 512-token context, tools/batch_decode_bench.py; profiles/decode_mega_r2.md):
//   per-kernel path 75.6-76.1 us per layer; this kernel 152.9 us; with the FMA loop removed 121.9; with the reductions removed
//   as well (stream + handshake + barriers + staging + attention only) 93.4.  In isolation the same ring streams at
//   6.9-7.0 TB/s with one producer thread and 64 KB in flight per SM (tools/bulk_stream_probe.cu, tools/ring_probe.cu), and
//   the grid barrier costs 1.2-1.9 us.  What loses: (1) the consumers are instruction-bound — with one or two 16-byte chunks
//   per thread per slot the wait / bookkeeping / shuffle-reduce / release around them is three times the FMA work, and a
//   single resident CTA has 16 warps to hide it where the per-kernel GEMV has 32; (2) barrier + staging + the attention
//   chain (27 us per layer) exceed what the ring can buffer (4 us per boundary), so most of it stays exposed; (3) with the
//   shared-memory carve-out at its maximum the L1 is tiny: any local-memory traffic (a spilled array, a struct passed by
//   reference) costs an L2 round trip — the first versions lost 35 us per layer to that alone; (4) integer divisions in the
//   single producer thread (four per slot) capped the stream at 3.5 TB/s until they were replaced by running counters.
#include <stdio.h>
#include <stdlib.h>

#include "common.cuh"
#include "internal.h"
#include "tc_common.cuh"

#ifndef MEGA_PROFILE
#define MEGA_PROFILE 0
#endif

namespace b200 {
using namespace tc;

constexpr int kMgWarps = 16;                       // consumer warps
constexpr int kMgConsumers = kMgWarps * 32;        // 512
constexpr int kMgThreads = kMgConsumers + 32;      // + the producer warp
constexpr uint32_t kMgSlotBytes = 16384;
constexpr int kMgMaxSlots = 12;
constexpr int kMgAttnFloats = 32 * kHeadDim + 64 + 3 * kHeadDim + 8;   // half-warp partials + their (m, l) + q / new k / new v

struct MegaGeom { int nr, nseg, seg_elems; };
// how a projection with rows of K 16-bit elements maps to ring slots: nr whole rows per slot, or nseg segments per row
__host__ __device__ __forceinline__ MegaGeom mega_geom(int K) {
    MegaGeom g;
    const uint32_t row_bytes = 2u * (uint32_t)K;
    if (row_bytes <= kMgSlotBytes) {
        const int fit = (int)(kMgSlotBytes / row_bytes);
        g.nr = fit < 4 ? fit : 4;
        g.nseg = 1;
        g.seg_elems = K;
    } else {
        g.nr = 1;
        g.nseg = (int)((row_bytes + kMgSlotBytes - 1) / kMgSlotBytes);
        g.seg_elems = (((K + g.nseg - 1) / g.nseg) + 255) & ~255;   // whole 256-element blocks of the staged vector
    }
    return g;
}

__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ void consumer_sync() { asm volatile("bar.sync 1, %0;" ::"n"(kMgConsumers) : "memory"); }
__device__ __forceinline__ uint4 lds128(uint32_t addr) {
    uint4 r;
    asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "r"(addr));
    return r;
}
__device__ __forceinline__ float2 h2f2(uint32_t v) { return __half22float2(*reinterpret_cast<const __half2*>(&v)); }
__device__ __forceinline__ float dot8h(const uint4& w, const float4& xl, const float4& xh, float a) {
    const float2 w0 = h2f2(w.x), w1 = h2f2(w.y), w2 = h2f2(w.z), w3 = h2f2(w.w);
    a = fmaf(w0.x, xl.x, a); a = fmaf(w0.y, xl.y, a);
    a = fmaf(w1.x, xl.z, a); a = fmaf(w1.y, xl.w, a);
    a = fmaf(w2.x, xh.x, a); a = fmaf(w2.y, xh.y, a);
    a = fmaf(w3.x, xh.z, a); a = fmaf(w3.y, xh.w, a);
    return a;
}
__device__ __forceinline__ float dot8bf(const uint4& k, const float (&q)[8]) {
    float d = bf_lo(k.x) * q[0];
    d = fmaf(bf_hi(k.x), q[1], d);
    d = fmaf(bf_lo(k.y), q[2], d); d = fmaf(bf_hi(k.y), q[3], d);
    d = fmaf(bf_lo(k.z), q[4], d); d = fmaf(bf_hi(k.z), q[5], d);
    d = fmaf(bf_lo(k.w), q[6], d); d = fmaf(bf_hi(k.w), q[7], d);
    return d;
}

// The staged activation vector is kept in blocks of 256 elements = 32 chunks of 8: the first float4 of every chunk, then the
// second float4 of every chunk, so that the 32 lanes of a warp (one chunk each) read 512 contiguous bytes twice.
__device__ __forceinline__ int xs_f4(int i4) {   // natural float4 index -> index in that layout
    const int c = i4 >> 1;
    return ((c >> 5) << 6) + ((i4 & 1) << 5) + (c & 31);
}

// A wait that runs out (a protocol bug, a faulted copy) must never hang the GPU: the first one to give up records where
// (bar[2] = code, bar[3..5] = CTA, two details) and from then on every wait in every CTA returns at once, so the kernel ends;
// the host finds the code after the call (results are garbage then).
__device__ __forceinline__ bool mega_aborted(const unsigned int* bar) { return *reinterpret_cast<const volatile unsigned int*>(bar + 2) != 0u; }
__device__ __noinline__ void mega_fail(unsigned int* bar, unsigned int code, unsigned int a, unsigned int b) {
    if (atomicCAS(bar + 2, 0u, code) == 0u) {
        bar[3] = blockIdx.x; bar[4] = a; bar[5] = b;
        printf("decode_mega: wait %u timed out in CTA %u (warp %u): %u %u\n", code, blockIdx.x, threadIdx.x >> 5, a, b);
    }
}
__device__ __forceinline__ void mega_mbar_wait(uint32_t bar, uint32_t parity, unsigned int* gbar, unsigned int code, unsigned int info) {
    uint32_t done = 0;
    for (uint32_t spin = 0; !done; ++spin) {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done) : "r"(bar), "r"(parity) : "memory");
        if (!done && (spin & 1023u) == 1023u) {
            if (mega_aborted(gbar)) return;
            if (spin > (1u << 22)) { mega_fail(gbar, code, info, parity); return; }
        }
    }
}

enum MegaEpi { MG_RAW = 0, MG_RESID = 1 };

struct MegaShared {
    uint32_t ring, full0, empty0;   // shared-memory addresses
    int n_slots, mode;
    unsigned int* gbar;
};

// One projection: this CTA's rows of W [N,K] against the staged vector xs; scale multiplies every result (1 / rms).
// ALL consumer warps work on the oldest slot of the ring together (thread t owns the 16-byte chunks t, t + 512, ... of every
// row in it), so slots are consumed strictly in the order they were requested and live in shared memory only for as long as
// their bytes are in flight — the whole ring is bytes in flight, as the registers of the per-kernel GEMV were (and every
// thread waits for every slot in order, so a parity wait can never be taken for an earlier use of the barrier).
template <int EPI>
__device__ __forceinline__ void mega_gemv(const MegaShared& sh, uint32_t& cidx, uint32_t& cpar, const float* xs, float* part, int& buf, int N,
                                       int K, float scale, float* out, int tid, int warp, int lane) {
    const MegaGeom gm = mega_geom(K);
    const int rb = (int)((int64_t)N * blockIdx.x / gridDim.x), re = (int)((int64_t)N * (blockIdx.x + 1) / gridDim.x);
    const int n_units = (re - rb + gm.nr - 1) / gm.nr;               // producer's units: nr rows (or one long row) each
    // A reduction group is up to 16 rows.  Inside a group the warps run free: every warp reduces its own share of a slot's
    // rows with shuffles and adds it to its line of a 16 x 16 array; the named barrier, the sum over the warps and the
    // epilogue come once per group (128 KB of weights), so the warps' waits and FMA chains overlap one another.
    const int per = gm.nseg == 1 ? 16 / gm.nr : (8 / gm.nseg > 0 ? 8 / gm.nseg : 1);
    const int rows_per_group = per * gm.nr;
    const float4* x4 = reinterpret_cast<const float4*>(xs);
    for (int u0 = 0; u0 < n_units; u0 += per) {
        float* pw = part + buf * (kMgWarps * 16) + warp * 16;
        if (lane < 16) pw[lane] = 0.f;
        // the finishing threads fetch their residual value now; it has arrived by the time the group is summed
        const int my_row = rb + u0 * gm.nr + tid;
        float resid = 0.f;
        if (EPI == MG_RESID && tid < rows_per_group && my_row < re) resid = __ldcg(out + my_row);
        __syncwarp();
        for (int u = 0; u < per && u0 + u < n_units; ++u) {
            const int row0 = rb + (u0 + u) * gm.nr;
            const int rows = (re - row0) < gm.nr ? (re - row0) : gm.nr;
            float acc[4] = {0.f, 0.f, 0.f, 0.f};
            for (int k0 = 0; k0 < K; k0 += gm.seg_elems) {
                // ring position and phase bit are carried along (no divisions in this loop or in the producer's: one
                // thread that divides four times per 16 KB slot cannot request 45 GB/s)
                const uint32_t idx = cidx;
                mega_mbar_wait(sh.full0 + 8 * idx, cpar, sh.gbar, 1u, idx);
                if (++cidx == (uint32_t)sh.n_slots) { cidx = 0; cpar ^= 1u; }
                const int len = (K - k0) < gm.seg_elems ? (K - k0) : gm.seg_elems;
                const uint32_t base = sh.ring + idx * kMgSlotBytes;
                const uint32_t pitch = (uint32_t)len * 2u;
                const float4* xb = x4 + (k0 >> 2);
                for (int c = tid; c < (len >> 3) && sh.mode < 2; c += kMgConsumers) {
                    const int xi = ((c >> 5) << 6) + (c & 31);
                    const float4 xl = xb[xi], xh = xb[xi + 32];
                    const uint32_t a = base + (uint32_t)c * 16u;
                    acc[0] = dot8h(lds128(a), xl, xh, acc[0]);
                    if (rows > 1) acc[1] = dot8h(lds128(a + pitch), xl, xh, acc[1]);
                    if (rows > 2) acc[2] = dot8h(lds128(a + 2 * pitch), xl, xh, acc[2]);
                    if (rows > 3) acc[3] = dot8h(lds128(a + 3 * pitch), xl, xh, acc[3]);
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(sh.empty0 + 8 * idx);   // 16 arrivals (one per warp) free the slot
            }
            if (sh.mode < 3) {
#pragma unroll
                for (int r = 0; r < 4; ++r) {
                    if (r < rows) {
                        const float v = warp_sum(acc[r]);
                        if (lane == 0) pw[u * gm.nr + r] = v;
                    }
                }
            }
        }
        consumer_sync();
        if (tid < rows_per_group && my_row < re) {
            const float* pb = part + buf * (kMgWarps * 16);
            float v = 0.f;
#pragma unroll
            for (int w = 0; w < kMgWarps; ++w) v += pb[w * 16 + tid];
            v *= scale;
            out[my_row] = (EPI == MG_RESID) ? resid + v : v;
        }
        buf ^= 1;
    }
}

// sum over the 512 consumer threads, same value in every thread (fixed order)
__device__ __forceinline__ float consumer_sum(float v, float* red, int warp, int lane) {
    v = warp_sum(v);
    if (lane == 0) red[warp] = v;
    consumer_sync();
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < kMgWarps; ++w) t += red[w];
    return t;
}

// One arrival per CTA on a 64-bit counter that only ever grows (gridDim.x per barrier: no reset, no second word, and the
// arrival is a fire-and-forget reduction: one L2 round trip per barrier instead of two).
__device__ __forceinline__ unsigned long long gtime() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
__device__ __forceinline__ void grid_barrier(unsigned int* bar, unsigned long long target, int tid, unsigned long long* tm = nullptr) {
    consumer_sync();
    if (tid == 0) {
        if (tm) tm[0] = gtime();
        unsigned long long* ctr = reinterpret_cast<unsigned long long*>(bar + 6);
        asm volatile("red.release.gpu.global.add.u64 [%0], 1;" ::"l"(ctr) : "memory");
        if (tm) tm[1] = gtime();
        unsigned long long g;
        uint32_t spin = 0;
        do {
            asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(g) : "l"(ctr) : "memory");
            if ((++spin & 255u) == 0u) {
                if (mega_aborted(bar)) break;
                if (spin > (1u << 21)) { mega_fail(bar, 3u, (unsigned int)target, (unsigned int)g); break; }
            }
        } while (g < target);
        if (tm) { tm[2] = gtime(); tm[3] = spin; }
    }
    consumer_sync();
}

__global__ void __launch_bounds__(kMgThreads, 1) decode_mega_kernel(const MegaParams p) {
    extern __shared__ __align__(1024) uint8_t mg_smem[];
    __shared__ float red_s[kMgWarps];
    __shared__ int s_last;
    __shared__ unsigned long long s_gen0;
    __shared__ float part_s[2 * kMgWarps * 16];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    MegaShared sh;
    sh.n_slots = p.n_slots;
    sh.gbar = p.bar;
    sh.mode = p.profile;
    sh.ring = s_u32(mg_smem);
    float* xs = reinterpret_cast<float*>(mg_smem + (size_t)p.n_slots * kMgSlotBytes);
    const uint32_t bars = sh.ring + (uint32_t)p.n_slots * kMgSlotBytes + (uint32_t)p.xs_floats * 4u;
    sh.full0 = bars;
    sh.empty0 = bars + 8u * kMgMaxSlots;
    if (tid == 0) {
        for (int i = 0; i < p.n_slots; ++i) { mbar_init(sh.full0 + 8 * i, 1); mbar_init(sh.empty0 + 8 * i, kMgWarps); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        // the counter stood at a multiple of the grid size when the launch began; CTAs that are already waiting at the first
        // barrier have added less than one grid's worth
        const unsigned long long v = *reinterpret_cast<volatile unsigned long long*>(p.bar + 6);
        s_gen0 = v - v % gridDim.x;

    }
    __syncthreads();
    const int H = p.H, I = p.I, nq = p.n_heads * kHeadDim;

    if (warp == kMgWarps) {  // ===== producer: this CTA's share of every projection, in consumption order
        if (lane == 0) {
            uint32_t idx = 0, par = 1;   // ring position; parity that says "the previous use of this slot was consumed"
            for (int l = 0; l < p.n_layers; ++l) {
                const MegaLayer L = p.layers[l];
                for (int ph = 0; ph < 4; ++ph) {
                    const __half* W = ph == 0 ? L.wqkv : ph == 1 ? L.wo : ph == 2 ? L.wgu : L.wdown;
                    const int N = ph == 0 ? p.qkv_rows : ph == 2 ? 2 * I : H;
                    const int K = ph == 1 ? nq : ph == 3 ? I : H;
                    const MegaGeom gm = mega_geom(K);
                    const int rb = (int)((int64_t)N * blockIdx.x / gridDim.x), re = (int)((int64_t)N * (blockIdx.x + 1) / gridDim.x);
                    const __half* src = W + (size_t)rb * K;
                    if (gm.nseg == 1) {
                        const uint32_t full_bytes = (uint32_t)gm.nr * 2u * (uint32_t)K;
                        for (int left = re - rb; left > 0; left -= gm.nr) {
                            const uint32_t bytes = left >= gm.nr ? full_bytes : (uint32_t)left * 2u * (uint32_t)K;
                            mega_mbar_wait(sh.empty0 + 8 * idx, par, p.bar, 2u, idx);
                            mbar_expect_tx(sh.full0 + 8 * idx, bytes);
                            bulk_g2s(sh.ring + idx * kMgSlotBytes, src, bytes, sh.full0 + 8 * idx);
                            src += (size_t)gm.nr * K;
                            if (++idx == (uint32_t)p.n_slots) { idx = 0; par ^= 1u; }
                        }
                    } else {
                        for (int row = rb; row < re; ++row, src += K) {
                            for (int k0 = 0; k0 < K; k0 += gm.seg_elems) {
                                const uint32_t bytes = 2u * (uint32_t)((K - k0) < gm.seg_elems ? (K - k0) : gm.seg_elems);
                                mega_mbar_wait(sh.empty0 + 8 * idx, par, p.bar, 2u, idx);
                                mbar_expect_tx(sh.full0 + 8 * idx, bytes);
                                bulk_g2s(sh.ring + idx * kMgSlotBytes, src + k0, bytes, sh.full0 + 8 * idx);
                                if (++idx == (uint32_t)p.n_slots) { idx = 0; par ^= 1u; }
                            }
                        }
                    }
                }
            }
        }
        return;
    }

    // ===== consumers
    unsigned long long gen = s_gen0;
    uint32_t cidx = 0, cpar = 0;
    int pbuf = 0;
    const int T = p.t_end[0];                       // keys the new token attends to (itself included)
    const int group = p.n_heads / p.n_kv;
    const int32_t* bt = p.block_table + (size_t)p.slot * p.max_pages_per_slot;
    const int64_t page_stride = (int64_t)p.n_kv * p.head_stride;
    const int l16 = lane & 15, grp = warp * 2 + (lane >> 4);   // 32 half-warps, 16 lanes x 8 dims each
    const int n_items = p.n_heads * p.n_splits;
    const int chunk = (T + p.n_splits - 1) / p.n_splits;
    float* g_acc = xs;                              // attention scratch aliases the activation vector
    float* g_m = xs + 32 * kHeadDim;
    float* g_l = g_m + 32;
    float* q_s = g_l + 32;
    float* knew_s = q_s + kHeadDim;
    float* vnew_s = knew_s + kHeadDim;
    constexpr int U = 4;                            // keys in flight per half-warp
    const float scale_log2 = 0.08838834764831845f * 1.4426950408889634f;  // 128^-0.5 * log2(e)

#if MEGA_PROFILE
    // timestamps of layer 1 (local arrays: they live in local memory and slow the kernel down; measurement builds only)
    unsigned long long tmark[16], tmb[4] = {0, 0, 0, 0};
    int nmark = 0;
    auto mark = [&](int l) {
        if (p.profile == 1 && l == 1 && tid == 0 && nmark < 16) {
            unsigned long long t;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
            tmark[nmark++] = t;
        }
    };
    unsigned long long* tmb_p = tmb;
#else
    auto mark = [](int) {};
    unsigned long long* tmb_p = nullptr;
#endif
    for (int l = 0; l < p.n_layers; ++l) {
        const MegaLayer L = p.layers[l];
        mark(l);
        __nv_bfloat16* kp = p.k_pool + (int64_t)l * p.layer_stride;
        __nv_bfloat16* vp = p.v_pool + (int64_t)l * p.layer_stride;
        // ---------------------------------------------------------------- RMSNorm (modeling_llama.py:62-67) + q/k/v projections
        {
            float ss = 0.f;
            for (int i = tid; i < H / 4; i += kMgConsumers) {
                const float4 v = __ldcg(reinterpret_cast<const float4*>(p.x) + i);
                const uint2 w = *reinterpret_cast<const uint2*>(L.norm1 + 4 * i);
                ss += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
                reinterpret_cast<float4*>(xs)[xs_f4(i)] = make_float4(v.x * bf_lo(w.x), v.y * bf_hi(w.x), v.z * bf_lo(w.y), v.w * bf_hi(w.y));
            }
            ss = consumer_sum(ss, red_s, warp, lane);
            mark(l);
            mega_gemv<MG_RAW>(sh, cidx, cpar, xs, part_s, pbuf, p.qkv_rows, H, rsqrtf(ss / (float)H + p.eps), p.qkv_raw, tid, warp, lane);
            mark(l);
        }
        // first window of cached keys of this CTA's first attention item: requested before the barrier, landed after it
        uint4 kk[U], vv[U];
        {
            const int item = blockIdx.x;
            if (item < n_items) {
                const int h = item / p.n_splits, split = item % p.n_splits;
                const int t0 = split * chunk, t1 = min(T, t0 + chunk);
                const int t_pool = min(t1, T - 1);
                const int64_t lane_base = (int64_t)(h / group) * p.head_stride + l16 * 8;
#pragma unroll
                for (int u = 0; u < U; ++u) {
                    const int t = t0 + grp + 32 * u;
                    if (t < t_pool) {
                        const int64_t off = (int64_t)bt[t / p.page_size] * page_stride + lane_base + (int64_t)(t % p.page_size) * kHeadDim;
                        kk[u] = *reinterpret_cast<const uint4*>(kp + off);
                        vv[u] = *reinterpret_cast<const uint4*>(vp + off);
                    } else {
                        kk[u] = make_uint4(0, 0, 0, 0);
                        vv[u] = make_uint4(0, 0, 0, 0);
                    }
                }
            }
        }
        grid_barrier(p.bar, gen += gridDim.x, tid, (p.profile == 1 && l == 1) ? tmb_p : nullptr);
        mark(l);
        // ---------------------------------------------------------------- attention: RoPE (modeling_llama.py:146-168), K/V append
        // (cache_utils.py:102-121 as a paged write), softmax(q K^T / sqrt(d)) V over every cached key (:199-221), split over CTAs
        for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
            const int h = item / p.n_splits, split = item % p.n_splits, kvh = h / group;
            const int t0 = split * chunk, t1 = min(T, t0 + chunk);
            const bool has_new = (t1 == T) && (t0 < T);      // this split owns the new token's own key
            const int t_pool = has_new ? T - 1 : t1;         // keys [t0, t_pool) come from the pool
            if (tid < 64) {
                const float a = __ldcg(p.qkv_raw + h * kHeadDim + tid), b = __ldcg(p.qkv_raw + h * kHeadDim + tid + 64);
                const float c = p.cos[tid], s = p.sin[tid];
                q_s[tid] = (a * c - b * s) * scale_log2;
                q_s[tid + 64] = (b * c + a * s) * scale_log2;
            } else if (tid < 128) {
                if (has_new) {
                    const int j = tid - 64;
                    const float* kr = p.qkv_raw + (size_t)(p.n_heads + kvh) * kHeadDim;
                    const float a = __ldcg(kr + j), b = __ldcg(kr + j + 64);
                    const float c = p.cos[j], s = p.sin[j];
                    const __nv_bfloat16 lo = __float2bfloat16_rn(a * c - b * s), hi = __float2bfloat16_rn(b * c + a * s);
                    knew_s[j] = __bfloat162float(lo);
                    knew_s[j + 64] = __bfloat162float(hi);
                    if (h % group == 0) {
                        __nv_bfloat16* kd = kp + p.kv_off[0] + (int64_t)kvh * p.head_stride + j;
                        kd[0] = lo;
                        kd[64] = hi;
                    }
                }
            } else if (tid < 192) {
                if (has_new) {
                    const int j = tid - 128;
                    const float* vr = p.qkv_raw + (size_t)(p.n_heads + p.n_kv + kvh) * kHeadDim;
                    const __nv_bfloat16 lo = __float2bfloat16_rn(__ldcg(vr + j)), hi = __float2bfloat16_rn(__ldcg(vr + j + 64));
                    vnew_s[j] = __bfloat162float(lo);
                    vnew_s[j + 64] = __bfloat162float(hi);
                    if (h % group == 0) {
                        __nv_bfloat16* vd = vp + p.kv_off[0] + (int64_t)kvh * p.head_stride + j;
                        vd[0] = lo;
                        vd[64] = hi;
                    }
                }
            }
            consumer_sync();
            float q[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) q[e] = q_s[l16 * 8 + e];
            float m_run = -INFINITY, l_run = 0.f, acc[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) acc[e] = 0.f;
            const int64_t lane_base = (int64_t)kvh * p.head_stride + l16 * 8;
            auto absorb = [&](float d, const float (&v)[8]) {   // one key: online softmax in base 2
                const float m_new = fmaxf(m_run, d);
                const float alpha = exp2f(m_run - m_new), w = exp2f(d - m_new);
                l_run = l_run * alpha + w;
#pragma unroll
                for (int e = 0; e < 8; ++e) acc[e] = fmaf(w, v[e], acc[e] * alpha);
                m_run = m_new;
            };
            for (int tb = t0; tb < t_pool; tb += 32 * U) {   // trip count uniform across the CTA (the shuffles need full warps)
                if (!(item == (int)blockIdx.x && tb == t0)) {
#pragma unroll
                    for (int u = 0; u < U; ++u) {
                        const int t = tb + grp + 32 * u;
                        if (t < t_pool) {
                            const int64_t off = (int64_t)bt[t / p.page_size] * page_stride + lane_base + (int64_t)(t % p.page_size) * kHeadDim;
                            kk[u] = *reinterpret_cast<const uint4*>(kp + off);
                            vv[u] = *reinterpret_cast<const uint4*>(vp + off);
                        }
                    }
                }
                float d[U];
#pragma unroll
                for (int u = 0; u < U; ++u) d[u] = dot8bf(kk[u], q);
#pragma unroll
                for (int o = 8; o > 0; o >>= 1)
#pragma unroll
                    for (int u = 0; u < U; ++u) d[u] += __shfl_xor_sync(0xffffffffu, d[u], o);
#pragma unroll
                for (int u = 0; u < U; ++u) {
                    if (tb + grp + 32 * u < t_pool) {
                        const uint4 v = vv[u];
                        const float vf[8] = {bf_lo(v.x), bf_hi(v.x), bf_lo(v.y), bf_hi(v.y), bf_lo(v.z), bf_hi(v.z), bf_lo(v.w), bf_hi(v.w)};
                        absorb(d[u], vf);
                    }
                }
            }
            if (has_new && warp == 0) {   // the token's own key and value, straight from the projection (rounded to bf16 like the cache)
                float d = 0.f;
#pragma unroll
                for (int e = 0; e < 8; ++e) d = fmaf(knew_s[l16 * 8 + e], q[e], d);
#pragma unroll
                for (int o = 8; o > 0; o >>= 1) d += __shfl_xor_sync(0xffffffffu, d, o);
                if (grp == 0) {
                    float vf[8];
#pragma unroll
                    for (int e = 0; e < 8; ++e) vf[e] = vnew_s[l16 * 8 + e];
                    absorb(d, vf);
                }
            }
            consumer_sync();   // q_s / knew_s / vnew_s were read by everyone; g_acc does not overlap them
#pragma unroll
            for (int e = 0; e < 8; ++e) g_acc[grp * kHeadDim + l16 * 8 + e] = acc[e];
            if (l16 == 0) { g_m[grp] = m_run; g_l[grp] = l_run; }
            consumer_sync();
            float num = 0.f, den = 0.f, mx = -INFINITY;
            if (tid < kHeadDim) {
                for (int g = 0; g < 32; ++g) mx = fmaxf(mx, g_m[g]);
                for (int g = 0; g < 32; ++g) {
                    const float w = (g_m[g] == -INFINITY) ? 0.f : exp2f(g_m[g] - mx);
                    num = fmaf(w, g_acc[g * kHeadDim + tid], num);
                    den = fmaf(w, g_l[g], den);
                }
            }
            if (p.n_splits == 1) {
                if (tid < kHeadDim) p.attn_out[h * kHeadDim + tid] = num / den;
            } else {
                float* part = p.part + ((size_t)h * p.n_splits + split) * 130;
                if (tid < kHeadDim) {
                    part[tid] = num;
                    if (tid == 0) { part[128] = mx; part[129] = den; }
                    __threadfence();
                }
                consumer_sync();
                if (tid == 0) {
                    const unsigned int prev = atomicAdd(p.tickets + h, 1u);
                    s_last = (prev == (unsigned int)p.n_splits - 1u);
                    if (s_last) p.tickets[h] = 0u;
                }
                consumer_sync();
                if (s_last && tid < kHeadDim) {
                    __threadfence();
                    const float* all = p.part + (size_t)h * p.n_splits * 130;
                    float gmx = -INFINITY;
                    for (int s = 0; s < p.n_splits; ++s) gmx = fmaxf(gmx, __ldcg(all + s * 130 + 128));
                    float n2 = 0.f, d2 = 0.f;
                    for (int s = 0; s < p.n_splits; ++s) {
                        const float pm = __ldcg(all + s * 130 + 128);
                        const float w = (pm == -INFINITY) ? 0.f : exp2f(pm - gmx);
                        n2 = fmaf(w, __ldcg(all + s * 130 + tid), n2);
                        d2 = fmaf(w, __ldcg(all + s * 130 + 129), d2);
                    }
                    p.attn_out[h * kHeadDim + tid] = n2 / d2;
                }
            }
            consumer_sync();   // the scratch is reused by the next item / the next staging
        }
        mark(l);
        grid_barrier(p.bar, gen += gridDim.x, tid);
        mark(l);
        // ---------------------------------------------------------------- o_proj + residual (modeling_llama.py:325)
        for (int i = tid; i < nq / 4; i += kMgConsumers) reinterpret_cast<float4*>(xs)[xs_f4(i)] = __ldcg(reinterpret_cast<const float4*>(p.attn_out) + i);
        consumer_sync();
        mark(l);
        mega_gemv<MG_RESID>(sh, cidx, cpar, xs, part_s, pbuf, H, nq, 1.f, p.x, tid, warp, lane);
        mark(l);
        grid_barrier(p.bar, gen += gridDim.x, tid);
        mark(l);
        // ---------------------------------------------------------------- RMSNorm + gate / up projections
        {
            float ss = 0.f;
            for (int i = tid; i < H / 4; i += kMgConsumers) {
                const float4 v = __ldcg(reinterpret_cast<const float4*>(p.x) + i);
                const uint2 w = *reinterpret_cast<const uint2*>(L.norm2 + 4 * i);
                ss += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
                reinterpret_cast<float4*>(xs)[xs_f4(i)] = make_float4(v.x * bf_lo(w.x), v.y * bf_hi(w.x), v.z * bf_lo(w.y), v.w * bf_hi(w.y));
            }
            ss = consumer_sum(ss, red_s, warp, lane);
            mark(l);
            mega_gemv<MG_RAW>(sh, cidx, cpar, xs, part_s, pbuf, 2 * I, H, rsqrtf(ss / (float)H + p.eps), p.gu_raw, tid, warp, lane);
            mark(l);
        }
        grid_barrier(p.bar, gen += gridDim.x, tid);
        mark(l);
        // ---------------------------------------------------------------- SiLU(gate) * up (modeling_llama.py:183), down_proj + residual (:331)
        for (int i = tid; i < I / 4; i += kMgConsumers) {   // gate rows of a 128-row block are its first 64, the up rows the other 64
            const int j = 4 * i;
            const float* gp = p.gu_raw + ((j >> 6) << 7) + (j & 63);
            const float4 g4 = __ldcg(reinterpret_cast<const float4*>(gp)), u4 = __ldcg(reinterpret_cast<const float4*>(gp + 64));
            reinterpret_cast<float4*>(xs)[xs_f4(i)] = make_float4(silu(g4.x) * u4.x, silu(g4.y) * u4.y, silu(g4.z) * u4.z, silu(g4.w) * u4.w);
        }
        consumer_sync();
        mark(l);
        mega_gemv<MG_RESID>(sh, cidx, cpar, xs, part_s, pbuf, H, I, 1.f, p.x, tid, warp, lane);
        mark(l);
        if (l + 1 < p.n_layers) grid_barrier(p.bar, gen += gridDim.x, tid);
        mark(l);
    }
#if MEGA_PROFILE
    if (p.profile == 1 && tid == 0 && nmark == 15) {
        // absolute ns (low digits) at: start | staged, qkv done, barrier passed | attention done, barrier | staged, o done, barrier |
        // staged, gate/up done, barrier | staged, down done, barrier
        const unsigned long long b = tmark[0] - (tmark[0] % 1000000ull);
        printf("mega %d: %llu %llu %llu %llu %llu %llu %llu %llu %llu %llu %llu %llu %llu %llu %llu\n", (int)blockIdx.x, tmark[0] - b, tmark[1] - b,
               tmark[2] - b, tmark[3] - b, tmark[4] - b, tmark[5] - b, tmark[6] - b, tmark[7] - b, tmark[8] - b, tmark[9] - b, tmark[10] - b,
               tmark[11] - b, tmark[12] - b, tmark[13] - b, tmark[14] - b);
        printf("megab %d: %llu %llu %llu %llu\n", (int)blockIdx.x, tmb[0] - b, tmb[1] - b, tmb[2] - b, tmb[3]);
    }
#endif
}

static char g_mega_err[256];
const char* decode_mega_last_error() { return g_mega_err; }

// shared memory: ring + activation vector (also the attention scratch) + mbarriers
static size_t mega_smem(int n_slots, int xs_floats) { return (size_t)n_slots * kMgSlotBytes + (size_t)xs_floats * 4 + 16 * kMgMaxSlots; }

// 0 = launched, 1 = shape not served (use the per-kernel path), -1 = error
int launch_decode_mega(const MegaParams& p_in, int num_sms, cudaStream_t st) {
    MegaParams p = p_in;
    const int nq = p.n_heads * kHeadDim;
    if (p.H % 8 != 0 || p.I % 64 != 0 || p.n_heads % p.n_kv != 0 || p.n_layers < 1) return 1;
    int xs = p.H > p.I ? p.H : p.I;
    if (nq > xs) xs = nq;
    if (kMgAttnFloats > xs) xs = kMgAttnFloats;
    xs = (xs + 255) & ~255;
    int n_slots = kMgMaxSlots;
    while (n_slots >= 4 && mega_smem(n_slots, xs) > 226 * 1024) --n_slots;   // 227 KB per CTA, less the kernel's static variables
    if (n_slots < 4) return 1;
    { const char* e = getenv("B200SHARD_MEGA_SLOTS"); if (e != nullptr && atoi(e) >= 2 && atoi(e) < n_slots) n_slots = atoi(e); }   // A/B: shallower ring
    p.n_slots = n_slots;
    p.xs_floats = xs;
    int ns = num_sms / p.n_heads;          // attention: one (head, split) item per CTA where the heads allow it
    if (ns < 1) ns = 1;
    if (ns > kMaxSplits) ns = kMaxSplits;
    p.n_splits = ns;
    // measurement modes: 1 = timestamps (MEGA_PROFILE builds only), 2 = no FMA loop, 3 = no FMA loop and no reductions (results are
    // garbage in modes 2 and 3; they exist for the breakdown in profiles/decode_mega_r2.md)
    { const char* e = getenv("B200SHARD_MEGA_PROFILE"); p.profile = e != nullptr ? atoi(e) : 0; }
    const size_t smem = mega_smem(n_slots, xs);
    if (cudaFuncSetAttribute(decode_mega_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) {
        snprintf(g_mega_err, sizeof(g_mega_err), "decode_mega: cannot raise the shared-memory limit");
        cudaGetLastError();
        return -1;
    }
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(num_sms);
    cfg.blockDim = dim3(kMgThreads);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeCooperative;
    attr[0].val.cooperative = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    const cudaError_t err = cudaLaunchKernelEx(&cfg, decode_mega_kernel, p);
    if (err != cudaSuccess) {
        snprintf(g_mega_err, sizeof(g_mega_err), "decode_mega launch failed: %s", cudaGetErrorString(err));
        cudaGetLastError();
        return -1;
    }
    return 0;
}

}  // namespace b200
