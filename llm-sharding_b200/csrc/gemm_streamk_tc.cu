// Weight-streaming projection GEMM for batched decode and short prompts, 4..32 token rows, second generation:
// out[m][n] = sum_k W[n][k] * x[m][k], persistent CTAs with an even ("stream-K") split of the weight bytes.
// Replaces the cuBLAS calls under q/k/v/o_proj and gate/up/down_proj (transformers modeling_llama.py:182-184,
// 262-264,287-288, reached from utils/shard_loader.py:66-74) when a shard is asked for one token of each of a few
// dozen requests (BASELINE configs[2]: micro-batched batch-32 decode).
//
// Why a second kernel: the first one (gemm_stream_tc.cu) gave every CTA one 128-row tile (or 1/S of it, S CTAs of a
// cluster) and launched tiles x S CTAs.  Tile counts (96 / 32 / 172 / 32 for a Llama-2-7B layer) do not divide by the
// 148 SMs, so some SMs streamed twice the bytes of others, every CTA made only two or three trips round a 4-stage ring,
// and the split-K ones paid a cluster launch and a DSMEM reduce-scatter with nothing overlapped: 2.1 - 4.1 TB/s
// (profiles/results_r1.md).  Here the unit of work is one k-step (128 weight rows x 64 k = 16 KB) of one tile; the
// tiles x k-steps units are numbered tile-major and cut into one CONTIGUOUS, EQUAL range per SM, so every SM streams
// the same number of bytes through one deep ring that never drains between tiles.
//
//   warp 0     TMA producer: walks the CTA's unit range; 128 x 64 weight tiles + (16|32) x 64 token tiles, 128-byte
//              swizzle, deep mbarrier ring (~200 KB in flight per SM).  The ring is filled with weight tiles BEFORE
//              griddepcontrol.wait (weights do not depend on the previous kernel).  Every tensor map covers memory of
//              this GPU only (the shard's weights and workspace).
//   warp 1     one thread issues tcgen05.mma (UMMA 128 x MP x 16, fp16 x fp16 -> fp32).  Two 32-column accumulators in
//              tensor memory: the epilogue of one tile overlaps the MMAs of the next.
//   warps 2-5  epilogue: one weight row per thread out of TMEM.
//
// A range that starts or ends inside a tile leaves that tile's sum split over consecutive CTAs.  The CTA that holds the
// tile's FIRST k-step owns it; the others hand their partial 128 x MP block to the owner through a per-CTA slot of a
// global workspace followed by a release of the slot's flag.  A contributor's fragment is always the first thing in its
// range and is published without waiting for anybody; the owner meets the unfinished tile at the END of its range and
// then acquires the flags — all CTAs are resident at once (grid <= number of SMs), so the wait is short and cannot
// deadlock.  Partials are added in CTA order: the result is bit-reproducible.  The owner then runs the fused epilogue of
// the first kernel: fp32 store, fp32 residual add in place (modeling_llama.py:325,331), SiLU * mul -> fp16 (:183), or
// RoPE + paged K/V append + fp32 q (:146-168, cache_utils.py:102-121).
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>

#include <mutex>

#include "common.cuh"
#include "internal.h"
#include "tc_common.cuh"

namespace b200 {
using namespace tc;

constexpr int kSkThreads = 192;
constexpr int kSkRows = 128;                         // weight rows per tile (UMMA M)
constexpr int kSkK = 64;                             // k per unit
constexpr uint32_t kSkStageW = kSkRows * kSkK * 2;   // 16 KB
constexpr uint32_t kSkTmemCols = 64;                 // two accumulators of up to 32 columns

__device__ __forceinline__ void sk_tmem_ld16(uint32_t taddr, uint32_t* r) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr));
}
__device__ __forceinline__ void sk_st_release(unsigned int* p, unsigned int v) {
    asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned int sk_ld_acquire(const unsigned int* p) {
    unsigned int v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void sk_epi_sync() { asm volatile("bar.sync 1, 128;" ::: "memory"); }  // the four epilogue warps

template <int MP>
constexpr uint32_t sk_stage_bytes() { return kSkStageW + MP * kSkK * 2; }
template <int MP, int NS>
constexpr size_t sk_smem_bytes() { return 1024 + (size_t)NS * sk_stage_bytes<MP>() + (size_t)kSkRows * MP * 4 + 256 + 128; }

// first unit of CTA c when `total` units are cut into `grid` equal contiguous ranges
__device__ __forceinline__ int64_t sk_range_begin(int64_t total, int grid, int c) { return total * c / grid; }

template <int MP, int EPI, int NS>
__global__ void __launch_bounds__(kSkThreads, NS > 5 ? 1 : 2)
gemm_streamk_tc_kernel(const __grid_constant__ CUtensorMap map_w, const __grid_constant__ CUtensorMap map_x, const StreamGemmParams p) {
    constexpr uint32_t kStage = sk_stage_bytes<MP>();
    extern __shared__ uint8_t smem_raw[];
    const uint32_t raw = s_u32(smem_raw);
    const uint32_t base = (raw + 1023u) & ~1023u;  // swizzled tiles need 1024-byte alignment
    uint8_t* gbase = smem_raw + (base - raw);
    float* red = reinterpret_cast<float*>(gbase + NS * kStage);   // [2 halves][MP][64]: one finished tile, rows j and j + 64 side by side
    const uint32_t bars = base + NS * kStage + kSkRows * MP * 4;
    const uint32_t full0 = bars, empty0 = bars + 8 * NS, tfull0 = bars + 16 * NS, tempty0 = tfull0 + 16;
    volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(gbase + NS * kStage + kSkRows * MP * 4 + 16 * NS + 32);
    float* rstd_s = reinterpret_cast<float*>(gbase + NS * kStage + kSkRows * MP * 4 + 256);   // [MP] 1 / rms of every token row (folded RMSNorm)

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int cta = blockIdx.x, grid = gridDim.x;
    const int ks = (p.K + kSkK - 1) / kSkK;
    const int n_tiles = (p.N + kSkRows - 1) / kSkRows;
    const int64_t total = (int64_t)n_tiles * ks;
    const int64_t u0 = sk_range_begin(total, grid, cta), u1 = sk_range_begin(total, grid, cta + 1);
    const int n_units = (int)(u1 - u0);
    constexpr uint32_t idesc = umma_idesc(kSkRows, MP, 0, 0, 0);

    if (warp == 0 && lane == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_w) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_x) : "memory");
        for (int i = 0; i < NS; ++i) { mbar_init(full0 + 8 * i, 1); mbar_init(empty0 + 8 * i, 1); }
        for (int i = 0; i < 2; ++i) { mbar_init(tfull0 + 8 * i, 1); mbar_init(tempty0 + 8 * i, 4); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(s_u32((const void*)tmem_slot)), "n"(kSkTmemCols) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    fence_before();
    __syncthreads();
    fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        if (lane == 0) {  // ===== TMA producer: one continuous stream over the CTA's units, whatever tiles they belong to
            const int pre = n_units < NS ? n_units : NS;
            for (int i = 0; i < pre; ++i) {  // weights first: they do not depend on the previous kernel
                const int64_t u = u0 + i;
                mbar_expect_tx(full0 + 8 * i, kStage);
                tma_load_2d(&map_w, full0 + 8 * i, base + i * kStage, (int)(u % ks) * kSkK, (int)(u / ks) * kSkRows);
            }
            asm volatile("griddepcontrol.wait;" ::: "memory");
            asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
            for (int i = 0; i < pre; ++i) tma_load_2d(&map_x, full0 + 8 * i, base + i * kStage + kSkStageW, (int)((u0 + i) % ks) * kSkK, 0);
            for (int i = pre; i < n_units; ++i) {
                const int st = i % NS;
                const int64_t u = u0 + i;
                mbar_wait(empty0 + 8 * st, ((uint32_t)(i / NS) & 1u) ^ 1u);
                mbar_expect_tx(full0 + 8 * st, kStage);
                tma_load_2d(&map_w, full0 + 8 * st, base + st * kStage, (int)(u % ks) * kSkK, (int)(u / ks) * kSkRows);
                tma_load_2d(&map_x, full0 + 8 * st, base + st * kStage + kSkStageW, (int)(u % ks) * kSkK, 0);
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {  // ===== MMA issuer: one accumulator per run of units of the same tile, alternating between two
            int i = 0;
            uint32_t si = 0;
            for (int64_t u = u0; u < u1; ++si) {
                const int kb0 = (int)(u % ks);
                const int n = (int)((ks - kb0) < (u1 - u) ? (ks - kb0) : (u1 - u));
                const uint32_t buf = si & 1u;
                mbar_wait(tempty0 + 8 * buf, ((si >> 1) & 1u) ^ 1u);   // the epilogue has read this accumulator's previous tile
                fence_after();
                const uint32_t acc = tmem_base + buf * 32;
                for (int k = 0; k < n; ++k, ++i) {
                    const int st = i % NS;
                    mbar_wait(full0 + 8 * st, (uint32_t)(i / NS) & 1u);
                    fence_after();
                    const uint32_t sa = base + st * kStage, sb = sa + kSkStageW;
#pragma unroll
                    for (int kk = 0; kk < kSkK / 16; ++kk)
                        umma_f16(acc, umma_desc_sw128(sa + kk * 32, 16, 1024), umma_desc_sw128(sb + kk * 32, 16, 1024), idesc,
                                 (k > 0 || kk > 0) ? 1u : 0u);
                    umma_commit(empty0 + 8 * st);
                }
                umma_commit(tfull0 + 8 * buf);
                u += n;
            }
        }
    } else {  // ===== epilogue warps: TMEM lane = weight row inside the tile
        asm volatile("griddepcontrol.wait;" ::: "memory");   // the residual stream / KV pages they touch belong to the previous kernel
        const int quad = warp & 3;
        const int t = quad * 32 + lane;
        const int j = t & 63, half = t >> 6;
        if ((EPI == EPI_QKV || EPI == EPI_SWIGLU || EPI == EPI_STORE) && p.x_res != nullptr) {   // folded RMSNorm, see gemm_stream_tc.cu
            for (int m = quad; m < p.M; m += 4) {
                const float4* xr = reinterpret_cast<const float4*>(p.x_res + (size_t)m * p.K);
                float ss = 0.f;
                for (int i = lane; i < p.K / 4; i += 32) {
                    const float4 x4 = xr[i];
                    ss += x4.x * x4.x + x4.y * x4.y + x4.z * x4.z + x4.w * x4.w;
                }
                ss = warp_sum(ss);
                if (lane == 0) rstd_s[m] = rsqrtf(ss / (float)p.K + p.eps);
            }
            sk_epi_sync();
        }
        uint32_t si = 0;
        for (int64_t u = u0; u < u1; ++si) {
            const int tile = (int)(u / ks), kb0 = (int)(u % ks);
            const int n = (int)((ks - kb0) < (u1 - u) ? (ks - kb0) : (u1 - u));
            u += n;
            const uint32_t buf = si & 1u;
            mbar_wait(tfull0 + 8 * buf, (si >> 1) & 1u);
            fence_after();
            float v[MP];
            {
                uint32_t r[MP];
                const uint32_t taddr = tmem_base + buf * 32 + ((uint32_t)(quad * 32) << 16);
                if constexpr (MP == 32) tmem_ld32_nowait(taddr, r); else sk_tmem_ld16(taddr, r);
                tmem_ld_wait();
#pragma unroll
                for (int m = 0; m < MP; ++m) v[m] = __uint_as_float(r[m]);
            }
            fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(tempty0 + 8 * buf);   // the issuer may start the tile after next in this accumulator
            if (kb0 != 0) {
                // ---- contributor: the tile began in an earlier CTA; hand the partial block to its owner and move on
                float* slot = p.ws + (size_t)cta * MP * kSkRows;
#pragma unroll
                for (int m = 0; m < MP; ++m) slot[m * kSkRows + t] = v[m];
                __threadfence();
                sk_epi_sync();
                if (t == 0) sk_st_release(p.ws_flags + cta, 1u);
                continue;
            }
            if (kb0 + n < ks) {
                // ---- owner of a tile that continues in the following CTAs: wait for all of them at once (one thread per
                // contributor), then add their partials in CTA order with every load of the gather in flight together
                const int64_t tile_end = (int64_t)(tile + 1) * ks;
                int n_c = 0;
                while (cta + 1 + n_c < grid && sk_range_begin(total, grid, cta + 1 + n_c) < tile_end) ++n_c;
                if (t < n_c) {
                    const long long t0 = clock64();
                    while (sk_ld_acquire(p.ws_flags + cta + 1 + t) == 0u)
                        if (clock64() - t0 > (1ll << 31)) __trap();   // ~1 s: a protocol bug must not hang the GPU
                }
                sk_epi_sync();
                constexpr int Q = 64 / MP;   // contributors' blocks per pass: 64 independent loads per thread
                for (int c0 = 0; c0 < n_c; c0 += Q) {
                    float w[Q][MP];
#pragma unroll
                    for (int q = 0; q < Q; ++q) {
                        const float* slot = p.ws + (size_t)(cta + 1 + c0 + q) * MP * kSkRows;
#pragma unroll
                        for (int m = 0; m < MP; ++m) w[q][m] = (c0 + q < n_c) ? __ldcg(slot + m * kSkRows + t) : 0.f;
                    }
#pragma unroll
                    for (int q = 0; q < Q; ++q)
#pragma unroll
                        for (int m = 0; m < MP; ++m) v[m] += w[q][m];
                }
                sk_epi_sync();
                if (t < n_c) p.ws_flags[cta + 1 + t] = 0u;   // consumed: the slots are free for the next launch on this stream
            }
            // ---- finished tile: rows j and j + 64 (a RoPE pair / a gate-up pair) meet in shared memory
#pragma unroll
            for (int m = 0; m < MP; ++m) red[(half * MP + m) * 64 + j] = v[m];
            sk_epi_sync();
            const int n_heads = p.n_heads, n_kv = p.n_kv;
            for (int e = t; e < 64 * p.M; e += 128) {
                const int jj = e & 63, m = e >> 6;
                float a = red[(0 * MP + m) * 64 + jj], b = red[(1 * MP + m) * 64 + jj];
                if ((EPI == EPI_QKV || EPI == EPI_SWIGLU || EPI == EPI_STORE) && p.x_res != nullptr) { a *= rstd_s[m]; b *= rstd_s[m]; }
                const int row0 = tile * kSkRows + jj, row1 = row0 + 64;
                if (EPI == EPI_STORE || EPI == EPI_RESID) {
                    if (row0 < p.N) {
                        float* o = p.out + (size_t)m * p.N + row0;
                        const float r = (EPI == EPI_RESID) ? (*o + a) : a;
                        *o = r;
                        if (EPI == EPI_RESID && p.xw_out != nullptr)
                            p.xw_out[(size_t)m * p.N + row0] = __float2half_rn(fminf(fmaxf(r * __bfloat162float(p.next_norm_w[row0]), -65504.f), 65504.f));
                    }
                    if (row1 < p.N) {
                        float* o = p.out + (size_t)m * p.N + row1;
                        const float r = (EPI == EPI_RESID) ? (*o + b) : b;
                        *o = r;
                        if (EPI == EPI_RESID && p.xw_out != nullptr)
                            p.xw_out[(size_t)m * p.N + row1] = __float2half_rn(fminf(fmaxf(r * __bfloat162float(p.next_norm_w[row1]), -65504.f), 65504.f));
                    }
                } else if (EPI == EPI_SWIGLU) {   // tile = 64 gate rows + the matching 64 up rows
                    if (row0 < p.N) p.act_h[(size_t)m * (p.N >> 1) + tile * 64 + jj] = __float2half_rn(fminf(fmaxf(silu(a) * b, -65504.f), 65504.f));
                } else {                          // EPI_QKV: tile == head, rows jj and jj + 64 are a rotate_half pair
                    const int head = tile;
                    if (head < n_heads + n_kv) {
                        const float c = p.cos[m * kHalfDim + jj], sn = p.sin[m * kHalfDim + jj];
                        const float lo = a * c - b * sn, hi = b * c + a * sn;
                        if (head < n_heads) {
                            float* q = p.q_out + (size_t)m * n_heads * kHeadDim + head * kHeadDim + jj;
                            q[0] = lo;
                            q[kHalfDim] = hi;
                        } else {
                            __nv_bfloat16* kd = p.k_pool + p.kv_off[m] + (int64_t)(head - n_heads) * p.head_stride + jj;
                            kd[0] = __float2bfloat16_rn(lo);
                            kd[kHalfDim] = __float2bfloat16_rn(hi);
                        }
                    } else {
                        __nv_bfloat16* vd = p.v_pool + p.kv_off[m] + (int64_t)(head - n_heads - n_kv) * p.head_stride + jj;
                        vd[0] = __float2bfloat16_rn(a);
                        vd[kHalfDim] = __float2bfloat16_rn(b);
                    }
                }
            }
            sk_epi_sync();   // `red` is free for the next tile
        }
    }
    __syncthreads();
    if (warp == 1) {
        fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(kSkTmemCols) : "memory");
    }
}

// ---- host side
static char g_sk_err[256];
const char* gemm_streamk_last_error() { return g_sk_err; }

size_t gemm_streamk_ws_floats(int num_sms) { return (size_t)2 * num_sms * kMaxStreamRows * kSkRows; }  // up to two CTAs per SM

// Workspace of the stand-alone entry point (b200_gemm_stream_f16): one per device, allocated on first use and kept.  A shard
// passes its own (StreamGemmParams::ws), so shards on different streams never share partial slots.
static std::mutex g_sk_mu;
static float* g_sk_ws[64];
static unsigned int* g_sk_flags[64];
static int default_workspace(int num_sms, float** ws, unsigned int** flags) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return -1;
    std::lock_guard<std::mutex> lock(g_sk_mu);
    if (g_sk_ws[dev] == nullptr) {
        void *a = nullptr, *b = nullptr;
        if (cudaMalloc(&a, gemm_streamk_ws_floats(num_sms > 256 ? num_sms : 256) * sizeof(float)) != cudaSuccess) return -1;
        if (cudaMalloc(&b, 1024 * sizeof(unsigned int)) != cudaSuccess) { cudaFree(a); return -1; }
        cudaMemset(b, 0, 1024 * sizeof(unsigned int));
        cudaDeviceSynchronize();
        g_sk_ws[dev] = reinterpret_cast<float*>(a);
        g_sk_flags[dev] = reinterpret_cast<unsigned int*>(b);
    }
    *ws = g_sk_ws[dev];
    *flags = g_sk_flags[dev];
    return 0;
}

template <int MP, int EPI, int NS>
static int sk_launch_one(const CUtensorMap& mw, const CUtensorMap& mx, const StreamGemmParams& p, int grid, cudaStream_t st) {
    auto kfn = gemm_streamk_tc_kernel<MP, EPI, NS>;
    constexpr size_t smem = sk_smem_bytes<MP, NS>();
    static_assert(smem <= 227 * 1024, "ring does not fit in shared memory");
    cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(kSkThreads);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    int na = 0;
    const char* e = getenv("B200SHARD_PDL");
    if (!(e != nullptr && e[0] == '0')) {
        attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[na].val.programmaticStreamSerializationAllowed = 1;
        ++na;
    }
    cfg.attrs = attr;
    cfg.numAttrs = na;
    const cudaError_t err = cudaLaunchKernelEx(&cfg, kfn, mw, mx, p);
    if (err != cudaSuccess) {
        snprintf(g_sk_err, sizeof(g_sk_err), "gemm_streamk launch failed: %s", cudaGetErrorString(err));
        return -1;
    }
    return 0;
}

int launch_gemm_streamk(const StreamGemmParams& p_in, int epi, int num_sms, cudaStream_t st) {
    StreamGemmParams p = p_in;
    if (p.M < 1 || p.M > kMaxStreamRows || p.K % 8 != 0 || p.K < kSkK || p.N < 1) {
        snprintf(g_sk_err, sizeof(g_sk_err), "gemm_stream: need 1 <= M <= %d, K %% 8 == 0, K >= 64 (M=%d N=%d K=%d)", kMaxStreamRows,
                 p.M, p.N, p.K);
        return -1;
    }
    if ((epi == EPI_QKV || epi == EPI_SWIGLU) && p.N % kSkRows != 0) {
        snprintf(g_sk_err, sizeof(g_sk_err), "gemm_stream: paired epilogues need N %% 128 == 0 (N=%d)", p.N);
        return -1;
    }
    if (p.ws == nullptr || p.ws_flags == nullptr) {
        if (default_workspace(num_sms, &p.ws, &p.ws_flags) != 0) {
            snprintf(g_sk_err, sizeof(g_sk_err), "gemm_stream: could not allocate the partial-sum workspace");
            return -1;
        }
    }
    const int MP = p.M <= 16 ? 16 : 32;
    CUtensorMap mw, mx;
    if (tma_map_2d(&mw, p.W, p.N, p.K, kSkRows) != 0 || tma_map_2d(&mx, p.X, p.M, p.K, MP) != 0) {
        snprintf(g_sk_err, sizeof(g_sk_err), "gemm_stream: tensor map creation failed");
        return -1;
    }
    const int64_t total = (int64_t)((p.N + kSkRows - 1) / kSkRows) * ((p.K + kSkK - 1) / kSkK);
    static int deep = -1, waves = -1;
    if (deep < 0) {
        const char* e = getenv("B200SHARD_SK_DEEP");    // 1: one CTA per SM with a ~200 KB ring; 0: ~100 KB rings, two CTAs fit an SM
        deep = (e != nullptr && e[0] == '0') ? 0 : 1;
        e = getenv("B200SHARD_SK_WAVES");               // CTAs per SM of this grid (1 leaves the second slot to the next kernel's prologue)
        waves = (e != nullptr && e[0] == '2') ? 2 : 1;
    }
    const int64_t want = (int64_t)num_sms * (deep ? 1 : waves);
    const int grid = (int)(total < want ? total : want);
    // ring depth: 16 KB of weights + 2 or 4 KB of token rows per stage
#define B200_SK_CASE(E) (deep ? (MP == 16 ? sk_launch_one<16, E, 11>(mw, mx, p, grid, st) : sk_launch_one<32, E, 10>(mw, mx, p, grid, st)) \
                              : (MP == 16 ? sk_launch_one<16, E, 5>(mw, mx, p, grid, st) : sk_launch_one<32, E, 5>(mw, mx, p, grid, st)))
    switch (epi) {
        case EPI_STORE: return B200_SK_CASE(EPI_STORE);
        case EPI_RESID: return B200_SK_CASE(EPI_RESID);
        case EPI_QKV: return B200_SK_CASE(EPI_QKV);
        case EPI_SWIGLU: return B200_SK_CASE(EPI_SWIGLU);
    }
#undef B200_SK_CASE
    snprintf(g_sk_err, sizeof(g_sk_err), "gemm_stream: unknown epilogue %d", epi);
    return -1;
}

}  // namespace b200
