// Weight-streaming projection GEMM for batched decode, 9..32 token rows: out[m][n] = sum_k W[n][k] * x[m][k].
// Replaces the cuBLAS calls under q/k/v/o_proj and gate/up/down_proj (transformers modeling_llama.py:182-184,
// 262-264,287-288, reached from utils/shard_loader.py:66-74) when a shard is asked for one token of each of a few
// dozen requests (BASELINE configs[2]: micro-batched batch-32 decode).  At these sizes the layer is still bound by
// reading its weights once from HBM, so the kernel is built like the 1..8-row GEMVs — every CTA streams a slab of
// weights exactly once, all CTAs resident at the same time — but the multiply runs on the 5th-generation tensor
// cores with the operand roles swapped: the WEIGHT tile is the 128-row A operand, the token rows are the narrow
// (16 or 32 column) B operand, so no tensor-core work is spent on padding tokens up to a 128-row tile.
//
//   warp 0     TMA producer: 128 x 64 weight tiles + (16|32) x 64 activation tiles, 128-byte swizzle, 4-stage mbarrier
//              ring.  The first weight tiles are requested BEFORE griddepcontrol.wait (weights do not depend on the
//              previous kernel); the activation tiles follow once the wait has returned.
//   warp 1     one thread issues tcgen05.mma (UMMA 128 x MP x 16, fp16 x fp16 -> fp32 in 32 TMEM columns)
//   warps 2-5  epilogue: one weight row per thread out of TMEM (tcgen05.ld 32x32b)
//
// Split-K over a thread-block cluster: the GEMMs with few weight rows (o_proj, down_proj: 32 tiles of 128 rows) could
// not occupy 148 SMs, so S = 2 | 4 | 8 CTAs of a cluster share one tile, each streaming 1/S of K.  The partial sums are
// exchanged reduce-scatter style through distributed shared memory: the CTA of rank d gathers, from every rank, the
// partials of the 128/S rows it owns (rows j and j + 64 stay together: they are the two halves of a RoPE pair / the
// gate and up rows of a SwiGLU pair), one cluster barrier orders the stores, and each CTA sums its rows in rank order
// (deterministic) and runs the fused epilogue: fp32 store, fp32 residual add in place (modeling_llama.py:325,331),
// SiLU * mul -> fp16 (:183), or RoPE + paged K/V append + fp32 q (:146-168, cache_utils.py:102-121).
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>

#include "common.cuh"
#include "internal.h"
#include "tc_common.cuh"

namespace b200 {
using namespace tc;

constexpr int kSwThreads = 192;
constexpr int kSwRows = 128;                         // weight rows per tile (UMMA M)
constexpr int kSwK = 64;                             // k per pipeline stage
constexpr uint32_t kSwStageW = kSwRows * kSwK * 2;   // 16 KB
constexpr uint32_t kSwTmemCols = 32;

__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t* r) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr));
}

template <int MP>
constexpr uint32_t sw_stage_bytes() { return kSwStageW + MP * kSwK * 2; }
template <int MP, int kSwStages>
constexpr size_t sw_smem_bytes() { return 1024 + (size_t)kSwStages * sw_stage_bytes<MP>() + (size_t)kSwRows * MP * 4 + 128 + 128; }

template <int MP, int EPI, int kSwStages>
__global__ void __launch_bounds__(kSwThreads, 2)
gemm_stream_tc_kernel(const __grid_constant__ CUtensorMap map_w, const __grid_constant__ CUtensorMap map_x, const StreamGemmParams p) {
    constexpr uint32_t kStage = sw_stage_bytes<MP>();
    extern __shared__ uint8_t smem_raw[];
    const uint32_t raw = s_u32(smem_raw);
    const uint32_t base = (raw + 1023u) & ~1023u;  // swizzled tiles need 1024-byte alignment
    uint8_t* gbase = smem_raw + (base - raw);
    float* red = reinterpret_cast<float*>(gbase + kSwStages * kStage);           // [S][2][MP][64/S] partials gathered here
    const uint32_t red_u32 = base + kSwStages * kStage;
    const uint32_t bars = red_u32 + kSwRows * MP * 4;
    const uint32_t full0 = bars, empty0 = bars + 8 * kSwStages, tmem_full = bars + 16 * kSwStages;
    volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(gbase + kSwStages * kStage + kSwRows * MP * 4 + 16 * kSwStages + 8);
    float* rstd_s = reinterpret_cast<float*>(gbase + kSwStages * kStage + kSwRows * MP * 4 + 128);   // [MP] 1 / rms of every token row (folded RMSNorm)

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int S = p.split;
    uint32_t rank = 0;
    if (S > 1) {
        asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(rank));
        // distributed shared memory may be written only once the owning CTA runs: arrive now, wait before the first remote store
        asm volatile("barrier.cluster.arrive.relaxed.aligned;" ::: "memory");
    }
    const int tile = blockIdx.x / S;
    const int ks_total = (p.K + kSwK - 1) / kSwK;
    const int k0 = (int)((int64_t)ks_total * rank / S), k1 = (int)((int64_t)ks_total * (rank + 1) / S);
    const int nk = k1 - k0;
    constexpr uint32_t idesc = umma_idesc(kSwRows, MP, 0, 0, 0);

    if (warp == 0 && lane == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_w) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_x) : "memory");
        for (int i = 0; i < kSwStages; ++i) { mbar_init(full0 + 8 * i, 1); mbar_init(empty0 + 8 * i, 1); }
        mbar_init(tmem_full, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(s_u32((const void*)tmem_slot)), "n"(kSwTmemCols) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    fence_before();
    __syncthreads();
    fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        if (lane == 0) {  // ===== TMA producer
            const int pre = nk < kSwStages ? nk : kSwStages;
            for (int i = 0; i < pre; ++i) {  // weights first: they do not depend on the previous kernel
                mbar_expect_tx(full0 + 8 * i, kStage);
                tma_load_2d(&map_w, full0 + 8 * i, base + i * kStage, (k0 + i) * kSwK, tile * kSwRows);
            }
            asm volatile("griddepcontrol.wait;" ::: "memory");
            asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
            for (int i = 0; i < pre; ++i) tma_load_2d(&map_x, full0 + 8 * i, base + i * kStage + kSwStageW, (k0 + i) * kSwK, 0);
            for (int kb = pre; kb < nk; ++kb) {
                const int st = kb % kSwStages;
                mbar_wait(empty0 + 8 * st, ((uint32_t)(kb / kSwStages) & 1u) ^ 1u);
                mbar_expect_tx(full0 + 8 * st, kStage);
                tma_load_2d(&map_w, full0 + 8 * st, base + st * kStage, (k0 + kb) * kSwK, tile * kSwRows);
                tma_load_2d(&map_x, full0 + 8 * st, base + st * kStage + kSwStageW, (k0 + kb) * kSwK, 0);
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {  // ===== MMA issuer
            for (int kb = 0; kb < nk; ++kb) {
                const int st = kb % kSwStages;
                mbar_wait(full0 + 8 * st, (uint32_t)(kb / kSwStages) & 1u);
                fence_after();
                const uint32_t sa = base + st * kStage, sb = sa + kSwStageW;
#pragma unroll
                for (int k = 0; k < kSwK / 16; ++k)
                    umma_f16(tmem_base, umma_desc_sw128(sa + k * 32, 16, 1024), umma_desc_sw128(sb + k * 32, 16, 1024), idesc,
                             (kb > 0 || k > 0) ? 1u : 0u);
                umma_commit(empty0 + 8 * st);
            }
            umma_commit(tmem_full);
        }
    } else {  // ===== epilogue, phase 1: TMEM -> the owning rank's gather buffer
        asm volatile("griddepcontrol.wait;" ::: "memory");
        const int quad = warp & 3;
        const int t = quad * 32 + lane;                  // weight row inside the tile = TMEM lane
        if ((EPI == EPI_QKV || EPI == EPI_SWIGLU || EPI == EPI_STORE) && p.x_res != nullptr) {
            // folded RMSNorm: while the MMAs run, every CTA recomputes the row statistics from the fp32 residual stream (same
            // loads, same order in every CTA: identical bits everywhere); one warp per token row, xor-shuffle reduction
            // With row_ss the producer of the residual stream left per-CTA partial sums of squares: M x parts floats to add
            // instead of the M x K block to re-read.
            for (int m = quad; m < p.M; m += 4) {
                float ss = 0.f;
                if (p.row_ss != nullptr) {
                    const float* pr = p.row_ss + (size_t)m * p.row_ss_parts;
                    for (int i = lane; i < p.row_ss_parts; i += 32) ss += pr[i];
                } else {
                    const float4* xr = reinterpret_cast<const float4*>(p.x_res + (size_t)m * p.K);
                    for (int i = lane; i < p.K / 4; i += 32) {
                        const float4 x4 = xr[i];
                        ss += x4.x * x4.x + x4.y * x4.y + x4.z * x4.z + x4.w * x4.w;
                    }
                }
                ss = warp_sum(ss);
                if (lane == 0) rstd_s[m] = rsqrtf(ss / (float)p.K + p.eps);
            }
        }
        uint32_t v[MP];
        if (nk > 0) {
            mbar_wait(tmem_full, 0);
            fence_after();
            const uint32_t taddr = tmem_base + ((uint32_t)(quad * 32) << 16);
            if constexpr (MP == 32) tmem_ld32_nowait(taddr, v); else tmem_ld16(taddr, v);
            tmem_ld_wait();
        } else {
#pragma unroll
            for (int m = 0; m < MP; ++m) v[m] = 0u;
        }
        const int P = 64 / S;                            // row pairs (j, j + 64) owned by one rank
        const int j = t & 63, half = t >> 6;
        const uint32_t owner = (uint32_t)(j / P), jl = (uint32_t)(j % P);
        uint32_t dst = red_u32;
        if (S > 1) {
            asm volatile("barrier.cluster.wait.aligned;" ::: "memory");  // pairs with the arrive at kernel entry
            asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(dst) : "r"(red_u32), "r"(owner));
        }
        dst += 4u * ((rank * 2u + (uint32_t)half) * MP * P + jl);
#pragma unroll
        for (int m = 0; m < MP; ++m)
            asm volatile("st.shared::cluster.f32 [%0], %1;" ::"r"(dst + 4u * (uint32_t)(m * P)), "f"(__uint_as_float(v[m])) : "memory");
        fence_before();
    }
    // every thread of every CTA of the cluster meets here: the partials are in place
    if (S > 1) {
        if (warp < 2) asm volatile("barrier.cluster.wait.aligned;" ::: "memory");  // the producer / issuer warps complete the entry barrier here
        asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
        asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
    } else {
        __syncthreads();
    }
    if (warp == 1) {
        fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(kSwTmemCols) : "memory");
    }
    if (warp < 2) return;

    // ===== epilogue, phase 2: this rank's P row pairs x M tokens, summed over the ranks in rank order
    const int P = 64 / S;
    const int n_heads = p.n_heads, n_kv = p.n_kv;
    float* sq_s = reinterpret_cast<float*>(gbase);       // [M][P] squares of this CTA's new residual values; the ring is idle by now
    for (int e = (int)threadIdx.x - 64; e < P * p.M; e += kSwThreads - 64) {
        const int jl = e % P, m = e / P;
        float a = 0.f, b = 0.f;
        for (int s = 0; s < S; ++s) {
            a += red[((s * 2 + 0) * MP + m) * P + jl];
            b += red[((s * 2 + 1) * MP + m) * P + jl];
        }
        if ((EPI == EPI_QKV || EPI == EPI_SWIGLU || EPI == EPI_STORE) && p.x_res != nullptr) { a *= rstd_s[m]; b *= rstd_s[m]; }
        const int jj = (int)rank * P + jl;               // 0..63 inside the tile
        const int row0 = tile * kSwRows + jj, row1 = row0 + 64;
        if (EPI == EPI_STORE || EPI == EPI_RESID) {
            float sq = 0.f;
            if (row0 < p.N) {
                float* o = p.out + (size_t)m * p.N + row0;
                const float r = (EPI == EPI_RESID) ? (*o + a) : a;
                *o = r;
                sq = r * r;
                if (EPI == EPI_RESID && p.xw_out != nullptr)
                    p.xw_out[(size_t)m * p.N + row0] = __float2half_rn(fminf(fmaxf(r * __bfloat162float(p.next_norm_w[row0]), -65504.f), 65504.f));
            }
            if (row1 < p.N) {
                float* o = p.out + (size_t)m * p.N + row1;
                const float r = (EPI == EPI_RESID) ? (*o + b) : b;
                *o = r;
                sq = fmaf(r, r, sq);
                if (EPI == EPI_RESID && p.xw_out != nullptr)
                    p.xw_out[(size_t)m * p.N + row1] = __float2half_rn(fminf(fmaxf(r * __bfloat162float(p.next_norm_w[row1]), -65504.f), 65504.f));
            }
            if (EPI == EPI_RESID && p.row_ss_out != nullptr) sq_s[e] = sq;   // e = m * P + jl
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
    if (EPI == EPI_RESID && p.row_ss_out != nullptr) {
        // this CTA's share of every token row's sum of squares (folded RMSNorm of the next projection), added in a fixed order
        asm volatile("bar.sync 1, 128;" ::: "memory");   // the four epilogue warps
        const int m = (int)threadIdx.x - 64;
        if (m < p.M) {
            float ss = 0.f;
            for (int jl = 0; jl < P; ++jl) ss += sq_s[m * P + jl];
            p.row_ss_out[(size_t)m * gridDim.x + blockIdx.x] = ss;
        }
    }
}

static char g_stream_err[256];
const char* gemm_stream_last_error() { return g_stream_err; }

// cluster size of the split-K: as many CTAs as it takes to occupy the machine (all of them resident at once, two per
// SM), while every rank keeps at least four k-steps of its own
int gemm_stream_split(int N, int K, int num_sms) {
    const int tiles = (N + kSwRows - 1) / kSwRows, ks = (K + kSwK - 1) / kSwK;
    // measured at 32 rows (us per 7B layer): no split 225, at most 2-way 167, 4-way 147.5, 8-way 146.4; aiming for 200 or
    // 300 CTAs instead, or three shallower CTAs per SM, changes the layer time by < 4 %
    int S = 1;
    while (S < 8 && tiles * S < num_sms && ks / (2 * S) >= 4) S *= 2;
    return S;
}

int gemm_stream_parts(int N, int K, int num_sms) { return ((N + kSwRows - 1) / kSwRows) * gemm_stream_split(N, K, num_sms); }

template <int MP, int EPI, int NS>
static int launch_one(const CUtensorMap& mw, const CUtensorMap& mx, const StreamGemmParams& p, cudaStream_t st) {
    auto kfn = gemm_stream_tc_kernel<MP, EPI, NS>;
    constexpr size_t smem = sw_smem_bytes<MP, NS>();
    cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaLaunchConfig_t cfg{};
    const int tiles = (p.N + kSwRows - 1) / kSwRows;
    cfg.gridDim = dim3(tiles * p.split);
    cfg.blockDim = dim3(kSwThreads);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[2];
    int na = 0;
    if (p.split > 1) {
        attr[na].id = cudaLaunchAttributeClusterDimension;
        attr[na].val.clusterDim.x = p.split;
        attr[na].val.clusterDim.y = 1;
        attr[na].val.clusterDim.z = 1;
        ++na;
    }
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
        snprintf(g_stream_err, sizeof(g_stream_err), "gemm_stream launch failed: %s", cudaGetErrorString(err));
        return -1;
    }
    return 0;
}

static bool streamk_enabled() {  // B200SHARD_STREAMK=1: the persistent stream-K kernel of gemm_streamk_tc.cu (A/B against this file's)
    static int v = -1;
    if (v < 0) {
        const char* e = getenv("B200SHARD_STREAMK");
        v = (e != nullptr && e[0] == '1') ? 1 : 0;
    }
    return v == 1;
}

bool gemm_streamk_active() { return streamk_enabled(); }

int launch_gemm_stream(const StreamGemmParams& p_in, int epi, int num_sms, cudaStream_t st) {
    // Measured (tools/stream_gemm_bench.py, profiles/stream_gemm_r2.md): the persistent stream-K kernel wins where there are
    // at least as many 128-row tiles as SMs (gate/up: 172, lm_head: 250) — every SM streams the same bytes through one deep
    // ring; with fewer tiles a CTA's whole range lies inside one tile, the hand-off of partials sits at the very end of the
    // kernel, and the cluster split-K kernel below (DSMEM reduce, two CTAs per SM) is faster.
    static int sk_all = -1;
    if (sk_all < 0) { const char* e = getenv("B200SHARD_STREAMK_ALL"); sk_all = (e != nullptr && e[0] == '1') ? 1 : 0; }
    if (streamk_enabled() && p_in.split == 0 && (sk_all == 1 || (p_in.N + kSwRows - 1) / kSwRows >= num_sms)) {
        const int r = launch_gemm_streamk(p_in, epi, num_sms, st);
        if (r != 0) snprintf(g_stream_err, sizeof(g_stream_err), "%s", gemm_streamk_last_error());
        return r;
    }
    StreamGemmParams p = p_in;
    if (p.M < 1 || p.M > kMaxStreamRows || p.K % 8 != 0 || p.K < kSwK || p.N < 1) {
        snprintf(g_stream_err, sizeof(g_stream_err), "gemm_stream: need 1 <= M <= %d, K %% 8 == 0, K >= 64 (M=%d N=%d K=%d)", kMaxStreamRows,
                 p.M, p.N, p.K);
        return -1;
    }
    if ((epi == EPI_QKV || epi == EPI_SWIGLU) && p.N % kSwRows != 0) {
        snprintf(g_stream_err, sizeof(g_stream_err), "gemm_stream: paired epilogues need N %% 128 == 0 (N=%d)", p.N);
        return -1;
    }
    if (p.split < 1) p.split = gemm_stream_split(p.N, p.K, num_sms);
    const int MP = p.M <= 16 ? 16 : 32;
    CUtensorMap mw, mx;
    if (tma_map_2d(&mw, p.W, p.N, p.K, kSwRows) != 0 || tma_map_2d(&mx, p.X, p.M, p.K, MP) != 0) {
        snprintf(g_stream_err, sizeof(g_stream_err), "gemm_stream: tensor map creation failed");
        return -1;
    }
    // ring depth: 5 stages of 18 KB for the 16-column operand (two CTAs of 101 KB still share an SM; same-box A/B against 4
    // stages: layer at 4 / 8 / 16 rows 89.0 / 94.0 / 102.7 -> 87.7 / 92.9 / 101.1 us, lm_head 43.7 -> 42.8 us), 4 stages of 20 KB
    // for the 32-column one.  B200SHARD_SW_STAGES=4 restores the shallower ring (A/B).
    static int deep = -1;
    if (deep < 0) { const char* e = getenv("B200SHARD_SW_STAGES"); deep = (e != nullptr && atoi(e) == 4) ? 0 : 1; }
#define B200_SW_CASE(E) (MP == 16 ? (deep ? launch_one<16, E, 5>(mw, mx, p, st) : launch_one<16, E, 4>(mw, mx, p, st)) : launch_one<32, E, 4>(mw, mx, p, st))
    switch (epi) {
        case EPI_STORE: return B200_SW_CASE(EPI_STORE);
        case EPI_RESID: return B200_SW_CASE(EPI_RESID);
        case EPI_QKV: return B200_SW_CASE(EPI_QKV);
        case EPI_SWIGLU: return B200_SW_CASE(EPI_SWIGLU);
    }
#undef B200_SW_CASE
    snprintf(g_stream_err, sizeof(g_stream_err), "gemm_stream: unknown epilogue %d", epi);
    return -1;
}

}  // namespace b200
