// Dense projection GEMMs of the shard path for more than a handful of token rows (prefill and
// batched decode): C[M,N] = A[M,K] * W[N,K]^T, bf16 operands, fp32 accumulation in tensor memory.
// Replaces the cuBLAS GEMMs under q/k/v/o_proj (transformers modeling_llama.py:262-264,287-288)
// and gate/up/down_proj (:182-184) that the reference reaches from utils/shard_loader.py:66-74.
//
// sm_100a design: one CTA per 128x128 output tile; warp 0 = TMA producer (cp.async.bulk.tensor,
// 128B-swizzled 128x64 tiles of A and W through a 6-stage mbarrier ring), warp 1 = single-thread
// tcgen05.mma issuer (UMMA 128x128x16, accumulator = 128 TMEM columns), warps 2..5 = epilogue
// (tcgen05.ld 32x32b, one accumulator row per thread) with the fused epilogues:
//   QKV_ROPE (RoPE + paged K/V append + hi/lo split of q), STORE_BF16, RESID_F32 (o_proj / down_proj + residual add, modeling_llama.py:325,331),
//   SWIGLU_BF16 (silu(gate)*up on the paired gate/up weight layout, :183), STORE_F32 (lm_head).
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "common.cuh"
#include "internal.h"
#include "tc_common.cuh"

namespace b200 {
using namespace tc;

constexpr int kBlkM = 128, kBlkN = 128, kBlkK = 64, kStages = 6;
constexpr int kGemmThreads = 192;
constexpr uint32_t kStageBytesA = kBlkM * kBlkK * 2, kStageBytesB = kBlkN * kBlkK * 2;
constexpr uint32_t kStageBytes = kStageBytesA + kStageBytesB;
constexpr size_t kGemmSmem = 1024 /*align slack*/ + (size_t)kStages * kStageBytes + 256 /*barriers*/;

// K-major operand tile, 128-byte swizzle, rows of 64 x 16-bit (= 128 B); 8-row groups are 1024 B apart
__device__ __forceinline__ uint64_t umma_desc(uint32_t smem_addr) { return umma_desc_sw128(smem_addr, 16, 1024); }
// kind::f16 instruction descriptor: D=f32, both operands K-major, N=128, M=128.  Operand format (a_format bits 7..9,
// b_format bits 10..12): 1 = bf16, 0 = fp16; the two must agree (a mixed fp16 x bf16 descriptor is an illegal
// instruction on sm_100a — measured).  The shard path runs fp16 x fp16: the bf16 model weights are held as fp16
// (exact for every bf16 value in fp16's normal range) so that activations can carry 10 mantissa bits — 8x less
// rounding than bf16 — into the tensor cores at no cost.
__device__ __forceinline__ uint32_t pack_f16(float a, float b) {
    a = fminf(fmaxf(a, -65504.f), 65504.f);
    b = fminf(fmaxf(b, -65504.f), 65504.f);
    __half2 t = __floats2half2_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&t);
}

// One accumulator row per thread: `trow` = TMEM address of this thread's lane at the first column of a 128-column
// logical tile, `n_blk` = index of that tile along N (one head for QKV_ROPE, one gate/up pair block for SWIGLU).
template <int EPI, int BN = kBlkN>   // BN: columns of the tile for the plain store / residual epilogues (the paired ones are 128 wide)
__device__ __forceinline__ void gemm_epilogue(const GemmParams& p, const uint32_t trow, const int row, const int n_blk) {
    const bool row_ok = row < p.M;
    if (EPI == GEMM_SWIGLU_BF16) {
        __nv_bfloat16* out = reinterpret_cast<__nv_bfloat16*>(p.out);
        const int No = p.N / 2;
#pragma unroll 1
        for (int c = 0; c < 2; ++c) {
            uint32_t g[32], u[32];
            tmem_ld32(trow + c * 32, g);
            tmem_ld32(trow + 64 + c * 32, u);
            const int col0 = n_blk * 64 + c * 32;
            if (row_ok) {
#pragma unroll
                for (int v = 0; v < 4; ++v) {
                    if (col0 + v * 8 < No) {
                        uint32_t w[4];
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            const float g0 = __uint_as_float(g[v * 8 + 2 * e]), g1 = __uint_as_float(g[v * 8 + 2 * e + 1]);
                            const float u0 = __uint_as_float(u[v * 8 + 2 * e]), u1 = __uint_as_float(u[v * 8 + 2 * e + 1]);
                            w[e] = p.out_f16 ? pack_f16(silu(g0) * u0, silu(g1) * u1) : pack_bf16(silu(g0) * u0, silu(g1) * u1);
                        }
                        *reinterpret_cast<uint4*>(out + (size_t)row * No + col0 + v * 8) = make_uint4(w[0], w[1], w[2], w[3]);
                    }
                }
            }
        }
    } else if (EPI == GEMM_QKV_ROPE) {
        // one 128-column tile = one head: apply_rotary_pos_emb (modeling_llama.py:146-168) in fp32 on the
        // accumulator, then q -> (hi, lo) bf16 pair, k / v -> bf16 straight into the KV pages
        // (DynamicCache.update, cache_utils.py:102-121, as a paged write).  A single rounding per value.
        const int head = n_blk;
        const bool is_q = head < p.n_heads, is_k = !is_q && head < p.n_heads + p.n_kv;
#pragma unroll 1
        for (int c = 0; c < 2; ++c) {
            uint32_t a[32], b[32];
            tmem_ld32(trow + c * 32, a);
            tmem_ld32(trow + 64 + c * 32, b);
            if (row_ok) {
                const int j0 = c * 32;
                float lo[32], hi[32];
                if (is_q || is_k) {
                    const float4* cs4 = reinterpret_cast<const float4*>(p.cos + (size_t)row * kHalfDim + j0);
                    const float4* sn4 = reinterpret_cast<const float4*>(p.sin + (size_t)row * kHalfDim + j0);
#pragma unroll
                    for (int v = 0; v < 8; ++v) {
                        const float4 cs = cs4[v], sn = sn4[v];
                        const float cc[4] = {cs.x, cs.y, cs.z, cs.w}, ss[4] = {sn.x, sn.y, sn.z, sn.w};
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            const float x = __uint_as_float(a[v * 4 + e]), y = __uint_as_float(b[v * 4 + e]);
                            lo[v * 4 + e] = x * cc[e] - y * ss[e];
                            hi[v * 4 + e] = y * cc[e] + x * ss[e];
                        }
                    }
                } else {
#pragma unroll
                    for (int e = 0; e < 32; ++e) { lo[e] = __uint_as_float(a[e]); hi[e] = __uint_as_float(b[e]); }
                }
                __nv_bfloat16* dst;
                if (is_q) dst = p.q_hi + ((size_t)row * p.n_heads + head) * kHeadDim + j0;
                else if (is_k) dst = p.k_pool + p.kv_off[row] + (int64_t)(head - p.n_heads) * p.head_stride + j0;
                else dst = p.v_pool + p.kv_off[row] + (int64_t)(head - p.n_heads - p.n_kv) * p.head_stride + j0;
#pragma unroll
                for (int v = 0; v < 4; ++v) {
                    *reinterpret_cast<uint4*>(dst + v * 8) = make_uint4(pack_bf16(lo[v * 8 + 0], lo[v * 8 + 1]), pack_bf16(lo[v * 8 + 2], lo[v * 8 + 3]),
                                                                        pack_bf16(lo[v * 8 + 4], lo[v * 8 + 5]), pack_bf16(lo[v * 8 + 6], lo[v * 8 + 7]));
                    *reinterpret_cast<uint4*>(dst + kHalfDim + v * 8) = make_uint4(pack_bf16(hi[v * 8 + 0], hi[v * 8 + 1]), pack_bf16(hi[v * 8 + 2], hi[v * 8 + 3]),
                                                                                   pack_bf16(hi[v * 8 + 4], hi[v * 8 + 5]), pack_bf16(hi[v * 8 + 6], hi[v * 8 + 7]));
                }
                if (is_q) {  // residual of the bf16 rounding, so that attention sees q with ~16 mantissa bits
                    __nv_bfloat16* dlo = p.q_lo + ((size_t)row * p.n_heads + head) * kHeadDim + j0;
#pragma unroll
                    for (int e = 0; e < 32; ++e) {
                        lo[e] -= __bfloat162float(__float2bfloat16_rn(lo[e]));
                        hi[e] -= __bfloat162float(__float2bfloat16_rn(hi[e]));
                    }
#pragma unroll
                    for (int v = 0; v < 4; ++v) {
                        *reinterpret_cast<uint4*>(dlo + v * 8) = make_uint4(pack_bf16(lo[v * 8 + 0], lo[v * 8 + 1]), pack_bf16(lo[v * 8 + 2], lo[v * 8 + 3]),
                                                                            pack_bf16(lo[v * 8 + 4], lo[v * 8 + 5]), pack_bf16(lo[v * 8 + 6], lo[v * 8 + 7]));
                        *reinterpret_cast<uint4*>(dlo + kHalfDim + v * 8) = make_uint4(pack_bf16(hi[v * 8 + 0], hi[v * 8 + 1]), pack_bf16(hi[v * 8 + 2], hi[v * 8 + 3]),
                                                                                       pack_bf16(hi[v * 8 + 4], hi[v * 8 + 5]), pack_bf16(hi[v * 8 + 6], hi[v * 8 + 7]));
                    }
                }
            }
        }
    } else {
#pragma unroll 1
        for (int c = 0; c < BN / 32; ++c) {
            uint32_t r[32];
            tmem_ld32(trow + c * 32, r);
            const int col0 = n_blk * BN + c * 32;
            if (row_ok) {
                if (EPI == GEMM_STORE_BF16) {
                    __nv_bfloat16* out = reinterpret_cast<__nv_bfloat16*>(p.out) + (size_t)row * p.N + col0;
#pragma unroll
                    for (int v = 0; v < 4; ++v)
                        if (col0 + v * 8 < p.N)
                            *reinterpret_cast<uint4*>(out + v * 8) = make_uint4(
                                pack_bf16(__uint_as_float(r[v * 8 + 0]), __uint_as_float(r[v * 8 + 1])),
                                pack_bf16(__uint_as_float(r[v * 8 + 2]), __uint_as_float(r[v * 8 + 3])),
                                pack_bf16(__uint_as_float(r[v * 8 + 4]), __uint_as_float(r[v * 8 + 5])),
                                pack_bf16(__uint_as_float(r[v * 8 + 6]), __uint_as_float(r[v * 8 + 7])));
                } else {
                    float* out = reinterpret_cast<float*>(p.out) + (size_t)row * p.N + col0;
#pragma unroll
                    for (int v = 0; v < 8; ++v)
                        if (col0 + v * 4 < p.N) {
                            float4 acc = make_float4(__uint_as_float(r[v * 4 + 0]), __uint_as_float(r[v * 4 + 1]),
                                                     __uint_as_float(r[v * 4 + 2]), __uint_as_float(r[v * 4 + 3]));
                            if (EPI == GEMM_RESID_F32) {
                                const float4 o = *reinterpret_cast<const float4*>(out + v * 4);
                                acc.x += o.x; acc.y += o.y; acc.z += o.z; acc.w += o.w;
                            }
                            *reinterpret_cast<float4*>(out + v * 4) = acc;
                        }
                }
            }
        }
    }
}

// BN x NSTG: 128 columns behind a 6-stage ring (one CTA per SM), or 64 columns behind a 4-stage ring — 97 KB, two CTAs per SM —
// for the GEMMs that have too few 128 x 128 tiles to give every SM more than one (o_proj / down_proj of a 512-token prompt:
// 128 tiles for 148 SMs, nothing to overlap a tile's ramp and epilogue with; 256 half-width tiles, two resident per SM, do).
template <int EPI, int BN, int NSTG>
__global__ void __launch_bounds__(kGemmThreads, (BN == 128 ? 1 : 2))
gemm_tcgen05_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b,
                    const GemmParams p) {
    constexpr uint32_t kStB = BN * kBlkK * 2, kSt = kStageBytesA + kStB;   // bytes per stage: W tile, both tiles
    extern __shared__ uint8_t smem_raw[];
    const uint32_t raw = s_u32(smem_raw);
    const uint32_t base = (raw + 1023u) & ~1023u;  // SWIZZLE_128B tiles need 1024-byte alignment
    uint8_t* gen_base = smem_raw + (base - raw);
    const uint32_t bars = base + NSTG * kSt;
    const uint32_t full0 = bars, empty0 = bars + 8 * NSTG, tmem_full = bars + 16 * NSTG;
    volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(gen_base + NSTG * kSt + 16 * NSTG + 8);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int n_blk = blockIdx.x, m_blk = blockIdx.y;
    const int num_k = (p.K + kBlkK - 1) / kBlkK;
    const uint32_t idesc = (1u << 4) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(kBlkM >> 4) << 24) | (p.ab_f16 ? 0u : ((1u << 7) | (1u << 10)));

    if (warp == 0 && lane == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(&tma_a) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&tma_b) : "memory");
        for (int i = 0; i < NSTG; ++i) { mbar_init(full0 + 8 * i, 1); mbar_init(empty0 + 8 * i, 1); }
        mbar_init(tmem_full, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(s_u32((const void*)tmem_slot)), "n"(BN) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        if (lane == 0) {  // ===== TMA producer
            // programmatic dependent launch: the weight tiles of the first trip round the ring do not depend on the previous
            // kernel and are requested before griddepcontrol.wait; the activation tiles follow it
            const int pre = num_k < NSTG ? num_k : NSTG;
            for (int kb = 0; kb < pre; ++kb) {
                mbar_expect_tx(full0 + 8 * kb, kSt);
                tma_load_2d(&tma_b, full0 + 8 * kb, base + kb * kSt + kStageBytesA, kb * kBlkK, n_blk * BN);
            }
            asm volatile("griddepcontrol.wait;" ::: "memory");
            asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
            for (int kb = 0; kb < pre; ++kb) tma_load_2d(&tma_a, full0 + 8 * kb, base + kb * kSt, kb * kBlkK, m_blk * kBlkM);
            for (int kb = pre; kb < num_k; ++kb) {
                const int st = kb % NSTG;
                const uint32_t ph = (uint32_t)(kb / NSTG) & 1u;
                mbar_wait(empty0 + 8 * st, ph ^ 1u);
                mbar_expect_tx(full0 + 8 * st, kSt);
                const uint32_t sa = base + st * kSt, sb = sa + kStageBytesA;
                tma_load_2d(&tma_a, full0 + 8 * st, sa, kb * kBlkK, m_blk * kBlkM);
                tma_load_2d(&tma_b, full0 + 8 * st, sb, kb * kBlkK, n_blk * BN);
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {  // ===== MMA issuer (one thread drives the tensor core for the CTA)
            for (int kb = 0; kb < num_k; ++kb) {
                const int st = kb % NSTG;
                const uint32_t ph = (uint32_t)(kb / NSTG) & 1u;
                mbar_wait(full0 + 8 * st, ph);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint32_t sa = base + st * kSt, sb = sa + kStageBytesA;
#pragma unroll
                for (int k = 0; k < kBlkK / 16; ++k)
                    umma_f16(tmem_base, umma_desc(sa + k * 32), umma_desc(sb + k * 32), idesc, (kb > 0 || k > 0) ? 1u : 0u);
                umma_commit(empty0 + 8 * st);  // frees the smem slot when these MMAs retire
            }
            umma_commit(tmem_full);            // accumulator complete
        }
    } else {  // ===== epilogue: warp w may only touch TMEM lanes 32*(w%4) .. +31
        asm volatile("griddepcontrol.wait;" ::: "memory");   // the residual stream / RoPE table / KV pages belong to earlier kernels
        mbar_wait(tmem_full, 0);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const int quad = warp & 3;
        const int row = m_blk * kBlkM + quad * 32 + lane;
        const uint32_t trow = tmem_base + ((uint32_t)(quad * 32) << 16);
        gemm_epilogue<EPI, BN>(p, trow, row, n_blk);
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    }
    __syncthreads();
    if (warp == 1) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(BN) : "memory");
    }
}

// ================================================================================================================
// CTA-pair variant for M > 128: one 256 x 256 output tile per pair of CTAs (cluster of 2 on one TPC),
// tcgen05.mma.cta_group::2 with UMMA M = 256.  Each CTA stages its own 128 rows of A and only HALF of the 256-row B
// tile (16 KB + 16 KB per k-step for 128 x 256 outputs per CTA): twice the arithmetic intensity of the single-CTA
// kernel above, which ncu shows bound by L2 -> shared-memory traffic (profiles/prefill_gemm_r1_summary.md).
// Protocol: both CTAs run a TMA producer; every TMA signals the LEADER's (cluster rank 0) full barrier; the leader's
// single MMA thread issues for the pair and commits with a multicast arrive onto both CTAs' empty / tmem_full
// barriers; each CTA drains its own 128 TMEM lanes x 256 columns through the shared epilogue (two 128-column halves).
constexpr int kBlkN2 = 256;
constexpr uint32_t kStageBytesB2 = (kBlkN2 / 2) * kBlkK * 2;         // this CTA's half of the B tile
constexpr uint32_t kStageBytes2 = kStageBytesA + kStageBytesB2;      // 32 KB per CTA per stage

constexpr uint32_t kPeerMask = 0xFEFFFFFFu;                          // shared::cluster address of the same offset in CTA rank 0

__device__ __forceinline__ void tma_load_2d_pair(const CUtensorMap* map, uint32_t leader_bar, uint32_t dst, int x, int y) {
    asm volatile(
        "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(dst), "l"(map), "r"(leader_bar), "r"(x), "r"(y) : "memory");
}
__device__ __forceinline__ void umma_f16_pair(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t accum) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(da), "l"(db), "r"(idesc), "r"(accum) : "memory");
}
__device__ __forceinline__ void umma_commit_pair(uint32_t bar) {  // arrives on `bar` in BOTH CTAs of the pair
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                 ::"r"(bar), "h"((uint16_t)3) : "memory");
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}

// Remote arrive on the barrier at the same shared-memory offset in the leader CTA (cluster rank 0).
__device__ __forceinline__ void mbar_arrive_leader(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(bar & kPeerMask) : "memory");
}

// Persistent: one CTA pair per SM pair walks over the 256 x 256 tiles (m fastest, so that concurrently running pairs
// share the weight tile in L2).  The accumulator is double buffered in tensor memory (2 x 256 columns): the epilogue of
// tile i (TMEM -> registers -> global, with the fused RoPE / SwiGLU / residual work) overlaps the MMAs of tile i+1.
// BN2: columns of the pair's tile — 256, or 128 (with NSTG2 = 8 stages of 24 KB) for the residual GEMMs of mid-sized prompts, where 256 x 256
// tiles number fewer than the SM pairs and 128 x 128 single-CTA tiles move a third more bytes through shared memory per flop.
template <int EPI, int BN2, int NSTG2>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kGemmThreads, 1)
gemm_pair_tcgen05_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b,
                         const GemmParams p) {
    constexpr uint32_t kStB2 = (BN2 / 2) * kBlkK * 2, kSt2 = kStageBytesA + kStB2;   // this CTA's half of the B tile; both tiles
    extern __shared__ uint8_t smem_raw[];
    const uint32_t raw = s_u32(smem_raw);
    const uint32_t base = (raw + 1023u) & ~1023u;
    uint8_t* gen_base = smem_raw + (base - raw);
    const uint32_t bars = base + NSTG2 * kSt2;
    const uint32_t full0 = bars, empty0 = bars + 8 * NSTG2, tfull0 = bars + 16 * NSTG2, tempty0 = tfull0 + 16;
    volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(gen_base + NSTG2 * kSt2 + 16 * NSTG2 + 32);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    uint32_t cta_rank;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(cta_rank));
    const bool leader = cta_rank == 0;
    const int pair_id = blockIdx.x >> 1, n_pairs = gridDim.x >> 1;
    const int n_m = (p.M + 255) / 256, n_n = (p.N + BN2 - 1) / BN2;
    const int n_tiles = n_m * n_n;
    const int num_k = (p.K + kBlkK - 1) / kBlkK;
    const uint32_t idesc = (1u << 4) | ((uint32_t)(BN2 >> 3) << 17) | ((uint32_t)(256 >> 4) << 24) | (p.ab_f16 ? 0u : ((1u << 7) | (1u << 10)));

    if (warp == 0 && lane == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(&tma_a) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&tma_b) : "memory");
        for (int i = 0; i < NSTG2; ++i) { mbar_init(full0 + 8 * i, 1); mbar_init(empty0 + 8 * i, 1); }
        for (int i = 0; i < 2; ++i) { mbar_init(tfull0 + 8 * i, 1); mbar_init(tempty0 + 8 * i, 8); }  // 4 epilogue warps x 2 CTAs
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {  // the same warp of both CTAs allocates for the pair: all 512 columns (two accumulators)
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(s_u32((const void*)tmem_slot)), "n"(512) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    cluster_sync_all();  // barriers of both CTAs are initialised before anybody signals across the pair
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        if (lane == 0) {  // ===== TMA producer (both CTAs): own A rows + own half of B, completion counted on the leader
            asm volatile("griddepcontrol.wait;" ::: "memory");   // A comes from the previous kernel (the prologue above overlapped its tail)
            asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
            uint32_t it = 0;
            for (int t = pair_id; t < n_tiles; t += n_pairs) {
                const int m_row0 = (t % n_m) * 256 + (int)cta_rank * kBlkM;
                const int n_row0 = (t / n_m) * BN2 + (int)cta_rank * (BN2 / 2);
                for (int kb = 0; kb < num_k; ++kb, ++it) {
                    const uint32_t st = it % NSTG2, ph = (it / NSTG2) & 1u;
                    mbar_wait(empty0 + 8 * st, ph ^ 1u);
                    if (leader) mbar_expect_tx(full0 + 8 * st, 2 * kSt2);
                    const uint32_t sa = base + st * kSt2, sb = sa + kStageBytesA;
                    const uint32_t lbar = (full0 + 8 * st) & kPeerMask;
                    tma_load_2d_pair(&tma_a, lbar, sa, kb * kBlkK, m_row0);
                    tma_load_2d_pair(&tma_b, lbar, sb, kb * kBlkK, n_row0);
                }
            }
        }
    } else if (warp == 1) {
        if (leader && lane == 0) {  // ===== MMA issuer: one thread for the pair
            uint32_t it = 0, ti = 0;
            for (int t = pair_id; t < n_tiles; t += n_pairs, ++ti) {
                const uint32_t buf = ti & 1u;
                mbar_wait(tempty0 + 8 * buf, ((ti >> 1) & 1u) ^ 1u);  // both CTAs' epilogues have drained this accumulator
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint32_t acc = tmem_base + buf * BN2;
                for (int kb = 0; kb < num_k; ++kb, ++it) {
                    const uint32_t st = it % NSTG2, ph = (it / NSTG2) & 1u;
                    mbar_wait(full0 + 8 * st, ph);
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                    const uint32_t sa = base + st * kSt2, sb = sa + kStageBytesA;
#pragma unroll
                    for (int k = 0; k < kBlkK / 16; ++k)
                        umma_f16_pair(acc, umma_desc(sa + k * 32), umma_desc(sb + k * 32), idesc, (kb > 0 || k > 0) ? 1u : 0u);
                    umma_commit_pair(empty0 + 8 * st);
                }
                umma_commit_pair(tfull0 + 8 * buf);
            }
        }
    } else {  // ===== epilogue (both CTAs): own 128 lanes, 256 columns = two logical 128-column tiles
        asm volatile("griddepcontrol.wait;" ::: "memory");
        const int quad = warp & 3;
        uint32_t ti = 0;
        for (int t = pair_id; t < n_tiles; t += n_pairs, ++ti) {
            const uint32_t buf = ti & 1u;
            const int m_row0 = (t % n_m) * 256 + (int)cta_rank * kBlkM, n_blk = t / n_m;
            mbar_wait(tfull0 + 8 * buf, (ti >> 1) & 1u);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const int row = m_row0 + quad * 32 + lane;
            const uint32_t trow = tmem_base + buf * BN2 + ((uint32_t)(quad * 32) << 16);
#pragma unroll 1
            for (int sub = 0; sub < BN2 / kBlkN; ++sub)
                if ((n_blk * (BN2 / kBlkN) + sub) * kBlkN < p.N) gemm_epilogue<EPI>(p, trow + sub * kBlkN, row, n_blk * (BN2 / kBlkN) + sub);
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            __syncwarp();
            if (lane == 0) mbar_arrive_leader(tempty0 + 8 * buf);  // this warp no longer reads accumulator `buf`
        }
    }
    __syncthreads();
    cluster_sync_all();  // neither CTA may free TMEM / leave while the other still reads its shared memory or TMEM
    if (warp == 1) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(512) : "memory");
    }
}

// ---- host side: tensor maps through the driver entry point (no link-time libcuda dependency)
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn g_encode = nullptr;
static thread_local char g_gemm_err[256] = "";
const char* gemm_error() { return g_gemm_err; }

static int make_map(CUtensorMap* map, const void* ptr, int rows, int K, int box_rows) {
    if (g_encode == nullptr) {
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult q;
        cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
        if (e != cudaSuccess || fn == nullptr) {
            snprintf(g_gemm_err, sizeof(g_gemm_err), "cuTensorMapEncodeTiled entry point unavailable: %s", cudaGetErrorString(e));
            return -1;
        }
        g_encode = reinterpret_cast<EncodeTiledFn>(fn);
    }
    const cuuint64_t dims[2] = {(cuuint64_t)K, (cuuint64_t)rows};
    const cuuint64_t strides[1] = {(cuuint64_t)K * 2};
    const cuuint32_t box[2] = {(cuuint32_t)kBlkK, (cuuint32_t)box_rows};
    const cuuint32_t estr[2] = {1, 1};
    CUresult r = g_encode(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), dims, strides, box, estr,
                          CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        snprintf(g_gemm_err, sizeof(g_gemm_err), "cuTensorMapEncodeTiled failed (%d) rows=%d K=%d ptr=%p", (int)r, rows, K, ptr);
        return -1;
    }
    return 0;
}

// 2-D row-major 16-bit tensor [rows, cols], box = 64 columns x box_rows rows, 128-byte swizzle
int tma_map_2d(void* map_out, const void* ptr, int rows, int cols, int box_rows) {
    return make_map(reinterpret_cast<CUtensorMap*>(map_out), ptr, rows, cols, box_rows);
}

// launch with programmatic stream serialization (B200SHARD_PDL=0 turns it off for A/B runs)
template <typename K>
static void launch_pdl(K kfn, dim3 grid, size_t smem, cudaStream_t st, const CUtensorMap& ma, const CUtensorMap& mb, const GemmParams& p) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = grid;
    cfg.blockDim = dim3(kGemmThreads);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    static int pdl = -1;
    if (pdl < 0) { const char* e = getenv("B200SHARD_PDL"); pdl = (e != nullptr && e[0] == '0') ? 0 : 1; }
    cfg.numAttrs = pdl;
    cudaLaunchKernelEx(&cfg, kfn, ma, mb, p);
}

int launch_gemm(const GemmParams& p, cudaStream_t st) {
    if (p.K % 8 != 0 || p.N % 8 != 0 || p.M <= 0) {
        snprintf(g_gemm_err, sizeof(g_gemm_err), "gemm: need K%%8==0, N%%8==0, M>0 (M=%d N=%d K=%d)", p.M, p.N, p.K);
        return -1;
    }
    if (p.epi == GEMM_QKV_ROPE && (p.N != (p.n_heads + 2 * p.n_kv) * 128 || p.q_hi == nullptr || p.q_lo == nullptr)) {
        snprintf(g_gemm_err, sizeof(g_gemm_err), "gemm qkv+rope: N must be (n_heads + 2 n_kv) * 128 (N=%d)", p.N);
        return -1;
    }
    if (p.epi == GEMM_SWIGLU_BF16 && p.N % 128 != 0) {
        snprintf(g_gemm_err, sizeof(g_gemm_err), "gemm swiglu: packed N must be a multiple of 128 (N=%d)", p.N);
        return -1;
    }
    CUtensorMap ma, mb;
    if (make_map(&ma, p.A, p.M, p.K, kBlkM) != 0) return -1;
    if (make_map(&mb, p.W, p.N, p.K, kBlkN) != 0) return -1;
    static int pair_mode = -1;
    if (pair_mode < 0) {
        const char* e = getenv("B200SHARD_GEMM_PAIR");
        pair_mode = (e != nullptr && e[0] == '0') ? 0 : 1;
    }
    // CTA pairs (256 x 256 tiles, half the L2 -> shared-memory traffic per flop) once there are enough tiles to occupy
    // most SM pairs; small problems keep the finer 128 x 128 tiling (measured: profiles/gemm_bench_r1.md)
    // Residual / plain-store GEMMs whose 256 x 256 tiles would not fill the SM pairs: 256 x 128 pair tiles, where there are more
    // of those than SM pairs so that the persistent pairs overlap one tile's epilogue with the next tile's main loop (513..768
    // rows for N = 4096: per layer 406 -> 384 us at 768 tokens; with a single tile per pair, 512 rows, it loses 264 -> 268 us,
    // same box).  B200SHARD_GEMM_PAIR_N128=0 keeps the single-CTA kernels there (A/B).
    static int pair_n128 = -1;
    if (pair_n128 < 0) { const char* e = getenv("B200SHARD_GEMM_PAIR_N128"); pair_n128 = (e != nullptr && e[0] == '0') ? 0 : 1; }
    const int64_t tiles256 = (int64_t)((p.M + 255) / 256) * ((p.N + kBlkN2 - 1) / kBlkN2);
    if (pair_mode == 1 && pair_n128 == 1 && p.M > kBlkM && tiles256 < 64 && p.N % 128 == 0 && 2 * tiles256 > 74 &&
        (p.epi == GEMM_RESID_F32 || p.epi == GEMM_STORE_F32 || p.epi == GEMM_STORE_BF16)) {
        int n_sms = 148, dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&n_sms, cudaDevAttrMultiProcessorCount, dev);
        const int64_t tiles = (int64_t)((p.M + 255) / 256) * (p.N / 128);
        const int pairs = (int)(tiles < n_sms / 2 ? tiles : n_sms / 2);
        CUtensorMap mb64;
        if (make_map(&mb64, p.W, p.N, p.K, 64) != 0) return -1;
        constexpr size_t smem3 = 1024 + (size_t)8 * (kStageBytesA + 64 * kBlkK * 2) + 256;
#define B200_GEMM3_LAUNCH(E)                                                                                \
    {                                                                                                       \
        auto kfn = gemm_pair_tcgen05_kernel<E, 128, 8>;                                                     \
        cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem3);                 \
        launch_pdl(kfn, dim3(2 * pairs), smem3, st, ma, mb64, p);                                           \
    }
        if (p.epi == GEMM_RESID_F32) B200_GEMM3_LAUNCH(GEMM_RESID_F32)
        else if (p.epi == GEMM_STORE_F32) B200_GEMM3_LAUNCH(GEMM_STORE_F32)
        else B200_GEMM3_LAUNCH(GEMM_STORE_BF16)
#undef B200_GEMM3_LAUNCH
        return 0;
    }
    if (pair_mode == 1 && p.M > kBlkM && tiles256 >= 64) {
        int n_sms = 148, dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&n_sms, cudaDevAttrMultiProcessorCount, dev);
        const int64_t tiles2 = (int64_t)((p.M + 255) / 256) * ((p.N + kBlkN2 - 1) / kBlkN2);
        const int pairs = (int)(tiles2 < n_sms / 2 ? tiles2 : n_sms / 2);
        dim3 grid2(2 * pairs);
        const size_t smem2 = 1024 + (size_t)kStages * kStageBytes2 + 256;
#define B200_GEMM2_LAUNCH(E)                                                                                \
    {                                                                                                       \
        auto kfn = gemm_pair_tcgen05_kernel<E, kBlkN2, kStages>;                                                             \
        cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem2);                 \
        launch_pdl(kfn, grid2, smem2, st, ma, mb, p);                                                       \
    }
        switch (p.epi) {
            case GEMM_STORE_BF16: B200_GEMM2_LAUNCH(GEMM_STORE_BF16) break;
            case GEMM_RESID_F32: B200_GEMM2_LAUNCH(GEMM_RESID_F32) break;
            case GEMM_SWIGLU_BF16: B200_GEMM2_LAUNCH(GEMM_SWIGLU_BF16) break;
            case GEMM_QKV_ROPE: B200_GEMM2_LAUNCH(GEMM_QKV_ROPE) break;
            default: B200_GEMM2_LAUNCH(GEMM_STORE_F32) break;
        }
#undef B200_GEMM2_LAUNCH
        return 0;
    }
    // Few 128 x 128 tiles (less than 1.5 per SM) and an epilogue that does not pair columns: half-width tiles, two CTAs per SM.
    static int n64 = -1;   // B200SHARD_GEMM_N64=0: always 128-column tiles (A/B)
    if (n64 < 0) { const char* e = getenv("B200SHARD_GEMM_N64"); n64 = (e != nullptr && e[0] == '0') ? 0 : 1; }
    const int64_t tiles128 = (int64_t)((p.N + kBlkN - 1) / kBlkN) * ((p.M + kBlkM - 1) / kBlkM);
    const bool plain = p.epi == GEMM_STORE_BF16 || p.epi == GEMM_RESID_F32 || p.epi == GEMM_STORE_F32;
    if (n64 == 1 && plain && tiles128 < 222 && p.N % 64 == 0) {
        CUtensorMap mb64;
        if (make_map(&mb64, p.W, p.N, p.K, 64) != 0) return -1;
        dim3 grid64((p.N + 63) / 64, (p.M + kBlkM - 1) / kBlkM);
        constexpr size_t smem64 = 1024 + (size_t)4 * (kStageBytesA + 64 * kBlkK * 2) + 256;
#define B200_GEMM64_LAUNCH(E)                                                                               \
    {                                                                                                       \
        auto kfn = gemm_tcgen05_kernel<E, 64, 4>;                                                           \
        cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem64);                \
        launch_pdl(kfn, grid64, smem64, st, ma, mb64, p);                                                   \
    }
        if (p.epi == GEMM_STORE_BF16) B200_GEMM64_LAUNCH(GEMM_STORE_BF16)
        else if (p.epi == GEMM_RESID_F32) B200_GEMM64_LAUNCH(GEMM_RESID_F32)
        else B200_GEMM64_LAUNCH(GEMM_STORE_F32)
#undef B200_GEMM64_LAUNCH
        return 0;
    }
    dim3 grid((p.N + kBlkN - 1) / kBlkN, (p.M + kBlkM - 1) / kBlkM);
#define B200_GEMM_LAUNCH(E)                                                                                 \
    {                                                                                                       \
        auto kfn = gemm_tcgen05_kernel<E, kBlkN, kStages>;                                                  \
        cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kGemmSmem);             \
        launch_pdl(kfn, grid, kGemmSmem, st, ma, mb, p);                                                    \
    }
    switch (p.epi) {
        case GEMM_STORE_BF16: B200_GEMM_LAUNCH(GEMM_STORE_BF16) break;
        case GEMM_RESID_F32: B200_GEMM_LAUNCH(GEMM_RESID_F32) break;
        case GEMM_SWIGLU_BF16: B200_GEMM_LAUNCH(GEMM_SWIGLU_BF16) break;
        case GEMM_QKV_ROPE: B200_GEMM_LAUNCH(GEMM_QKV_ROPE) break;
        default: B200_GEMM_LAUNCH(GEMM_STORE_F32) break;
    }
#undef B200_GEMM_LAUNCH
    return 0;
}

}  // namespace b200
