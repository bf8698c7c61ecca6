// Prefill attention on the 5th-generation tensor cores: softmax_fp32(q k^T / sqrt(d) [+ causal mask]) v over the paged
// KV cache (eager_attention_forward, transformers modeling_llama.py:199-221, reached from utils/shard_loader.py:66-74;
// the reference passes no mask, utils/node_worker.py:253-258, so causal == 0 is the parity mode).
//
// One CTA = 128 query rows of one (batch row, head).  Warp-specialised:
//   warp 0     TMA producer: Q (hi and lo halves) once, then 128-key K/V tiles (two 64-token pages, looked up in the
//              block table) through a 2-stage mbarrier ring; every tile is 128-byte swizzled.
//   warp 1     single-thread tcgen05.mma issuer: S = Q K^T (UMMA 128x128x16, q as a hi+lo bf16 pair -> two passes)
//              into tensor memory, then O_tile = P V with V read as an MN-major operand straight from its natural
//              [token][dim] page layout.
//   warps 2-5  softmax: one query row per thread, S read from TMEM (tcgen05.ld), fp32 online softmax (running max and
//              sum, ex2 on the SFU), P written to shared memory as the bf16 A operand of the second MMA.
//   warps 6-9  accumulate the output row (128 fp32 registers per thread) from O_tile with the rescale factor the
//              softmax warps publish — one tile behind the softmax, so neither group waits for the other's MMA.
// The scores live in two TMEM buffers: S of tile j+2 is issued as soon as the softmax has consumed tile j, and the K
// and V rings are released independently (K after its scores, V after its P V product).
// The [S,T] score matrix of the reference never exists; K and V are read once per 128 queries.
#include <stdio.h>
#include <stdlib.h>

#include "common.cuh"
#include "internal.h"
#include "tc_common.cuh"

namespace b200 {
using namespace tc;

__device__ __forceinline__ float ex2_fast(float x) {  // 2^x on the SFU; ex2(-inf) = 0
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

constexpr int kQT = 128, kKT = 128;              // queries / keys per tile
constexpr int kTcThreads = 320;
constexpr uint32_t kHalfTile = 128 * 128;        // 128 rows x 64 16-bit elements = 16 KB (one 128B-swizzled half)
constexpr uint32_t kOffQhi = 0, kOffQlo = 2 * kHalfTile, kOffK = 4 * kHalfTile, kOffV = 8 * kHalfTile, kOffP = 12 * kHalfTile;
constexpr uint32_t kOffBar = 14 * kHalfTile;     // barriers (128 B) + alpha[3][128]
constexpr size_t kTcSmem = 1024 + 14 * (size_t)kHalfTile + 128 + 3 * 128 * sizeof(float);

__global__ void __launch_bounds__(kTcThreads, 1)
attn_prefill_tc_kernel(const __grid_constant__ CUtensorMap map_qhi, const __grid_constant__ CUtensorMap map_qlo,
                       const __grid_constant__ CUtensorMap map_k, const __grid_constant__ CUtensorMap map_v,
                       const AttnPrefillParams p) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t raw = s_u32(smem_raw);
    const uint32_t base = (raw + 1023u) & ~1023u;
    uint8_t* gbase = smem_raw + (base - raw);
    const uint32_t bars = base + kOffBar;
    const uint32_t q_full = bars, k_full0 = bars + 8, k_empty0 = bars + 24, v_full0 = bars + 40, v_empty0 = bars + 56,
                   s_full0 = bars + 72, p_full = bars + 88, o_full = bars + 96, o_empty = bars + 104;
    volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(gbase + kOffBar + 112);
    float* alpha_s = reinterpret_cast<float*>(gbase + kOffBar + 128);   // [3][128]: rescale factor of tile j in slot j % 3

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int qt = blockIdx.x, h = blockIdx.y, b = blockIdx.z;
    const int q0 = qt * kQT;
    const int base_len = p.call_base[b];
    const int T = base_len + p.S;
    const int slot = p.slots.v[b];
    const int kvh = h / (p.n_heads / p.n_kv);
    const int32_t* bt = p.block_table + (size_t)slot * p.max_pages_per_slot;
    int n_kt = (T + kKT - 1) / kKT;
    if (p.causal) n_kt = (min(T, base_len + q0 + kQT) + kKT - 1) / kKT;

    if (warp == 0 && lane == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_qhi) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_k) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_v) : "memory");
        mbar_init(q_full, 1);
        for (int i = 0; i < 2; ++i) {
            mbar_init(k_full0 + 8 * i, 1); mbar_init(k_empty0 + 8 * i, 1);
            mbar_init(v_full0 + 8 * i, 1); mbar_init(v_empty0 + 8 * i, 1);
            mbar_init(s_full0 + 8 * i, 1);
        }
        mbar_init(p_full, 4);   // one arrive per softmax warp
        mbar_init(o_full, 1);
        mbar_init(o_empty, 4);  // one arrive per accumulating warp
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {  // two score buffers (2 x 128 columns) + one output tile (128 columns); 512 is the next power of two
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(s_u32((const void*)tmem_slot)), "n"(512) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    fence_before();
    __syncthreads();
    fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const uint32_t tmem_o = tmem_base + 256;

    if (warp == 0) {
        if (lane == 0) {  // ===== TMA producer: K and V rings are independent (a K stage frees up as soon as its scores are issued)
            asm volatile("griddepcontrol.wait;" ::: "memory");   // q and the new K/V rows come from the QKV projection just before
            asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
            const int qrow = b * p.S + q0;
            mbar_expect_tx(q_full, 4 * kHalfTile);
            for (int half = 0; half < 2; ++half) {
                tma_load_2d(&map_qhi, q_full, base + kOffQhi + half * kHalfTile, h * kHeadDim + half * 64, qrow);
                tma_load_2d(&map_qlo, q_full, base + kOffQlo + half * kHalfTile, h * kHeadDim + half * 64, qrow);
            }
            auto tile_rows = [&](int j, int pg) -> int {
                const int t0 = j * kKT + pg * 64;
                const int page = bt[(t0 < T ? t0 : 0) / p.page_size];  // a tile half past the end reads a valid page; it is masked
                return (page * p.n_kv + kvh) * p.page_size + (t0 < T ? t0 % p.page_size : 0);
            };
            auto load = [&](const CUtensorMap* map, uint32_t bar, uint32_t dst, int j) {
                mbar_expect_tx(bar, 2 * kHalfTile);
                for (int pg = 0; pg < 2; ++pg) {
                    const int row0 = tile_rows(j, pg);
                    for (int half = 0; half < 2; ++half)
                        tma_load_2d(map, bar, dst + half * kHalfTile + pg * (kHalfTile / 2), half * 64, row0);
                }
            };
            for (int j = 0; j < n_kt; ++j) {
                const int st = j & 1;
                const uint32_t ph = (uint32_t)(j >> 1) & 1u;
                mbar_wait(k_empty0 + 8 * st, ph ^ 1u);
                load(&map_k, k_full0 + 8 * st, base + kOffK + (uint32_t)st * 2 * kHalfTile, j);
                mbar_wait(v_empty0 + 8 * st, ph ^ 1u);
                load(&map_v, v_full0 + 8 * st, base + kOffV + (uint32_t)st * 2 * kHalfTile, j);
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {  // ===== MMA issuer
            constexpr uint32_t idesc_qk = umma_idesc(128, 128, 1, 0, 0);  // bf16, both K-major
            constexpr uint32_t idesc_pv = umma_idesc(128, 128, 1, 0, 1);  // bf16, P K-major, V MN-major
            auto issue_qk = [&](int j) {  // S[j & 1] = (q_hi + q_lo) K_j^T
                const int st = j & 1;
                mbar_wait(k_full0 + 8 * st, (uint32_t)(j >> 1) & 1u);
                fence_after();
                const uint32_t kb = base + kOffK + (uint32_t)st * 2 * kHalfTile;
                uint32_t acc = 0;
#pragma unroll
                for (int part = 0; part < 2; ++part) {
                    const uint32_t qb = base + (part == 0 ? kOffQhi : kOffQlo);
#pragma unroll
                    for (int half = 0; half < 2; ++half)
#pragma unroll
                        for (int k = 0; k < 4; ++k) {
                            umma_f16(tmem_base + st * 128, umma_desc_sw128(qb + half * kHalfTile + k * 32, 16, 1024),
                                     umma_desc_sw128(kb + half * kHalfTile + k * 32, 16, 1024), idesc_qk, acc);
                            acc = 1;
                        }
                }
                umma_commit(s_full0 + 8 * st);
                umma_commit(k_empty0 + 8 * st);
            };
            mbar_wait(q_full, 0);
            issue_qk(0);
            if (n_kt > 1) issue_qk(1);
            for (int j = 0; j < n_kt; ++j) {
                const int st = j & 1;
                mbar_wait(p_full, (uint32_t)j & 1u);                          // P_j written, S[j & 1] consumed
                mbar_wait(v_full0 + 8 * st, (uint32_t)(j >> 1) & 1u);
                mbar_wait(o_empty, ((uint32_t)j & 1u) ^ 1u);                  // O_tile of tile j-1 folded into registers
                fence_after();
                const uint32_t pb = base + kOffP, vb = base + kOffV + (uint32_t)st * 2 * kHalfTile;
#pragma unroll
                for (int kk = 0; kk < 8; ++kk)  // 16 keys per MMA
                    umma_f16(tmem_o, umma_desc_sw128(pb + (kk >> 2) * kHalfTile + (kk & 3) * 32, 16, 1024),
                             umma_desc_sw128(vb + kk * 2048, kHalfTile, 1024), idesc_pv, kk > 0 ? 1u : 0u);
                umma_commit(o_full);
                umma_commit(v_empty0 + 8 * st);
                if (j + 2 < n_kt) issue_qk(j + 2);                            // reuses S[j & 1]
            }
        }
    } else if (warp < 6) {  // ===== softmax warps: one query row per thread, never wait for the second MMA's result
        const int quad = warp & 3;
        const int r = quad * 32 + lane;
        const int qpos = base_len + q0 + r;
        const uint32_t lane_addr = (uint32_t)(quad * 32) << 16;
        const float scale_log2 = 0.08838834764831845f * 1.4426950408889634f;
        float m_run = -INFINITY, l_run = 0.f;
        for (int j = 0; j < n_kt; ++j) {
            const uint32_t tmem_s = tmem_base + (j & 1) * 128 + lane_addr;
            mbar_wait(s_full0 + 8 * (j & 1), (uint32_t)(j >> 1) & 1u);
            fence_after();
            const int key0 = j * kKT;
            const bool full = (key0 + kKT <= T) && (!p.causal || key0 + kKT - 1 <= base_len + q0);
            // the whole score row (128 values) is read from tensor memory ONCE and stays in registers for both passes
            uint32_t s[128];
#pragma unroll
            for (int c = 0; c < 4; ++c) tmem_ld32_nowait(tmem_s + c * 32, s + c * 32);
            tmem_ld_wait();
            float mx = -INFINITY;
            if (full) {
#pragma unroll
                for (int i = 0; i < 128; ++i) mx = fmaxf(mx, __uint_as_float(s[i]));
            } else {
#pragma unroll
                for (int i = 0; i < 128; ++i) {
                    const int key = key0 + i;
                    if (key < T && (!p.causal || key <= qpos)) mx = fmaxf(mx, __uint_as_float(s[i]));
                }
            }
            mx *= scale_log2;
            const float m_new = fmaxf(m_run, mx);
            const float m_use = (m_new == -INFINITY) ? 0.f : m_new;
            const float alpha = ex2_fast(m_run - m_use);
            alpha_s[(j % 3) * 128 + r] = alpha;
            float rs = 0.f;
            if (full) {
#pragma unroll
                for (int i = 0; i < 128; i += 2) {
                    const float p0 = ex2_fast(fmaf(__uint_as_float(s[i]), scale_log2, -m_use));
                    const float p1 = ex2_fast(fmaf(__uint_as_float(s[i + 1]), scale_log2, -m_use));
                    rs += p0 + p1;
                    s[i >> 1] = pack_bf16(p0, p1);
                }
            } else {
#pragma unroll
                for (int i = 0; i < 128; i += 2) {
                    const int key = key0 + i;
                    const bool ok0 = key < T && (!p.causal || key <= qpos);
                    const bool ok1 = key + 1 < T && (!p.causal || key + 1 <= qpos);
                    const float p0 = ok0 ? ex2_fast(fmaf(__uint_as_float(s[i]), scale_log2, -m_use)) : 0.f;
                    const float p1 = ok1 ? ex2_fast(fmaf(__uint_as_float(s[i + 1]), scale_log2, -m_use)) : 0.f;
                    rs += p0 + p1;
                    s[i >> 1] = pack_bf16(p0, p1);
                }
            }
            if (j > 0) mbar_wait(o_full, (uint32_t)(j - 1) & 1u);  // the second MMA of tile j-1 no longer reads P
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                uint8_t* prow = gbase + kOffP + (c >> 1) * kHalfTile + r * 128;
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const int chunk = (c & 1) * 4 + k;
                    *reinterpret_cast<uint4*>(prow + ((chunk ^ (r & 7)) << 4)) =
                        make_uint4(s[c * 16 + 4 * k], s[c * 16 + 4 * k + 1], s[c * 16 + 4 * k + 2], s[c * 16 + 4 * k + 3]);
                }
            }
            l_run = l_run * alpha + rs;
            m_run = m_new;
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // P must be visible to the tensor core's reads
            fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(p_full);
        }
        // 1/l goes through the alpha slot that the accumulating warps are guaranteed to be done with
        alpha_s[(n_kt % 3) * 128 + r] = l_run > 0.f ? 1.f / l_run : 0.f;
        asm volatile("bar.sync 1, 256;" ::: "memory");
    } else {  // ===== accumulating warps: fold O_tile = P_j V_j into registers, one tile behind the softmax
        const int quad = warp & 3;
        const int r = quad * 32 + lane;
        const uint32_t lane_addr = (uint32_t)(quad * 32) << 16;
        float o[128];
#pragma unroll
        for (int i = 0; i < 128; ++i) o[i] = 0.f;
        for (int j = 0; j < n_kt; ++j) {
            mbar_wait(o_full, (uint32_t)j & 1u);
            fence_after();
            const float alpha = alpha_s[(j % 3) * 128 + r];
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                uint32_t t[32];
                tmem_ld32(tmem_o + lane_addr + c * 32, t);
#pragma unroll
                for (int i = 0; i < 32; ++i) o[c * 32 + i] = fmaf(o[c * 32 + i], alpha, __uint_as_float(t[i]));
            }
            fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(o_empty);
        }
        asm volatile("bar.sync 1, 256;" ::: "memory");
        const float inv = alpha_s[(n_kt % 3) * 128 + r];
        if (q0 + r < p.S) {
            uint8_t* dst = reinterpret_cast<uint8_t*>(p.out) + (((size_t)b * p.S + q0 + r) * p.n_heads + h) * kHeadDim * 2;
#pragma unroll
            for (int v = 0; v < 16; ++v) {
                uint32_t w[4];
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    const float a = o[v * 8 + 2 * e] * inv, c2 = o[v * 8 + 2 * e + 1] * inv;
                    if (p.out_f16) {
                        __half2 t = __floats2half2_rn(a, c2);
                        w[e] = *reinterpret_cast<uint32_t*>(&t);
                    } else {
                        w[e] = pack_bf16(a, c2);
                    }
                }
                *reinterpret_cast<uint4*>(dst + v * 16) = make_uint4(w[0], w[1], w[2], w[3]);
            }
        }
    }
    fence_before();
    __syncthreads();
    if (warp == 1) {
        fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(512) : "memory");
    }
}

int launch_attn_prefill_tc(const AttnPrefillParams& p, int64_t pool_rows, cudaStream_t st) {
    CUtensorMap mqh, mql, mk, mv;
    const int q_rows = p.B * p.S, q_cols = p.n_heads * kHeadDim;
    if (tma_map_2d(&mqh, p.q, q_rows, q_cols, 128) != 0) return -1;
    if (tma_map_2d(&mql, p.q_lo, q_rows, q_cols, 128) != 0) return -1;
    if (tma_map_2d(&mk, p.k_pool, (int)pool_rows, kHeadDim, 64) != 0) return -1;
    if (tma_map_2d(&mv, p.v_pool, (int)pool_rows, kHeadDim, 64) != 0) return -1;
    cudaFuncSetAttribute(attn_prefill_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kTcSmem);
    dim3 grid((p.S + kQT - 1) / kQT, p.n_heads, p.B);
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = grid;
    cfg.blockDim = dim3(kTcThreads);
    cfg.dynamicSmemBytes = kTcSmem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    static int pdl = -1;
    if (pdl < 0) { const char* e = getenv("B200SHARD_PDL"); pdl = (e != nullptr && e[0] == '0') ? 0 : 1; }
    cfg.numAttrs = pdl;
    cudaLaunchKernelEx(&cfg, attn_prefill_tc_kernel, mqh, mql, mk, mv, p);
    return 0;
}

}  // namespace b200
