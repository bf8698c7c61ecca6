// Batched decode attention on the 5th-generation tensor cores with TMA-staged K/V pages:
// softmax_fp32(q k^T / sqrt(d)) v for ONE new token of each of many sequences (eager_attention_forward, transformers
// modeling_llama.py:199-221, reached from utils/shard_loader.py:66-74 with S = 1 and a batch of requests in flight).
//
// Why: the CUDA-core kernel (kernels_stream.cu, one half-warp per key) reads K/V through registers — one warp
// instruction per 8.9 bytes at 15 warps per SM, 62 % of HBM at 32 rows x 704 keys (profiles/results_r1.md).  Here no
// thread ever touches a K or V element: pages go HBM -> shared memory by TMA and are consumed by tcgen05.mma in place.
//
// One CTA = one (row, kv head): all n_heads / n_kv query heads that share the K/V (1 for Llama-2, 4 for Llama-3.x).
// The roles of the operands are swapped so that the 128-wide M side of the MMA is always full although there is a
// single query token:
//   S^T [128 keys x 16] = K_tile [128 keys x 128 d]  (A, K-major, straight from the page layout)
//                         x Q^T  [128 d x 16]         (B, K-major)
//   O^T [128 d x 16]   += V_tile^T [128 d x 128 keys] (A, MN-major: the natural [token][dim] page layout, no transpose)
//                         x P^T  [128 keys x 16]      (B, K-major)
// The 16 columns of the narrow operand carry PRECISION, not padding: fp32 q is split into bf16 parts (hi, mid, lo) and
// the parts of head h sit in columns h, g + h, 2g + h of the same operand, so ONE pass over the K tile yields the partial
// scores of every part and the softmax thread adds its columns; the probabilities travel the same way as hi and lo
// columns and the two output columns are added at the end.  (Measured before this layout, with one MMA pass per part: 40
// MMAs per 128 keys made the kernel MMA-issue bound, 64-81 % of HBM; 16 MMAs reach 92-94 %, profiles/attn_decode_tc_r2.md.)
// Two passes over the CTA's keys, so no accumulator is ever rescaled: pass 1 streams the K pages and parks the
// scaled scores in shared memory (4 bytes per key and head) while a running maximum is kept; pass 2 streams the V pages,
// turns scores into probabilities against the final maximum and accumulates O^T in tensor memory.  K and V tiles share
// one TMA ring (K0 .. Kn-1, V0 .. Vn-1): the V pages are already in flight while the maximum is reduced.
//   warp 0     TMA producer (tiles of 128 keys = two 64-token pages looked up in the block table; every tile that lies
//              below the length cached BEFORE this call is requested ahead of griddepcontrol.wait; every tensor map
//              covers this GPU's K/V pool only)
//   warp 1     single-thread tcgen05.mma issuer
//   warps 2-5  softmax (TMEM lane = key in pass 1, = output dim at the end), P written as the B operand
// Keys past a row's length inside its last tile get probability exactly 0; what their K/V rows hold must be finite
// (the pool is zero-initialised and only ever holds values the model wrote).
#include <stdio.h>
#include <stdlib.h>

#include "common.cuh"
#include "internal.h"
#include "tc_common.cuh"

namespace b200 {
using namespace tc;

constexpr int kDtThreads = 192;
constexpr uint32_t kDtTile = 128 * 128 * 2;       // one 128-key K or V tile: 32 KB (two 128B-swizzled halves of 64 dims)
constexpr uint32_t kDtHalf = kDtTile / 2;
constexpr uint32_t kDtQBytes = 2 * 2 * 1024;      // q operand: 2 dim-halves x two 8-row swizzle atoms (pass 1)
constexpr uint32_t kDtPBytes = 2 * 2 * 2 * 1024;  // P operand: 2 buffers x 2 key-halves x two atoms (pass 2, same space)
static_assert(kDtQBytes <= kDtPBytes, "the q operand lives in the P buffers' space");
constexpr int kDtMaxGroup = 8;                    // query heads per kv head (2 probability columns each)

__device__ __forceinline__ void dt_tmem_ld16(uint32_t taddr, uint32_t* r) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void dt_tmem_ld8(uint32_t taddr, uint32_t* r) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                 : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void dt_sync() { asm volatile("bar.sync 1, 128;" ::: "memory"); }   // the four softmax warps
// byte offset of element (operand row r < 16, column c < 64) inside a 16-row K-major SWIZZLE_128B block: two 8-row atoms
__device__ __forceinline__ uint32_t dt_op_off(int r, int c) {
    return (uint32_t)((r >> 3) * 1024 + (r & 7) * 128 + ((((c >> 3) ^ r) & 7) << 4) + (c & 7) * 2);
}
__device__ __forceinline__ float dt_ex2(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

struct DtParams {
    AttnDecodeParams a;
    int sc_stride;   // keys the score buffer holds per head (a multiple of 128)
    int q_parts;     // bf16 parts of q: 3 while 3 * group <= 16, else 2
};

// kDtN: columns of the narrow operand — 8 while the parts of the group fit (3 * g <= 8, e.g. MHA), else 16
template <int NS, int kDtN>
__global__ void __launch_bounds__(kDtThreads, NS <= 3 ? 2 : 1)
attn_decode_tc_kernel(const __grid_constant__ CUtensorMap map_k, const __grid_constant__ CUtensorMap map_v, const DtParams dp) {
    const AttnDecodeParams& p = dp.a;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t raw = s_u32(smem_raw);
    const uint32_t base = (raw + 1023u) & ~1023u;
    uint8_t* gbase = smem_raw + (base - raw);
    constexpr uint32_t kOffQP = NS * kDtTile;                 // q parts in pass 1, P buffers in pass 2 (pass 2 starts after the last q MMA)
    constexpr uint32_t kOffSc = kOffQP + kDtPBytes;
    float* sc = reinterpret_cast<float*>(gbase + kOffSc);
    const uint32_t misc = base + kOffSc + (uint32_t)dp.sc_stride * 4u * (uint32_t)(p.n_heads / p.n_kv);
    uint8_t* gmisc = gbase + (misc - base);
    const uint32_t full0 = misc, empty0 = misc + 48, s_full0 = misc + 96, s_empty0 = misc + 112, p_full0 = misc + 128,
                   p_empty0 = misc + 144, o_full = misc + 160, q_full = misc + 168;
    volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(gmisc + 176);
    float* red_max = reinterpret_cast<float*>(gmisc + 192);   // [4 warps][8 heads]
    float* red_sum = reinterpret_cast<float*>(gmisc + 320);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int kvh = blockIdx.x, m = blockIdx.y;
    const int g = p.n_heads / p.n_kv;
    const int T = p.t_end[m];
    const int n_t = (T + 127) >> 7;
    const int b = m / p.S;
    const int slot = p.slots.v[b];
    const int32_t* bt = p.block_table + (size_t)slot * p.max_pages_per_slot;
    if (T > dp.sc_stride) __trap();   // the host sizes the score buffer from the cached length; never reached unless that logic breaks

    if (warp == 0 && lane == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_k) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_v) : "memory");
        for (int i = 0; i < NS; ++i) { mbar_init(full0 + 8 * i, 1); mbar_init(empty0 + 8 * i, 1); }
        for (int i = 0; i < 2; ++i) {
            mbar_init(s_full0 + 8 * i, 1); mbar_init(s_empty0 + 8 * i, 4);
            mbar_init(p_full0 + 8 * i, 4); mbar_init(p_empty0 + 8 * i, 1);
        }
        mbar_init(o_full, 1);
        mbar_init(q_full, 4);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {  // S^T buffers at columns 0 and 16, O^T at column 32
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(s_u32((const void*)tmem_slot)), "n"(64) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    fence_before();
    __syncthreads();
    fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const uint32_t tmem_o = tmem_base + 32;

    if (warp == 0) {
        if (lane == 0) {  // ===== TMA producer: K tiles of pass 1, then V tiles of pass 2, one ring
            const int t_old = (p.call_base != nullptr) ? min(T, p.call_base[b]) : 0;
            auto issue = [&](int i) {
                const int st = i % NS, tile = i < n_t ? i : i - n_t;
                if (i >= NS) mbar_wait(empty0 + 8 * st, ((uint32_t)(i / NS) & 1u) ^ 1u);
                mbar_expect_tx(full0 + 8 * st, kDtTile);
                const CUtensorMap* map = i < n_t ? &map_k : &map_v;
                const uint32_t dst = base + (uint32_t)st * kDtTile;
                for (int pg = 0; pg < 2; ++pg) {
                    const int t0 = tile * 128 + pg * 64;
                    const int page = bt[(t0 < T ? t0 : 0) / p.page_size];   // a half past the end reads a valid page: its keys are masked
                    const int row0 = (page * p.n_kv + kvh) * p.page_size + (t0 < T ? t0 % p.page_size : 0);
                    for (int half = 0; half < 2; ++half)
                        tma_load_2d(map, full0 + 8 * st, dst + half * kDtHalf + pg * (kDtHalf / 2), half * 64, row0);
                }
            };
            int i = 0;
            // tiles that lie entirely below the length cached before this call do not depend on the previous kernel
            for (; i < 2 * n_t && i < NS; ++i) {
                const int tile = i < n_t ? i : i - n_t;
                if ((tile + 1) * 128 > t_old) break;
                issue(i);
            }
            asm volatile("griddepcontrol.wait;" ::: "memory");
            asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
            for (; i < 2 * n_t; ++i) issue(i);
        }
    } else if (warp == 1) {
        if (lane == 0) {  // ===== MMA issuer
            constexpr uint32_t idesc_qk = umma_idesc(128, kDtN, 1, 0, 0);   // bf16; K tile K-major, q K-major
            constexpr uint32_t idesc_pv = umma_idesc(128, kDtN, 1, 1, 0);   // bf16; V tile MN-major (A), P K-major
            const uint32_t qp = base + kOffQP;
            mbar_wait(q_full, 0);
            fence_after();
            for (int j = 0; j < n_t; ++j) {
                const int st = j % NS;
                mbar_wait(full0 + 8 * st, (uint32_t)(j / NS) & 1u);
                mbar_wait(s_empty0 + 8 * (j & 1), (((uint32_t)j >> 1) & 1u) ^ 1u);
                fence_after();
                const uint32_t kt = base + (uint32_t)st * kDtTile, acc = tmem_base + (uint32_t)(j & 1) * kDtN;
#pragma unroll
                for (int half = 0; half < 2; ++half)
#pragma unroll
                    for (int k = 0; k < 4; ++k)   // every bf16 part of every head of the group rides in the 16 columns: one pass
                        umma_f16(acc, umma_desc_sw128(kt + half * kDtHalf + k * 32, 16, 1024),
                                 umma_desc_sw128(qp + half * 2048 + k * 32, 16, 1024), idesc_qk, (half > 0 || k > 0) ? 1u : 0u);
                umma_commit(s_full0 + 8 * (j & 1));
                umma_commit(empty0 + 8 * st);
            }
            for (int j = 0; j < n_t; ++j) {
                const int i = n_t + j, st = i % NS;
                mbar_wait(full0 + 8 * st, (uint32_t)(i / NS) & 1u);
                mbar_wait(p_full0 + 8 * (j & 1), ((uint32_t)j >> 1) & 1u);
                fence_after();
                const uint32_t vt = base + (uint32_t)st * kDtTile;
                const uint32_t pb = qp + (uint32_t)(j & 1) * 4096;
#pragma unroll
                for (int kk = 0; kk < 8; ++kk)   // 16 keys per MMA; hi and lo probabilities are different columns
                    umma_f16(tmem_o, umma_desc_sw128(vt + kk * 2048, kDtHalf, 1024),
                             umma_desc_sw128(pb + (kk >> 2) * 2048 + (kk & 3) * 32, 16, 1024), idesc_pv, (j > 0 || kk > 0) ? 1u : 0u);
                umma_commit(p_empty0 + 8 * (j & 1));
                umma_commit(empty0 + 8 * st);
            }
            umma_commit(o_full);
        }
    } else {  // ===== softmax warps
        asm volatile("griddepcontrol.wait;" ::: "memory");   // q (and this call's own K/V rows) come from the previous kernel
        const int quad = warp & 3;
        const int t = quad * 32 + lane;
        const uint32_t lane_addr = (uint32_t)(quad * 32) << 16;
        uint8_t* qp = gbase + kOffQP;
        // ---- q: fp32 -> bf16 parts; part i of head h is operand row i * g + h, the other rows are zero.  Thread t owns dim t.
        {
            uint8_t* qcol = qp + (uint32_t)(t >> 6) * 2048;
            const int c = t & 63;
            for (int r = 0; r < kDtN; ++r) *reinterpret_cast<__nv_bfloat16*>(qcol + dt_op_off(r, c)) = __float2bfloat16_rn(0.f);
            float qv[kDtMaxGroup];
#pragma unroll
            for (int h = 0; h < kDtMaxGroup; ++h)   // independent loads
                qv[h] = h < g ? p.q[((size_t)m * p.n_heads + kvh * g + h) * kHeadDim + t] : 0.f;
#pragma unroll
            for (int h = 0; h < kDtMaxGroup; ++h)
                if (h < g) {
                    float v = qv[h];
                    for (int part = 0; part < dp.q_parts; ++part) {
                        const __nv_bfloat16 o = __float2bfloat16_rn(v);
                        *reinterpret_cast<__nv_bfloat16*>(qcol + dt_op_off(part * g + h, c)) = o;
                        v -= __bfloat162float(o);
                    }
                }
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncwarp();
        if (lane == 0) mbar_arrive(q_full);
        const float scale_log2 = 0.08838834764831845f * 1.4426950408889634f;   // 128^-0.5 * log2(e)
        float m_run[kDtMaxGroup], l_run[kDtMaxGroup];
#pragma unroll
        for (int h = 0; h < kDtMaxGroup; ++h) { m_run[h] = -INFINITY; l_run[h] = 0.f; }
        // ---- pass 1: scores of every key of this row, scaled, parked in shared memory
        for (int j = 0; j < n_t; ++j) {
            mbar_wait(s_full0 + 8 * (j & 1), ((uint32_t)j >> 1) & 1u);
            fence_after();
            uint32_t r[16];
            if constexpr (kDtN == 16) dt_tmem_ld16(tmem_base + (uint32_t)(j & 1) * kDtN + lane_addr, r);
            else dt_tmem_ld8(tmem_base + (uint32_t)(j & 1) * kDtN + lane_addr, r);
            fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(s_empty0 + 8 * (j & 1));
            const int key = j * 128 + t;
#pragma unroll
            for (int h = 0; h < kDtMaxGroup; ++h)
                if (h < g) {
                    float dot = __uint_as_float(r[h]);   // columns h, g + h, 2g + h: the partial scores of the bf16 parts of q
                    if (dp.q_parts > 1) dot += __uint_as_float(r[g + h]);
                    if (dp.q_parts > 2) dot += __uint_as_float(r[2 * g + h]);
                    const float s = key < T ? dot * scale_log2 : -INFINITY;
                    sc[h * dp.sc_stride + key] = s;
                    m_run[h] = fmaxf(m_run[h], s);
                }
        }
#pragma unroll
        for (int h = 0; h < kDtMaxGroup; ++h)
            if (h < g) {
                const float v = warp_max(m_run[h]);
                if (lane == 0) red_max[quad * 8 + h] = v;
            }
        // the q parts are no longer read (every score MMA has completed): their space becomes the P buffers, rows >= g zero
        for (int idx = t; idx < (int)(kDtPBytes / 16); idx += 128) reinterpret_cast<uint4*>(qp)[idx] = make_uint4(0, 0, 0, 0);
        dt_sync();
#pragma unroll
        for (int h = 0; h < kDtMaxGroup; ++h)
            if (h < g) m_run[h] = fmaxf(fmaxf(red_max[h], red_max[8 + h]), fmaxf(red_max[16 + h], red_max[24 + h]));
        // ---- pass 2: probabilities against the final maximum, as bf16 hi + lo B operands
        for (int j = 0; j < n_t; ++j) {
            mbar_wait(p_empty0 + 8 * (j & 1), (((uint32_t)j >> 1) & 1u) ^ 1u);
            const int key = j * 128 + t;
            uint8_t* pb = qp + (uint32_t)(j & 1) * 4096 + (uint32_t)(t >> 6) * 2048;
#pragma unroll
            for (int h = 0; h < kDtMaxGroup; ++h)
                if (h < g) {
                    const float s = sc[h * dp.sc_stride + key];
                    const float pr = s == -INFINITY ? 0.f : dt_ex2(s - m_run[h]);
                    l_run[h] += pr;
                    const __nv_bfloat16 hi = __float2bfloat16_rn(pr);
                    const __nv_bfloat16 lo = __float2bfloat16_rn(pr - __bfloat162float(hi));
                    *reinterpret_cast<__nv_bfloat16*>(pb + dt_op_off(h, t & 63)) = hi;        // column h
                    *reinterpret_cast<__nv_bfloat16*>(pb + dt_op_off(g + h, t & 63)) = lo;    // column g + h
                }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncwarp();
            if (lane == 0) mbar_arrive(p_full0 + 8 * (j & 1));
        }
#pragma unroll
        for (int h = 0; h < kDtMaxGroup; ++h)
            if (h < g) {
                const float v = warp_sum(l_run[h]);
                if (lane == 0) red_sum[quad * 8 + h] = v;
            }
        dt_sync();
        // ---- output: TMEM lane = dim t of every head of the group
        mbar_wait(o_full, 0);
        fence_after();
        uint32_t o[16];
        if constexpr (kDtN == 16) dt_tmem_ld16(tmem_o + lane_addr, o);
        else dt_tmem_ld8(tmem_o + lane_addr, o);
#pragma unroll
        for (int h = 0; h < kDtMaxGroup; ++h)
            if (h < g) {
                const float l = (red_sum[h] + red_sum[8 + h]) + (red_sum[16 + h] + red_sum[24 + h]);
                const float v = (__uint_as_float(o[h]) + __uint_as_float(o[g + h])) / l;
                const size_t idx = ((size_t)m * p.n_heads + kvh * g + h) * kHeadDim + t;
                p.out[idx] = v;
                if (p.out_h != nullptr) p.out_h[idx] = __float2half_rn(v);
            }
        fence_before();
    }
    __syncthreads();
    if (warp == 1) {
        fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(64) : "memory");
    }
}

// returns 0 = launched, 1 = this shape is not served here (the caller keeps the CUDA-core kernel), -1 = error (gemm_error())
int launch_attn_decode_tc(const AttnDecodeParams& p, int M, int max_keys, int64_t pool_rows, cudaStream_t st) {
    const int g = p.n_heads / p.n_kv;
    if (g < 1 || g > kDtMaxGroup || p.n_heads % p.n_kv != 0 || p.page_size % 64 != 0 || max_keys < 1) return 1;
    DtParams dp;
    dp.a = p;
    const int n_cols = 3 * g <= 8 ? 8 : 16;
    dp.q_parts = 3 * g <= n_cols ? 3 : 2;   // 24 or 16 mantissa bits of q; the probabilities always travel as hi + lo (2 * g <= n_cols)
    dp.sc_stride = (max_keys + 1023) / 1024 * 1024;   // whole kilo-keys: a captured graph stays valid while the context grows inside its bucket
    const size_t fixed = 1024 + kDtPBytes + (size_t)g * dp.sc_stride * 4 + 512;
    int ns = 0;
    if (fixed + 3 * (size_t)kDtTile <= 113 * 1024) ns = 3;          // two CTAs per SM
    else if (fixed + 6 * (size_t)kDtTile <= 227 * 1024) ns = 6;
    else if (fixed + 4 * (size_t)kDtTile <= 227 * 1024) ns = 4;
    else return 1;
    CUtensorMap mk, mv;
    if (tma_map_2d(&mk, p.k_pool, (int)pool_rows, kHeadDim, 64) != 0) return -1;
    if (tma_map_2d(&mv, p.v_pool, (int)pool_rows, kHeadDim, 64) != 0) return -1;
    const size_t smem = fixed + (size_t)ns * kDtTile;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(p.n_kv, M);
    cfg.blockDim = dim3(kDtThreads);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    static int pdl = -1;
    if (pdl < 0) { const char* e = getenv("B200SHARD_PDL"); pdl = (e != nullptr && e[0] == '0') ? 0 : 1; }
    cfg.numAttrs = pdl;
    cudaError_t e;
#define B200_DT_LAUNCH(KFN)                                                                        \
    {                                                                                              \
        auto kfn_ = KFN;                                                                           \
        cudaFuncSetAttribute(kfn_, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);        \
        e = cudaLaunchKernelEx(&cfg, kfn_, mk, mv, dp);                                            \
    }
    if (n_cols == 8) {
        if (ns == 3) B200_DT_LAUNCH((attn_decode_tc_kernel<3, 8>))
        else if (ns == 6) B200_DT_LAUNCH((attn_decode_tc_kernel<6, 8>))
        else B200_DT_LAUNCH((attn_decode_tc_kernel<4, 8>))
    } else {
        if (ns == 3) B200_DT_LAUNCH((attn_decode_tc_kernel<3, 16>))
        else if (ns == 6) B200_DT_LAUNCH((attn_decode_tc_kernel<6, 16>))
        else B200_DT_LAUNCH((attn_decode_tc_kernel<4, 16>))
    }
#undef B200_DT_LAUNCH
    return e == cudaSuccess ? 0 : -1;
}

}  // namespace b200
