// Prefill attention over the paged KV cache: softmax_fp32(q k^T / sqrt(d) [+ causal mask]) v
// (eager_attention_forward, transformers modeling_llama.py:199-221, reached from
// utils/shard_loader.py:66-74).  The reference passes NO mask (utils/node_worker.py:253-258), so
// `causal == 0` is the parity mode; `causal == 1` is real-Llama masking.
//
// Flash-style: one CTA = 64 query rows of one (batch row, head); K/V tiles of 64 keys are
// staged into swizzled shared memory with 16-byte cp.async (double buffered), scores and P.V
// run on the tensor cores (bf16 mma, fp32 accumulate), the online softmax is fp32 with
// warp-shuffle row reductions.  The [S,T] score matrix the reference materialises never exists.
#include "common.cuh"
#include "internal.h"

namespace b200 {

constexpr int kBM = 64, kBN = 64;
constexpr int kPrefillThreads = 128;
constexpr int kTileBytes = 64 * 256;  // 64 rows x 128 bf16

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
// byte offset of 16-byte chunk `chunk` (0..15) of row `row` in a [64][128] bf16 tile, XOR-swizzled
__device__ __forceinline__ uint32_t swz(int row, int chunk) { return (uint32_t)(row * 256 + ((chunk ^ (row & 7)) << 4)); }

__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, bool valid) {
    const int sz = valid ? 16 : 0;  // src-size 0 => the 16 bytes are zero-filled
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(sz));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

__device__ __forceinline__ void ldsm_x4(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void mma_bf16(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

__global__ void __launch_bounds__(kPrefillThreads) attn_prefill_kernel(const AttnPrefillParams p) {
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t* sQ = smem;                      // 16 KB (reused to stage the output)
    uint8_t* sK = smem + kTileBytes;         // 2 x 16 KB
    uint8_t* sV = smem + 3 * kTileBytes;     // 2 x 16 KB
    uint8_t* sQlo = smem + 5 * kTileBytes;   // 16 KB, only when q arrives as a (hi, lo) bf16 pair
    const bool split_q = p.q_lo != nullptr;
    const int qt = blockIdx.x, h = blockIdx.y, b = blockIdx.z;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int q0 = qt * kBM;
    const int base = p.call_base[b];
    const int T = base + p.S;
    const int slot = p.slots.v[b];
    const int kvh = h / (p.n_heads / p.n_kv);
    const int32_t* bt = p.block_table + (size_t)slot * p.max_pages_per_slot;
    const int64_t head_stride = (int64_t)p.page_size * kHeadDim;
    const int64_t page_stride = (int64_t)p.n_kv * head_stride;
    int n_kt = (T + kBN - 1) / kBN;
    if (p.causal) {
        const int lim = min(T, base + q0 + kBM);
        n_kt = (lim + kBN - 1) / kBN;
    }

    // ---- Q tile
    for (int i = tid; i < 64 * 16; i += kPrefillThreads) {
        const int r = i >> 4, c = i & 15;
        const bool ok = (q0 + r) < p.S;
        const __nv_bfloat16* src = p.q + (((size_t)b * p.S + (ok ? q0 + r : 0)) * p.n_heads + h) * kHeadDim + c * 8;
        cp_async16(smem_u32(sQ) + swz(r, c), src, ok);
        if (split_q) cp_async16(smem_u32(sQlo) + swz(r, c), p.q_lo + (src - p.q), ok);
    }
    auto load_kv = [&](int kt, int stage) {
        const int t0 = kt * kBN;
        const int page = bt[t0 / p.page_size];
        const int64_t off = page * page_stride + kvh * head_stride + (int64_t)(t0 % p.page_size) * kHeadDim;
        const __nv_bfloat16* ksrc = p.k_pool + off;
        const __nv_bfloat16* vsrc = p.v_pool + off;
        const uint32_t kd = smem_u32(sK) + stage * kTileBytes, vd = smem_u32(sV) + stage * kTileBytes;
        for (int i = tid; i < 64 * 16; i += kPrefillThreads) {
            const int r = i >> 4, c = i & 15;
            const bool ok = (t0 + r) < T;
            cp_async16(kd + swz(r, c), ksrc + (ok ? r * kHeadDim + c * 8 : 0), ok);
            cp_async16(vd + swz(r, c), vsrc + (ok ? r * kHeadDim + c * 8 : 0), ok);
        }
    };
    load_kv(0, 0);
    cp_async_commit();

    float o[16][4];
#pragma unroll
    for (int i = 0; i < 16; ++i) { o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f; }
    float m_run[2] = {-INFINITY, -INFINITY}, l_run[2] = {0.f, 0.f};
    const float scale_log2 = 0.08838834764831845f * 1.4426950408889634f;
    const int row0 = q0 + warp * 16 + (lane >> 2);  // sequence index of this thread's first row (second: +8)

    for (int kt = 0; kt < n_kt; ++kt) {
        const int stage = kt & 1;
        if (kt + 1 < n_kt) load_kv(kt + 1, stage ^ 1);
        cp_async_commit();
        cp_async_wait<1>();
        __syncthreads();
        const uint32_t kb = smem_u32(sK) + stage * kTileBytes, vb = smem_u32(sV) + stage * kTileBytes;
        // ---- S = Q K^T  (Q fragments are re-read from shared memory per k-step pair: with the hi/lo split they
        //      would not fit in registers next to the 64 output accumulators)
        float s[8][4];
#pragma unroll
        for (int nt = 0; nt < 8; ++nt) s[nt][0] = s[nt][1] = s[nt][2] = s[nt][3] = 0.f;
#pragma unroll
        for (int ks = 0; ks < 8; ks += 2) {
            const int mi = lane >> 3, r = lane & 7;
            uint32_t qh0[4], qh1[4], ql0[4], ql1[4];
            const int qrow = warp * 16 + (mi & 1) * 8 + r;
            ldsm_x4(smem_u32(sQ) + swz(qrow, ks * 2 + (mi >> 1)), qh0[0], qh0[1], qh0[2], qh0[3]);
            ldsm_x4(smem_u32(sQ) + swz(qrow, ks * 2 + 2 + (mi >> 1)), qh1[0], qh1[1], qh1[2], qh1[3]);
            if (split_q) {
                ldsm_x4(smem_u32(sQlo) + swz(qrow, ks * 2 + (mi >> 1)), ql0[0], ql0[1], ql0[2], ql0[3]);
                ldsm_x4(smem_u32(sQlo) + swz(qrow, ks * 2 + 2 + (mi >> 1)), ql1[0], ql1[1], ql1[2], ql1[3]);
            }
#pragma unroll
            for (int nt = 0; nt < 8; ++nt) {
                uint32_t b0, b1, b2, b3;
                ldsm_x4(kb + swz(nt * 8 + r, ks * 2 + mi), b0, b1, b2, b3);
                mma_bf16(s[nt], qh0, b0, b1);
                mma_bf16(s[nt], qh1, b2, b3);
                if (split_q) {
                    mma_bf16(s[nt], ql0, b0, b1);
                    mma_bf16(s[nt], ql1, b2, b3);
                }
            }
        }
        // ---- mask + online softmax (fp32)
        float mx[2] = {-INFINITY, -INFINITY};
#pragma unroll
        for (int nt = 0; nt < 8; ++nt) {
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const int key = kt * kBN + nt * 8 + (lane & 3) * 2 + (e & 1);
                const int row = row0 + (e >> 1) * 8;
                const bool ok = key < T && (!p.causal || key <= base + row);
                const float v = ok ? s[nt][e] * scale_log2 : -INFINITY;
                s[nt][e] = v;
                mx[e >> 1] = fmaxf(mx[e >> 1], v);
            }
        }
        float alpha[2], m_use[2];
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            mx[i] = fmaxf(mx[i], __shfl_xor_sync(0xffffffffu, mx[i], 1));
            mx[i] = fmaxf(mx[i], __shfl_xor_sync(0xffffffffu, mx[i], 2));
            const float m_new = fmaxf(m_run[i], mx[i]);
            m_use[i] = (m_new == -INFINITY) ? 0.f : m_new;
            alpha[i] = exp2f(m_run[i] - m_use[i]);
            m_run[i] = m_new;
        }
        float rs[2] = {0.f, 0.f};
        uint32_t pf[4][4];  // P as A fragments, 4 k-steps of 16 keys
#pragma unroll
        for (int nt = 0; nt < 8; ++nt) {
            const float p0 = exp2f(s[nt][0] - m_use[0]), p1 = exp2f(s[nt][1] - m_use[0]);
            const float p2 = exp2f(s[nt][2] - m_use[1]), p3 = exp2f(s[nt][3] - m_use[1]);
            rs[0] += p0 + p1;
            rs[1] += p2 + p3;
            pf[nt >> 1][(nt & 1) * 2 + 0] = pack_bf16(p0, p1);
            pf[nt >> 1][(nt & 1) * 2 + 1] = pack_bf16(p2, p3);
        }
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            rs[i] += __shfl_xor_sync(0xffffffffu, rs[i], 1);
            rs[i] += __shfl_xor_sync(0xffffffffu, rs[i], 2);
            l_run[i] = l_run[i] * alpha[i] + rs[i];
        }
#pragma unroll
        for (int dt = 0; dt < 16; ++dt) {
            o[dt][0] *= alpha[0]; o[dt][1] *= alpha[0];
            o[dt][2] *= alpha[1]; o[dt][3] *= alpha[1];
        }
        // ---- O += P V
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) {
#pragma unroll
            for (int dc = 0; dc < 16; dc += 2) {
                uint32_t b0, b1, b2, b3;
                const int mi = lane >> 3, r = lane & 7;
                ldsm_x4_t(vb + swz(kk * 16 + (mi & 1) * 8 + r, dc + (mi >> 1)), b0, b1, b2, b3);
                mma_bf16(o[dc], pf[kk], b0, b1);
                mma_bf16(o[dc + 1], pf[kk], b2, b3);
            }
        }
        __syncthreads();  // everyone done with this stage before it is refilled
    }
    cp_async_wait<0>();

    // ---- normalise, stage through smem (Q tile is dead), coalesced 16-byte stores
    const float inv0 = l_run[0] > 0.f ? 1.f / l_run[0] : 0.f, inv1 = l_run[1] > 0.f ? 1.f / l_run[1] : 0.f;
    const int lr = warp * 16 + (lane >> 2);
#pragma unroll
    for (int dt = 0; dt < 16; ++dt) {
        const int col = dt * 8 + (lane & 3) * 2;  // element column
        uint32_t w0, w1;
        if (p.out_f16) {
            __half2 t0 = __floats2half2_rn(o[dt][0] * inv0, o[dt][1] * inv0), t1 = __floats2half2_rn(o[dt][2] * inv1, o[dt][3] * inv1);
            w0 = *reinterpret_cast<uint32_t*>(&t0);
            w1 = *reinterpret_cast<uint32_t*>(&t1);
        } else {
            w0 = pack_bf16(o[dt][0] * inv0, o[dt][1] * inv0);
            w1 = pack_bf16(o[dt][2] * inv1, o[dt][3] * inv1);
        }
        *reinterpret_cast<uint32_t*>(sQ + swz(lr, col >> 3) + (col & 7) * 2) = w0;
        *reinterpret_cast<uint32_t*>(sQ + swz(lr + 8, col >> 3) + (col & 7) * 2) = w1;
    }
    __syncthreads();
    for (int i = tid; i < 64 * 16; i += kPrefillThreads) {
        const int r = i >> 4, c = i & 15;
        if (q0 + r < p.S) {
            const uint4 v = *reinterpret_cast<const uint4*>(sQ + swz(r, c));
            *reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(p.out) + (((size_t)b * p.S + q0 + r) * p.n_heads + h) * kHeadDim + c * 8) = v;
        }
    }
}

void launch_attn_prefill(const AttnPrefillParams& p, cudaStream_t st) {
    const size_t smem = 6 * kTileBytes;
    cudaFuncSetAttribute(attn_prefill_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    dim3 grid((p.S + kBM - 1) / kBM, p.n_heads, p.B);
    attn_prefill_kernel<<<grid, kPrefillThreads, smem, st>>>(p);
}

}  // namespace b200
