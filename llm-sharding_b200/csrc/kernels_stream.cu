// HBM-bound kernels of the shard path: the weight-streaming skinny GEMM used for decode
// (fused RMSNorm prologue; RoPE + paged-KV append / residual / SwiGLU epilogues), RMSNorm,
// embedding gather, arg-max, KV page bookkeeping.  Replaces the per-op ATen launches under
// HF LlamaDecoderLayer (transformers modeling_llama.py:62-67,138-168,182-184,303-332) that the
// reference calls from utils/shard_loader.py:66-74.
#include <stdio.h>

#include <mutex>

#include "common.cuh"
#include "internal.h"

namespace b200 {

// =========================================================================== GEMV
// out[m][n] = sum_k W[n][k] * xin[m][k],  M <= 4 token rows, weights bf16 streamed once.
// One warp owns R weight rows at a time ("item"); lanes stride over K in 16-byte chunks.
// xin lives in shared memory as fp32 (optionally RMS-normalised on the way in), laid out
// so that each lane's two float4 reads per step are bank-conflict free:
//   element k -> chunk c=k/8, step=c/32, lane=c%32, half=(k%8)/4, e=k%4
//   smem float index = ((step*2+half)*32 + lane)*4 + e
constexpr int kGemvThreads = 512;

// two packed 16-bit weights -> floats (the shard path holds weights as fp16; the bf16 form serves the unit entry point)
template <bool WBF16>
__device__ __forceinline__ float2 unpack2(uint32_t v) {
    if (WBF16) return make_float2(bf_lo(v), bf_hi(v));
    return __half22float2(*reinterpret_cast<const __half2*>(&v));
}

__device__ __forceinline__ int xs_index(int k) {
    const int c = k >> 3, step = c >> 5, lane = c & 31, half = (k >> 2) & 1, e = k & 3;
    return (((step * 2 + half) * 32 + lane) << 2) + e;
}

template <int M, int R, int EPI, bool NORM, bool WBF16>
__global__ void __launch_bounds__(kGemvThreads, (M == 1 ? 2 : 1)) gemv_kernel(const GemvParams p) {
    extern __shared__ __align__(16) float xs[];  // [M][Kpad]
    __shared__ float red[33];
    const int K = p.K;
    const int nsteps = (K / 8 + 31) / 32;
    const int Kpad = nsteps * 256;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    // ---- prologue: stage (normalised) activations in shared memory
#pragma unroll
    for (int m = 0; m < M; ++m) {
        const float* xr = p.x + (size_t)m * K;
        float rstd = 1.f;
        if (NORM) {
            float ss = 0.f;
            for (int k = tid; k < K; k += kGemvThreads) { const float v = xr[k]; ss += v * v; }
            ss = block_sum(ss, red);
            rstd = rsqrtf(ss / (float)K + p.eps);
        }
        float* xm = xs + (size_t)m * Kpad;
        for (int k = tid; k < Kpad; k += kGemvThreads) {
            float v = 0.f;
            if (k < K) {
                v = xr[k];
                if (NORM) v = v * rstd * __bfloat162float(p.norm_w[k]);
            }
            xm[xs_index(k)] = v;
        }
    }
    __syncthreads();

    const int warps_per_cta = kGemvThreads / 32;
    const int gw = blockIdx.x * warps_per_cta + warp;
    const int total_warps = gridDim.x * warps_per_cta;
    constexpr bool PAIRED = (EPI == EPI_QKV || EPI == EPI_SWIGLU);
    const int n_items = PAIRED ? (p.N / 128) * (128 / R) : (p.N + R - 1) / R;
    const int kchunks = K / 8;
    const float4* xs4 = reinterpret_cast<const float4*>(xs);
    const int m_stride4 = Kpad / 4;

    for (int item = gw; item < n_items; item += total_warps) {
        int rows[R];
        if (PAIRED) {
            const int per_tile = 128 / R;
            const int base = (item / per_tile) * 128 + (item % per_tile) * (R / 2);
#pragma unroll
            for (int r = 0; r < R; ++r) rows[r] = base + (r >> 1) + 64 * (r & 1);
        } else {
#pragma unroll
            for (int r = 0; r < R; ++r) rows[r] = min(item * R + r, p.N - 1);
        }
        const uint4* wrow[R];
#pragma unroll
        for (int r = 0; r < R; ++r) wrow[r] = reinterpret_cast<const uint4*>(reinterpret_cast<const uint16_t*>(p.W) + (size_t)rows[r] * K);

        float acc[R][M];
#pragma unroll
        for (int r = 0; r < R; ++r)
#pragma unroll
            for (int m = 0; m < M; ++m) acc[r][m] = 0.f;

        constexpr int U = (R == 2) ? 4 : 2;  // steps batched per load group: U*R 16-byte loads in flight per lane
        for (int s0 = 0; s0 < nsteps; s0 += U) {
            uint4 w[U][R];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int c = (s0 + u) * 32 + lane;
                const bool ok = (s0 + u < nsteps) && (c < kchunks);
#pragma unroll
                for (int r = 0; r < R; ++r) w[u][r] = ok ? ld_stream(wrow[r] + c) : make_uint4(0, 0, 0, 0);
            }
#pragma unroll
            for (int u = 0; u < U; ++u) {
                if (s0 + u < nsteps) {
                    const int i4 = ((s0 + u) * 2) * 32 + lane;
#pragma unroll
                    for (int m = 0; m < M; ++m) {
                        const float4 xl = xs4[m * m_stride4 + i4];
                        const float4 xh = xs4[m * m_stride4 + i4 + 32];
#pragma unroll
                        for (int r = 0; r < R; ++r) {
                            const uint4 ww = w[u][r];
                            const float2 w0 = unpack2<WBF16>(ww.x), w1 = unpack2<WBF16>(ww.y), w2 = unpack2<WBF16>(ww.z),
                                         w3 = unpack2<WBF16>(ww.w);
                            float a = acc[r][m];
                            a = fmaf(w0.x, xl.x, a);
                            a = fmaf(w0.y, xl.y, a);
                            a = fmaf(w1.x, xl.z, a);
                            a = fmaf(w1.y, xl.w, a);
                            a = fmaf(w2.x, xh.x, a);
                            a = fmaf(w2.y, xh.y, a);
                            a = fmaf(w3.x, xh.z, a);
                            a = fmaf(w3.y, xh.w, a);
                            acc[r][m] = a;
                        }
                    }
                }
            }
        }
#pragma unroll
        for (int r = 0; r < R; ++r)
#pragma unroll
            for (int m = 0; m < M; ++m) acc[r][m] = warp_sum(acc[r][m]);

        // ---- epilogues (every lane holds every reduced value; lane (r*M+m) / (pair*M+m) writes)
        if (EPI == EPI_STORE || EPI == EPI_RESID) {
#pragma unroll
            for (int r = 0; r < R; ++r)
#pragma unroll
                for (int m = 0; m < M; ++m)
                    if (lane == r * M + m && item * R + r < p.N) {
                        float* o = p.out + (size_t)m * p.N + rows[r];
                        *o = (EPI == EPI_RESID) ? (*o + acc[r][m]) : acc[r][m];
                    }
        } else if (EPI == EPI_SWIGLU) {
#pragma unroll
            for (int pr = 0; pr < R / 2; ++pr)
#pragma unroll
                for (int m = 0; m < M; ++m)
                    if (lane == pr * M + m) {
                        const int tile = rows[2 * pr] >> 7, j = rows[2 * pr] & 63;
                        p.out[(size_t)m * (p.N / 2) + tile * 64 + j] = silu(acc[2 * pr][m]) * acc[2 * pr + 1][m];
                    }
        } else if (EPI == EPI_QKV) {
#pragma unroll
            for (int pr = 0; pr < R / 2; ++pr)
#pragma unroll
                for (int m = 0; m < M; ++m)
                    if (lane == pr * M + m) {
                        const int head = rows[2 * pr] >> 7, j = rows[2 * pr] & 63;
                        const float a = acc[2 * pr][m], b = acc[2 * pr + 1][m];
                        if (head < p.n_heads + p.n_kv) {  // q or k head: rotate (rotate_half convention)
                            const float c = p.cos[m * kHalfDim + j], s = p.sin[m * kHalfDim + j];
                            const float lo = a * c - b * s, hi = b * c + a * s;
                            if (head < p.n_heads) {
                                float* q = p.q_out + (size_t)m * p.n_heads * kHeadDim + head * kHeadDim + j;
                                q[0] = lo;
                                q[kHalfDim] = hi;
                            } else {
                                __nv_bfloat16* kd = p.k_pool + p.kv_off[m] + (int64_t)(head - p.n_heads) * p.head_stride + j;
                                kd[0] = __float2bfloat16_rn(lo);
                                kd[kHalfDim] = __float2bfloat16_rn(hi);
                            }
                        } else {
                            __nv_bfloat16* vd = p.v_pool + p.kv_off[m] + (int64_t)(head - p.n_heads - p.n_kv) * p.head_stride + j;
                            vd[0] = __float2bfloat16_rn(a);
                            vd[kHalfDim] = __float2bfloat16_rn(b);
                        }
                    }
        }
    }
}

template <int M, int R>
static void gemv_dispatch(const GemvParams& p, int epi, bool norm, int grid, size_t smem, cudaStream_t st) {
#define B200_GEMV_CASE(E, NRM, WB)                                                                         \
    {                                                                                                      \
        auto kfn = gemv_kernel<M, R, E, NRM, WB>;                                                          \
        cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);                 \
        kfn<<<grid, kGemvThreads, smem, st>>>(p);                                                          \
    }
    if (epi == EPI_STORE) {
        if (p.w_bf16) B200_GEMV_CASE(EPI_STORE, false, true)
        else if (norm) B200_GEMV_CASE(EPI_STORE, true, false)
        else B200_GEMV_CASE(EPI_STORE, false, false)
    }
    else if (epi == EPI_RESID) B200_GEMV_CASE(EPI_RESID, false, false)
    else if (epi == EPI_QKV) B200_GEMV_CASE(EPI_QKV, true, false)
    else B200_GEMV_CASE(EPI_SWIGLU, true, false)
#undef B200_GEMV_CASE
}

void launch_gemv(const GemvParams& p, int M, int epi, bool norm, int num_sms, cudaStream_t st) {
    const int nsteps = (p.K / 8 + 31) / 32;
    const size_t smem = (size_t)M * nsteps * 256 * sizeof(float);
    const int ctas_per_sm = (M == 1 && smem <= 100 * 1024) ? 2 : 1;
    const int grid = num_sms * ctas_per_sm;
    switch (M) {
        case 1: gemv_dispatch<1, 2>(p, epi, norm, grid, smem, st); break;
        case 2: gemv_dispatch<2, 4>(p, epi, norm, grid, smem, st); break;
        case 3: gemv_dispatch<3, 4>(p, epi, norm, grid, smem, st); break;
        default: gemv_dispatch<4, 4>(p, epi, norm, grid, smem, st); break;
    }
}

// =========================================================================== KV bookkeeping
__global__ void kv_init_kernel(KvDev kv) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < kv.n_pages) kv.free_stack[i] = kv.n_pages - 1 - i;  // pop order 0,1,2,...
    if (i < kv.max_slots) { kv.seq_lens[i] = 0; kv.pages_held[i] = 0; }
    if (i < kv.max_slots * kv.max_pages_per_slot) kv.block_table[i] = -1;
    if (i == 0) { *kv.free_top = kv.n_pages; *kv.error = 0; }
}
void launch_kv_init(const KvDev& kv, cudaStream_t st) {
    int n = kv.n_pages;
    if (kv.max_slots * kv.max_pages_per_slot > n) n = kv.max_slots * kv.max_pages_per_slot;
    kv_init_kernel<<<(n + 255) / 256, 256, 0, st>>>(kv);
}

// one thread: deterministic page hand-out for the rows of this call; commits the new lengths
__global__ void kv_alloc_kernel(KvDev kv, SlotList slots, int B, int S, int32_t* call_base) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    int top = *kv.free_top;
    for (int b = 0; b < B; ++b) {
        const int slot = slots.v[b];
        const int base = kv.seq_lens[slot];
        call_base[b] = base;
        const int need = (base + S + kv.page_size - 1) / kv.page_size;
        int held = kv.pages_held[slot];
        if (need > kv.max_pages_per_slot) { *kv.error = 2; continue; }
        while (held < need) {
            if (top <= 0) { *kv.error = 1; break; }
            kv.block_table[slot * kv.max_pages_per_slot + held] = kv.free_stack[--top];
            ++held;
        }
        kv.pages_held[slot] = held;
        kv.seq_lens[slot] = base + S;
    }
    *kv.free_top = top;
}
void launch_kv_alloc(const KvDev& kv, const SlotList& slots, int B, int S, int32_t* call_base, cudaStream_t st) {
    kv_alloc_kernel<<<1, 32, 0, st>>>(kv, slots, B, S, call_base);
}

__global__ void kv_reset_kernel(KvDev kv, int slot) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    int top = *kv.free_top;
    const int lo = slot < 0 ? 0 : slot, hi = slot < 0 ? kv.max_slots : slot + 1;
    for (int s = lo; s < hi; ++s) {
        for (int i = kv.pages_held[s] - 1; i >= 0; --i) {
            kv.free_stack[top++] = kv.block_table[s * kv.max_pages_per_slot + i];
            kv.block_table[s * kv.max_pages_per_slot + i] = -1;
        }
        kv.pages_held[s] = 0;
        kv.seq_lens[s] = 0;
    }
    *kv.free_top = top;
}
void launch_kv_reset(const KvDev& kv, int slot, cudaStream_t st) { kv_reset_kernel<<<1, 32, 0, st>>>(kv, slot); }

// per-token metadata: position (build_position_ids, utils/forwarding_utils.py:24), page address,
// visible key count, and the RoPE table (LlamaRotaryEmbedding.forward, modeling_llama.py:124-135,
// kept in fp32: the reference rounds cos/sin to the activation dtype, the fp32 oracle does not).
__global__ void prep_kernel(KvDev kv, SlotList slots, int B, int S, int causal, InvFreq inv, int n_kv,
                            TokMeta tm) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    const int M = B * S;
    const int m = idx / kHalfDim, j = idx % kHalfDim;
    if (m >= M) return;
    const int b = m / S, s = m % S;
    const int base = tm.call_base[b];
    const int pos = base + s;
    // inv_freq = 1 / theta^(2j/d) comes from the host (compute_default_rope_parameters,
    // modeling_llama.py:116-118); angle = inv_freq * pos in fp32 like the reference's matmul.
    const float ang = (float)pos * inv.v[j];
    float sn, cs;
    sincosf(ang, &sn, &cs);
    tm.cos[(size_t)m * kHalfDim + j] = cs;
    tm.sin[(size_t)m * kHalfDim + j] = sn;
    if (j == 0) {
        const int slot = slots.v[b];
        const int page = kv.block_table[slot * kv.max_pages_per_slot + pos / kv.page_size];
        tm.pos[m] = pos;
        tm.t_end[m] = causal ? pos + 1 : base + S;
        tm.kv_off[m] = ((int64_t)page * n_kv * kv.page_size + (pos % kv.page_size)) * kHeadDim;
    }
}
void launch_prep(const KvDev& kv, const SlotList& slots, int B, int S, int causal, const InvFreq& inv, int n_kv,
                 const TokMeta& tm, cudaStream_t st) {
    const int64_t n = (int64_t)B * S * kHalfDim;
    prep_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(kv, slots, B, S, causal, inv, n_kv, tm);
}

// =========================================================================== small bandwidth kernels
__global__ void load_residual_kernel(const __nv_bfloat16* __restrict__ in, float* __restrict__ x, int64_t n8) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; i < n8; i += stride) {
        const uint4 v = reinterpret_cast<const uint4*>(in)[i];
        float4* o = reinterpret_cast<float4*>(x) + 2 * i;
        o[0] = make_float4(bf_lo(v.x), bf_hi(v.x), bf_lo(v.y), bf_hi(v.y));
        o[1] = make_float4(bf_lo(v.z), bf_hi(v.z), bf_lo(v.w), bf_hi(v.w));
    }
}
void launch_load_residual(const __nv_bfloat16* in, float* x, int64_t n, cudaStream_t st) {
    const int64_t n8 = n / 8;
    int grid = (int)((n8 + 255) / 256);
    if (grid > 148 * 8) grid = 148 * 8;
    if (grid < 1) grid = 1;
    load_residual_kernel<<<grid, 256, 0, st>>>(in, x, n8);
}

// One CTA per row; fp32 in, bf16 out; 128-bit loads/stores.  n % 8 == 0.
// LlamaRMSNorm.forward (modeling_llama.py:62-67): mean of squares in fp32, rsqrt(var+eps), * weight.
__device__ __forceinline__ uint32_t pack_f16_sat(float a, float b) {
    a = fminf(fmaxf(a, -65504.f), 65504.f);
    b = fminf(fmaxf(b, -65504.f), 65504.f);
    __half2 t = __floats2half2_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&t);
}
template <bool F16OUT>
__global__ void __launch_bounds__(256) rmsnorm_bf16_kernel(const float* __restrict__ x,
                                                           const __nv_bfloat16* __restrict__ w, float eps, int n,
                                                           __nv_bfloat16* __restrict__ out) {
    __shared__ float red[33];
    const float* xr = x + (size_t)blockIdx.x * n;
    __nv_bfloat16* orow = out + (size_t)blockIdx.x * n;
    const int n8 = n / 8;
    float rstd = 1.f;
    if (w != nullptr) {
        float ss = 0.f;
        for (int i = threadIdx.x; i < n8; i += blockDim.x) {
            const float4 a = reinterpret_cast<const float4*>(xr)[2 * i], b = reinterpret_cast<const float4*>(xr)[2 * i + 1];
            ss += a.x * a.x + a.y * a.y + a.z * a.z + a.w * a.w + b.x * b.x + b.y * b.y + b.z * b.z + b.w * b.w;
        }
        ss = block_sum(ss, red);
        rstd = rsqrtf(ss / (float)n + eps);
    }
    for (int i = threadIdx.x; i < n8; i += blockDim.x) {
        const float4 a = reinterpret_cast<const float4*>(xr)[2 * i], b = reinterpret_cast<const float4*>(xr)[2 * i + 1];
        float v[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
        if (w != nullptr) {
            const uint4 ww = reinterpret_cast<const uint4*>(w)[i];
            const float wf[8] = {bf_lo(ww.x), bf_hi(ww.x), bf_lo(ww.y), bf_hi(ww.y),
                                 bf_lo(ww.z), bf_hi(ww.z), bf_lo(ww.w), bf_hi(ww.w)};
#pragma unroll
            for (int e = 0; e < 8; ++e) v[e] = v[e] * rstd * wf[e];
        }
        uint4 o;
        if (F16OUT) {
            o.x = pack_f16_sat(v[0], v[1]); o.y = pack_f16_sat(v[2], v[3]);
            o.z = pack_f16_sat(v[4], v[5]); o.w = pack_f16_sat(v[6], v[7]);
        } else {
            o.x = pack_bf16(v[0], v[1]); o.y = pack_bf16(v[2], v[3]);
            o.z = pack_bf16(v[4], v[5]); o.w = pack_bf16(v[6], v[7]);
        }
        reinterpret_cast<uint4*>(orow)[i] = o;
    }
}
void launch_rmsnorm_bf16(const float* x, const __nv_bfloat16* norm_w, float eps, int rows, int n,
                         void* out, cudaStream_t st, bool f16_out) {
    if (rows <= 0) return;
    if (f16_out) rmsnorm_bf16_kernel<true><<<rows, 256, 0, st>>>(x, norm_w, eps, n, reinterpret_cast<__nv_bfloat16*>(out));
    else rmsnorm_bf16_kernel<false><<<rows, 256, 0, st>>>(x, norm_w, eps, n, reinterpret_cast<__nv_bfloat16*>(out));
}

__global__ void __launch_bounds__(256) lastrow_norm_f32_kernel(const float* __restrict__ x,
                                                               const __nv_bfloat16* __restrict__ w, float eps, int S,
                                                               int n, float* __restrict__ out) {
    __shared__ float red[33];
    const float* xr = x + ((size_t)blockIdx.x * S + (S - 1)) * n;
    float rstd = 1.f;
    if (w != nullptr) {
        float ss = 0.f;
        for (int i = threadIdx.x; i < n; i += blockDim.x) ss += xr[i] * xr[i];
        ss = block_sum(ss, red);
        rstd = rsqrtf(ss / (float)n + eps);
    }
    for (int i = threadIdx.x; i < n; i += blockDim.x)
        out[(size_t)blockIdx.x * n + i] = (w != nullptr) ? xr[i] * rstd * __bfloat162float(w[i]) : xr[i];
}
void launch_lastrow_norm_f32(const float* x, const __nv_bfloat16* norm_w, float eps, int B, int S, int n,
                             float* out, cudaStream_t st) {
    lastrow_norm_f32_kernel<<<B, 256, 0, st>>>(x, norm_w, eps, S, n, out);
}

__global__ void lastrow_bf16_to_f32_kernel(const __nv_bfloat16* __restrict__ h, int S, int n, float* __restrict__ out) {
    const __nv_bfloat16* r = h + ((size_t)blockIdx.x * S + (S - 1)) * n;
    for (int i = threadIdx.x; i < n; i += blockDim.x) out[(size_t)blockIdx.x * n + i] = __bfloat162float(r[i]);
}
void launch_lastrow_bf16_to_f32(const __nv_bfloat16* h, int B, int S, int n, float* out, cudaStream_t st) {
    lastrow_bf16_to_f32_kernel<<<B, 256, 0, st>>>(h, S, n, out);
}

// torch.argmax(logits[:, -1, :], dim=-1) (utils/node_worker.py:264): first index of the maximum.
__global__ void __launch_bounds__(1024) argmax_kernel(const float* __restrict__ logits, int V, int64_t* tok,
                                                      float* margin) {
    __shared__ float sv[32], s2[32];
    __shared__ int si[32];
    const float* row = logits + (size_t)blockIdx.x * V;
    float best = -INFINITY, second = -INFINITY;
    int bi = 0x7fffffff;
    for (int i = threadIdx.x; i < V; i += blockDim.x) {
        const float v = row[i];
        if (v > best) { second = best; best = v; bi = i; }
        else if (v > second) second = v;
    }
    auto merge = [](float& b, int& i, float& s, float ob, int oi, float os) {
        if (ob > b || (ob == b && oi < i)) { s = fmaxf(fmaxf(b, os), s); b = ob; i = oi; }
        else s = fmaxf(s, ob);
    };
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float ob = __shfl_xor_sync(0xffffffffu, best, o);
        const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
        const float os = __shfl_xor_sync(0xffffffffu, second, o);
        merge(best, bi, second, ob, oi, os);
    }
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    if (lane == 0) { sv[w] = best; si[w] = bi; s2[w] = second; }
    __syncthreads();
    if (w == 0) {
        const int nw = blockDim.x >> 5;
        best = lane < nw ? sv[lane] : -INFINITY;
        bi = lane < nw ? si[lane] : 0x7fffffff;
        second = lane < nw ? s2[lane] : -INFINITY;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float ob = __shfl_xor_sync(0xffffffffu, best, o);
            const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
            const float os = __shfl_xor_sync(0xffffffffu, second, o);
            merge(best, bi, second, ob, oi, os);
        }
        if (lane == 0) {
            tok[blockIdx.x] = (int64_t)bi;
            if (margin != nullptr) margin[blockIdx.x] = best - second;
        }
    }
}
void launch_argmax(const float* logits, int B, int V, int64_t* tok, float* margin, cudaStream_t st) {
    argmax_kernel<<<B, 1024, 0, st>>>(logits, V, tok, margin);
}

// embed_tokens(ids) (utils/node_worker.py:215,302): row gather, 128-bit copies.
__global__ void embed_kernel(const __nv_bfloat16* __restrict__ table, const int64_t* __restrict__ ids, int H, int V,
                             __nv_bfloat16* __restrict__ out) {
    int64_t id = ids[blockIdx.x];
    if (id < 0) id = 0;
    if (id >= V) id = V - 1;
    const uint4* src = reinterpret_cast<const uint4*>(table + (size_t)id * H);
    uint4* dst = reinterpret_cast<uint4*>(out + (size_t)blockIdx.x * H);
    for (int i = threadIdx.x; i < H / 8; i += blockDim.x) dst[i] = src[i];
}
void launch_embed(const __nv_bfloat16* table, const int64_t* ids, int n, int H, int V, __nv_bfloat16* out,
                  cudaStream_t st) {
    if (n > 0) embed_kernel<<<n, 128, 0, st>>>(table, ids, H, V, out);
}

// ---- weight ingest: .pth storage dtype -> bf16, and the paired gate/up layout
__device__ __forceinline__ float load_as_float(const void* src, int dtype, int64_t i) {
    if (dtype == 0) return __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(src)[i]);
    if (dtype == 1) return reinterpret_cast<const float*>(src)[i];
    return __half2float(reinterpret_cast<const __half*>(src)[i]);
}
__device__ __forceinline__ __half to_half_sat(float v) { return __float2half_rn(fminf(fmaxf(v, -65504.f), 65504.f)); }
__global__ void convert_kernel(const void* src, int dtype, void* dst, int64_t n, bool f16_out) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; i < n; i += stride) {
        const float v = load_as_float(src, dtype, i);
        if (f16_out) reinterpret_cast<__half*>(dst)[i] = to_half_sat(v);
        else reinterpret_cast<__nv_bfloat16*>(dst)[i] = __float2bfloat16_rn(v);
    }
}
void launch_convert_to_bf16(const void* src, int dtype, void* dst, int64_t n, cudaStream_t st, bool f16_out) {
    int64_t g = (n + 255) / 256;
    if (g > 148 * 16) g = 148 * 16;
    if (g < 1) g = 1;
    convert_kernel<<<(int)g, 256, 0, st>>>(src, dtype, dst, n, f16_out);
}
__global__ void pack_gate_up_kernel(const void* src, int dtype, __half* dst, int rows, int cols, int half) {
    const int r = blockIdx.x;
    const int64_t drow = (int64_t)(r / 64) * 128 + half * 64 + (r % 64);
    for (int c = threadIdx.x; c < cols; c += blockDim.x)
        dst[drow * cols + c] = to_half_sat(load_as_float(src, dtype, (int64_t)r * cols + c));
}
void launch_pack_gate_up(const void* src, int dtype, void* dst, int rows, int cols, int half, cudaStream_t st) {
    pack_gate_up_kernel<<<rows, 256, 0, st>>>(src, dtype, reinterpret_cast<__half*>(dst), rows, cols, half);
}

// =========================================================================== decode attention (split-K)
// eager_attention_forward (modeling_llama.py:199-221) for a handful of query rows: for each
// (token m, head h) the cached keys [0, t_end[m]) are split into n_splits chunks; each CTA
// computes an un-normalised partial (running max, sum, acc[128]) with warp-shuffle reductions,
// the combine kernel merges them.  K/V are read once, 128-bit, from the paged pool.
constexpr int kAttnThreads = 128;
constexpr int kPartStride = kHeadDim + 2;

__global__ void __launch_bounds__(kAttnThreads) attn_decode_split_kernel(const AttnDecodeParams p) {
    extern __shared__ __align__(16) float sm[];  // q[128] | red[40] | scores[chunk]
    float* qs = sm;
    float* red = sm + kHeadDim;
    float* sc = sm + kHeadDim + 40;
    const int split = blockIdx.x, h = blockIdx.y, m = blockIdx.z;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int T = p.t_end[m];
    const int chunk = (T + p.n_splits - 1) / p.n_splits;
    const int t0 = split * chunk, t1 = min(T, t0 + chunk);
    float* part = p.part + (((size_t)m * p.n_heads + h) * p.n_splits + split) * kPartStride;
    if (t0 >= t1) {  // empty split
        if (tid < kHeadDim) part[tid] = 0.f;
        if (tid == 0) { part[kHeadDim] = -INFINITY; part[kHeadDim + 1] = 0.f; }
        return;
    }
    const int b = m / p.S;
    const int slot = p.slots.v[b];
    const int kvh = h / (p.n_heads / p.n_kv);
    const int32_t* bt = p.block_table + (size_t)slot * p.max_pages_per_slot;
    const int64_t head_stride = (int64_t)p.page_size * kHeadDim;
    const int64_t page_stride = (int64_t)p.n_kv * head_stride;
    qs[tid] = p.q[((size_t)m * p.n_heads + h) * kHeadDim + tid];
    __syncthreads();
    const float scale_log2 = 0.08838834764831845f * 1.4426950408889634f;  // 128^-0.5 * log2(e)

    // phase 1: one key per thread
    float lmax = -INFINITY;
    for (int t = t0 + tid; t < t1; t += kAttnThreads) {
        const int page = bt[t / p.page_size];
        const uint4* kr = reinterpret_cast<const uint4*>(p.k_pool + page * page_stride + kvh * head_stride +
                                                         (int64_t)(t % p.page_size) * kHeadDim);
        float d = 0.f;
#pragma unroll
        for (int c = 0; c < 16; ++c) {
            const uint4 kk = kr[c];
            const float4 q0 = reinterpret_cast<const float4*>(qs)[2 * c], q1 = reinterpret_cast<const float4*>(qs)[2 * c + 1];
            d = fmaf(bf_lo(kk.x), q0.x, d); d = fmaf(bf_hi(kk.x), q0.y, d);
            d = fmaf(bf_lo(kk.y), q0.z, d); d = fmaf(bf_hi(kk.y), q0.w, d);
            d = fmaf(bf_lo(kk.z), q1.x, d); d = fmaf(bf_hi(kk.z), q1.y, d);
            d = fmaf(bf_lo(kk.w), q1.z, d); d = fmaf(bf_hi(kk.w), q1.w, d);
        }
        d *= scale_log2;
        sc[t - t0] = d;
        lmax = fmaxf(lmax, d);
    }
    lmax = warp_max(lmax);
    if (lane == 0) red[warp] = lmax;
    __syncthreads();
    float mx = fmaxf(fmaxf(red[0], red[1]), fmaxf(red[2], red[3]));
    // phase 2: weights.  HF casts softmax probabilities to the model dtype before P.V; here they
    // stay fp32 (the fp32 oracle is the target).
    float lsum = 0.f;
    for (int t = t0 + tid; t < t1; t += kAttnThreads) {
        const float w = exp2f(sc[t - t0] - mx);
        sc[t - t0] = w;
        lsum += w;
    }
    lsum = warp_sum(lsum);
    __syncthreads();
    if (lane == 0) red[8 + warp] = lsum;
    __syncthreads();
    const float l = red[8] + red[9] + red[10] + red[11];
    // phase 3: P.V — thread = (key group g of 8, 16-byte dim chunk c of 16)
    const int g = tid >> 4, c = tid & 15;
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int t = t0 + g; t < t1; t += 8) {
        const int page = bt[t / p.page_size];
        const uint4 vv = reinterpret_cast<const uint4*>(p.v_pool + page * page_stride + kvh * head_stride +
                                                        (int64_t)(t % p.page_size) * kHeadDim)[c];
        const float w = sc[t - t0];
        acc[0] = fmaf(w, bf_lo(vv.x), acc[0]); acc[1] = fmaf(w, bf_hi(vv.x), acc[1]);
        acc[2] = fmaf(w, bf_lo(vv.y), acc[2]); acc[3] = fmaf(w, bf_hi(vv.y), acc[3]);
        acc[4] = fmaf(w, bf_lo(vv.z), acc[4]); acc[5] = fmaf(w, bf_hi(vv.z), acc[5]);
        acc[6] = fmaf(w, bf_lo(vv.w), acc[6]); acc[7] = fmaf(w, bf_hi(vv.w), acc[7]);
    }
    // reduce the 8 key groups: lanes (g even/odd within a warp) by shuffle, then across 4 warps in smem
#pragma unroll
    for (int e = 0; e < 8; ++e) acc[e] += __shfl_xor_sync(0xffffffffu, acc[e], 16);
    __syncthreads();  // scores no longer needed
    float* stage = sc;  // needs 4*128 floats: guaranteed by the launcher (chunk smem >= 512 floats)
    if (lane < 16) {
#pragma unroll
        for (int e = 0; e < 8; ++e) stage[warp * kHeadDim + c * 8 + e] = acc[e];
    }
    __syncthreads();
    if (tid < kHeadDim) part[tid] = stage[tid] + stage[kHeadDim + tid] + stage[2 * kHeadDim + tid] + stage[3 * kHeadDim + tid];
    if (tid == 0) { part[kHeadDim] = mx; part[kHeadDim + 1] = l; }
}

__global__ void __launch_bounds__(kHeadDim) attn_decode_combine_kernel(const AttnDecodeParams p) {
    const int h = blockIdx.x, m = blockIdx.y, tid = threadIdx.x;
    const float* part = p.part + ((size_t)m * p.n_heads + h) * p.n_splits * kPartStride;
    float mx = -INFINITY;
    for (int s = 0; s < p.n_splits; ++s) mx = fmaxf(mx, part[s * kPartStride + kHeadDim]);
    float num = 0.f, den = 0.f;
    for (int s = 0; s < p.n_splits; ++s) {
        const float pm = part[s * kPartStride + kHeadDim];
        const float w = (pm == -INFINITY) ? 0.f : exp2f(pm - mx);
        num = fmaf(w, part[s * kPartStride + tid], num);
        den = fmaf(w, part[s * kPartStride + kHeadDim + 1], den);
    }
    p.out[((size_t)m * p.n_heads + h) * kHeadDim + tid] = num / den;
}

void launch_attn_decode(const AttnDecodeParams& p, int M, cudaStream_t st) {
    // smem: q + red + scores for the largest possible chunk is unknown on the host when the call is
    // replayed from a graph (T grows), so size for the slot capacity.
    const int64_t max_T = (int64_t)p.max_pages_per_slot * p.page_size;
    int chunk = (int)((max_T + p.n_splits - 1) / p.n_splits);  // <= kMaxChunkKeys by the caller's choice of n_splits
    if (chunk < 4 * kHeadDim) chunk = 4 * kHeadDim;
    const size_t smem = (size_t)(kHeadDim + 40 + chunk) * sizeof(float);
    cudaFuncSetAttribute(attn_decode_split_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    dim3 grid(p.n_splits, p.n_heads, M);
    attn_decode_split_kernel<<<grid, kAttnThreads, smem, st>>>(p);
    dim3 grid2(p.n_heads, M);
    attn_decode_combine_kernel<<<grid2, kHeadDim, 0, st>>>(p);
}

// =========================================================================== decode attention, allocation-free
// b200_attn_decode_static: the same split-K decode attention as b200_attn_decode, for callers that cannot
// allocate device memory at all.
//
// Who that is: (1) a host runtime that has already handed every byte of HBM to its own pool (a serving process
// whose KV pages and weights fill the card, or a torch process whose caching allocator owns the memory) — a
// cudaMalloc for the split-K partials would either fail or have to come out of somebody else's budget;
// (2) a latency-sensitive loop (one call per generated token) that must not risk the device-wide
// synchronisation cudaMalloc / cudaFree imply; (3) a tool that probes attention over K/V pages it already owns
// (a profiler, a parity checker) and has no shard handle whose workspace it could borrow.
//
// How: the partials live in the file-scope __device__ array below instead of caller memory.  Every GPU that
// loads this library carries the array (kStaticScratchBytes, about 2 MiB) until the process exits.  The kernels
// are the ones b200_attn_decode launches; only the scratch pointer differs.
//
// Rules for callers:
//  * Capacity: rows * n_heads * n_splits <= kStaticPartials (4096).  Larger problems are refused with
//    B200_EINVAL and a message; use b200_attn_decode with your own scratch for those.
//  * The first call on each device resolves the array's device address (cudaGetSymbolAddress, with that device
//    current) and caches it.  Make that first call before you begin any stream capture on that device.
//  * The array is shared by every call in the process and keeps whatever the previous call left in it.
//    Nothing in it has to start at zero: each (row, head, split) partial is written by the split kernel —
//    empty splits included — before the combine kernel reads it, so there is nothing to clear between calls.
//  * Because it is shared, two calls whose kernels can overlap on the GPU (different streams, or different
//    host threads on different streams) race on it and corrupt each other.  Issue them on one stream, or order
//    them with events.
//  * No cudaMalloc / cudaFree / allocating library routine is called, here or in the launcher.
constexpr int kStaticPartials = 4096;
constexpr size_t kStaticScratchFloats = (size_t)kStaticPartials * kPartStride;
constexpr size_t kStaticScratchBytes = kStaticScratchFloats * sizeof(float);  // 2,129,920 bytes
static __device__ __align__(256) float g_attn_static_scratch[kStaticScratchFloats];

constexpr int kMaxDevices = 64;
static float* g_static_scratch_dev[kMaxDevices];  // per device: resolved DEVICE address of the array above
static std::mutex g_static_scratch_mu;

// In host code `g_attn_static_scratch` is only a handle for the symbol, not an address a kernel may use.
static cudaError_t static_scratch_for_current_device(float** out) {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= kMaxDevices) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(g_static_scratch_mu);
    if (g_static_scratch_dev[dev] == nullptr) {
        float* ptr = nullptr;
        e = cudaGetSymbolAddress((void**)&ptr, g_attn_static_scratch);
        if (e != cudaSuccess) return e;
        if (ptr == nullptr) return cudaErrorInvalidSymbol;
        g_static_scratch_dev[dev] = ptr;
    }
    *out = g_static_scratch_dev[dev];
    return cudaSuccess;
}

}  // namespace b200

extern "C" __attribute__((visibility("default"))) int b200_attn_decode_static(
    const float* q, const void* k_pool, const void* v_pool, const int32_t* block_table, const int32_t* t_end, int rows,
    int n_heads, int n_kv_heads, int page_size, int max_pages_per_row, int n_splits, float* out, void* stream) {
    using namespace b200;
    char msg[256];
    if (q == nullptr || k_pool == nullptr || v_pool == nullptr || block_table == nullptr || t_end == nullptr || out == nullptr) {
        set_last_error("[ERROR] null argument");
        return -1;
    }
    if (rows < 1 || rows > kMaxBatch || n_heads < 1 || n_kv_heads < 1 || n_heads % n_kv_heads != 0 || page_size < 64 ||
        page_size % 64 != 0 || max_pages_per_row < 1 || n_splits < 1 || n_splits > kMaxSplits ||
        ((int64_t)max_pages_per_row * page_size + n_splits - 1) / n_splits > kMaxChunkKeys) {
        snprintf(msg, sizeof(msg), "[ERROR] attn_decode_static: need 1 <= rows <= %d, n_heads %% n_kv_heads == 0, page_size %% 64 == 0, "
                 "1 <= n_splits <= %d and at most %d keys per split", kMaxBatch, kMaxSplits, kMaxChunkKeys);
        set_last_error(msg);
        return -1;
    }
    const int64_t partials = (int64_t)rows * n_heads * n_splits;
    if (partials > kStaticPartials) {
        snprintf(msg, sizeof(msg), "[ERROR] attn_decode_static: rows*n_heads*n_splits = %lld exceeds the built-in scratch (%d partials, "
                 "%zu bytes); use b200_attn_decode with caller-provided scratch", (long long)partials, kStaticPartials, kStaticScratchBytes);
        set_last_error(msg);
        return -1;
    }
    float* scratch = nullptr;
    cudaError_t e = static_scratch_for_current_device(&scratch);
    if (e != cudaSuccess) {
        snprintf(msg, sizeof(msg), "[ERROR] CUDA: could not resolve the built-in scratch array on this device: %s", cudaGetErrorString(e));
        set_last_error(msg);
        return -3;
    }
    AttnDecodeParams a{};
    a.q = q; a.k_pool = reinterpret_cast<const __nv_bfloat16*>(k_pool); a.v_pool = reinterpret_cast<const __nv_bfloat16*>(v_pool);
    a.block_table = block_table; a.t_end = t_end;
    for (int i = 0; i < rows; ++i) a.slots.v[i] = i;
    a.S = 1; a.n_heads = n_heads; a.n_kv = n_kv_heads; a.page_size = page_size; a.max_pages_per_slot = max_pages_per_row;
    a.n_splits = n_splits; a.part = scratch; a.out = out;
    launch_attn_decode(a, rows, (cudaStream_t)stream);
    e = cudaGetLastError();
    if (e != cudaSuccess) {
        snprintf(msg, sizeof(msg), "[ERROR] CUDA: %s (attn_decode_static launch)", cudaGetErrorString(e));
        set_last_error(msg);
        return -3;
    }
    return 0;
}

namespace b200 {

// =========================================================================== RoPE + KV append (GEMM path)
// apply_rotary_pos_emb (modeling_llama.py:146-168) on the fused QKV GEMM output, then
// DynamicCache.update (cache_utils.py:102-121) as a paged write instead of a re-allocating cat.
// One CTA per token; thread handles 8 consecutive dims of the first half and the matching 8 of the second.
__global__ void __launch_bounds__(256) rope_append_kernel(const RopeAppendParams p) {
    const int m = blockIdx.x;
    const int n_all = p.n_heads + 2 * p.n_kv;
    const __nv_bfloat16* row = p.qkv + (size_t)m * n_all * kHeadDim;
    const float* cs = p.cos + (size_t)m * kHalfDim;
    const float* sn = p.sin + (size_t)m * kHalfDim;
    const int64_t off = p.kv_off[m];
    for (int w = threadIdx.x; w < n_all * 8; w += blockDim.x) {
        const int head = w >> 3, j0 = (w & 7) * 8;
        const uint4 lo4 = *reinterpret_cast<const uint4*>(row + head * kHeadDim + j0);
        const uint4 hi4 = *reinterpret_cast<const uint4*>(row + head * kHeadDim + kHalfDim + j0);
        float a[8] = {bf_lo(lo4.x), bf_hi(lo4.x), bf_lo(lo4.y), bf_hi(lo4.y), bf_lo(lo4.z), bf_hi(lo4.z), bf_lo(lo4.w), bf_hi(lo4.w)};
        float b[8] = {bf_lo(hi4.x), bf_hi(hi4.x), bf_lo(hi4.y), bf_hi(hi4.y), bf_lo(hi4.z), bf_hi(hi4.z), bf_lo(hi4.w), bf_hi(hi4.w)};
        uint4 olo = lo4, ohi = hi4;
        if (head < p.n_heads + p.n_kv) {
            float lo[8], hi[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) {
                const float c = cs[j0 + e], s = sn[j0 + e];
                lo[e] = a[e] * c - b[e] * s;
                hi[e] = b[e] * c + a[e] * s;
            }
            olo = make_uint4(pack_bf16(lo[0], lo[1]), pack_bf16(lo[2], lo[3]), pack_bf16(lo[4], lo[5]), pack_bf16(lo[6], lo[7]));
            ohi = make_uint4(pack_bf16(hi[0], hi[1]), pack_bf16(hi[2], hi[3]), pack_bf16(hi[4], hi[5]), pack_bf16(hi[6], hi[7]));
        }
        __nv_bfloat16* dst;
        if (head < p.n_heads) dst = p.q_out + ((size_t)m * p.n_heads + head) * kHeadDim;
        else if (head < p.n_heads + p.n_kv) dst = p.k_pool + off + (int64_t)(head - p.n_heads) * p.head_stride;
        else dst = p.v_pool + off + (int64_t)(head - p.n_heads - p.n_kv) * p.head_stride;
        *reinterpret_cast<uint4*>(dst + j0) = olo;
        *reinterpret_cast<uint4*>(dst + kHalfDim + j0) = ohi;
    }
}
void launch_rope_append(const RopeAppendParams& p, int M, cudaStream_t st) {
    if (M > 0) rope_append_kernel<<<M, 256, 0, st>>>(p);
}

}  // namespace b200
