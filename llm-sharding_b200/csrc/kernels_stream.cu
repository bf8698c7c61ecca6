// HBM-bound kernels of the shard path: the weight-streaming skinny GEMM used for decode
// (fused RMSNorm prologue; RoPE + paged-KV append / residual / SwiGLU epilogues), RMSNorm,
// embedding gather, arg-max, KV page bookkeeping.  Replaces the per-op ATen launches under
// HF LlamaDecoderLayer (transformers modeling_llama.py:62-67,138-168,182-184,303-332) that the
// reference calls from utils/shard_loader.py:66-74.
#include <stdio.h>
#include <stdlib.h>

#include <mutex>

#include "../../include/b200shard.h"
#include "common.cuh"
#include "internal.h"

namespace b200 {

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
__device__ __forceinline__ float load_as_float(const void* src, int dtype, int64_t i) {
    if (dtype == 0) return __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(src)[i]);
    if (dtype == 1) return reinterpret_cast<const float*>(src)[i];
    return __half2float(reinterpret_cast<const __half*>(src)[i]);
}
// Everything that precedes the first layer of a single-token call for a few rows, in ONE launch instead of four: page
// hand-out (kv_alloc_kernel), positions / RoPE table / page addresses (prep_kernel), and the fp32 residual stream either
// from the incoming hidden state (load_residual) or straight from the embedding table (embed_tokens,
// utils/node_worker.py:302).  One CTA: thread 0 hands out the pages, the barrier publishes call_base to the rest.
__global__ void __launch_bounds__(1024) decode_prologue_kernel(KvDev kv, SlotList slots, int B, int causal, InvFreq inv, int n_kv,
                                                               TokMeta tm, const void* __restrict__ in, int in_dtype,
                                                               const int64_t* __restrict__ tok_in,
                                                               const __nv_bfloat16* __restrict__ table, int V, int H,
                                                               float* __restrict__ x) {
    if (threadIdx.x == 0) {
        int top = *kv.free_top;
        for (int b = 0; b < B; ++b) {
            const int slot = slots.v[b];
            const int base = kv.seq_lens[slot];
            tm.call_base[b] = base;
            const int need = (base + 1 + kv.page_size - 1) / kv.page_size;
            int held = kv.pages_held[slot];
            if (need > kv.max_pages_per_slot) { *kv.error = 2; continue; }
            while (held < need) {
                if (top <= 0) { *kv.error = 1; break; }
                kv.block_table[slot * kv.max_pages_per_slot + held] = kv.free_stack[--top];
                ++held;
            }
            kv.pages_held[slot] = held;
            kv.seq_lens[slot] = base + 1;
        }
        *kv.free_top = top;
    }
    __syncthreads();
    for (int idx = threadIdx.x; idx < B * kHalfDim; idx += blockDim.x) {   // prep_kernel for S == 1
        const int m = idx / kHalfDim, j = idx % kHalfDim;
        const int pos = tm.call_base[m];
        float sn, cs;
        sincosf((float)pos * inv.v[j], &sn, &cs);
        tm.cos[(size_t)m * kHalfDim + j] = cs;
        tm.sin[(size_t)m * kHalfDim + j] = sn;
        if (j == 0) {
            const int page = kv.block_table[slots.v[m] * kv.max_pages_per_slot + pos / kv.page_size];
            tm.pos[m] = pos;
            tm.t_end[m] = pos + 1;   // causal or not: a single new token sees every cached key and itself
            tm.kv_off[m] = ((int64_t)page * n_kv * kv.page_size + (pos % kv.page_size)) * kHeadDim;
        }
    }
    (void)causal;
    const int n8 = B * H / 8;
    for (int i = threadIdx.x; i < n8; i += blockDim.x) {
        float v[8];
        if (tok_in != nullptr) {
            const int m = (i * 8) / H, c = (i * 8) % H;
            int64_t id = tok_in[m];
            id = id < 0 ? 0 : (id >= V ? V - 1 : id);
            const uint4 w = *reinterpret_cast<const uint4*>(table + (size_t)id * H + c);
            v[0] = bf_lo(w.x); v[1] = bf_hi(w.x); v[2] = bf_lo(w.y); v[3] = bf_hi(w.y);
            v[4] = bf_lo(w.z); v[5] = bf_hi(w.z); v[6] = bf_lo(w.w); v[7] = bf_hi(w.w);
        } else {
#pragma unroll
            for (int e = 0; e < 8; ++e) v[e] = load_as_float(in, in_dtype, (int64_t)i * 8 + e);
        }
        float4* o = reinterpret_cast<float4*>(x) + 2 * i;
        o[0] = make_float4(v[0], v[1], v[2], v[3]);
        o[1] = make_float4(v[4], v[5], v[6], v[7]);
    }
}
void launch_decode_prologue(const KvDev& kv, const SlotList& slots, int B, int causal, const InvFreq& inv, int n_kv, const TokMeta& tm,
                            const void* in, int in_dtype, const int64_t* tok_in, const __nv_bfloat16* table, int V, int H, float* x,
                            cudaStream_t st) {
    decode_prologue_kernel<<<1, 1024, 0, st>>>(kv, slots, B, causal, inv, n_kv, tm, in, in_dtype, tok_in, table, V, H, x);
}

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
// same for fp16 / fp32 messages (the reference runs in whatever dtype the node was started with, start_node.py:16-17)
__global__ void load_residual_any_kernel(const void* __restrict__ in, int dtype, float* __restrict__ x, int64_t n) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; i < n; i += stride)
        x[i] = dtype == 1 ? reinterpret_cast<const float*>(in)[i] : __half2float(reinterpret_cast<const __half*>(in)[i]);
}
void launch_load_residual(const void* in_any, int dtype, float* x, int64_t n, cudaStream_t st) {
    if (dtype != 0) {
        int grid = (int)((n + 255) / 256);
        if (grid > 148 * 16) grid = 148 * 16;
        load_residual_any_kernel<<<grid < 1 ? 1 : grid, 256, 0, st>>>(in_any, dtype, x, n);
        return;
    }
    const __nv_bfloat16* in = reinterpret_cast<const __nv_bfloat16*>(in_any);
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
template <int OUT>  // element type of `out`: 0 bf16, 1 fp32, 2 fp16 (saturating) — the B200_DT_* codes
__global__ void __launch_bounds__(256) rmsnorm_bf16_kernel(const float* __restrict__ x,
                                                           const __nv_bfloat16* __restrict__ w, float eps, int n,
                                                           __nv_bfloat16* __restrict__ out, int normalize,
                                                           float* __restrict__ ss_out) {
    __shared__ float red[33];
    // programmatic dependent launch: x comes from the previous kernel; the next kernel (a weight-streaming GEMM in the
    // batched-decode path) may start requesting its weights while this one runs
    asm volatile("griddepcontrol.wait;" ::: "memory");
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    const float* xr = x + (size_t)blockIdx.x * n;
    __nv_bfloat16* orow = out + (size_t)blockIdx.x * n * (OUT == 1 ? 2 : 1);
    const int n8 = n / 8;
    float rstd = 1.f;
    if (w != nullptr && (normalize || ss_out != nullptr)) {
        float ss = 0.f;
        for (int i = threadIdx.x; i < n8; i += blockDim.x) {
            const float4 a = reinterpret_cast<const float4*>(xr)[2 * i], b = reinterpret_cast<const float4*>(xr)[2 * i + 1];
            ss += a.x * a.x + a.y * a.y + a.z * a.z + a.w * a.w + b.x * b.x + b.y * b.y + b.z * b.z + b.w * b.w;
        }
        ss = block_sum(ss, red);
        if (normalize) rstd = rsqrtf(ss / (float)n + eps);
        if (ss_out != nullptr && threadIdx.x == 0) ss_out[blockIdx.x] = ss;   // the row's sum of squares, as one partial (row_ss_parts == 1)
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
        if (OUT == 1) {
            float4* of = reinterpret_cast<float4*>(orow) + 2 * i;
            of[0] = make_float4(v[0], v[1], v[2], v[3]);
            of[1] = make_float4(v[4], v[5], v[6], v[7]);
            continue;
        } else if (OUT == 2) {
            o.x = pack_f16_sat(v[0], v[1]); o.y = pack_f16_sat(v[2], v[3]);
            o.z = pack_f16_sat(v[4], v[5]); o.w = pack_f16_sat(v[6], v[7]);
        } else {
            o.x = pack_bf16(v[0], v[1]); o.y = pack_bf16(v[2], v[3]);
            o.z = pack_bf16(v[4], v[5]); o.w = pack_bf16(v[6], v[7]);
        }
        reinterpret_cast<uint4*>(orow)[i] = o;
    }
}
static void launch_rmsnorm_any(const float* x, const __nv_bfloat16* norm_w, float eps, int rows, int n,
                               void* out, cudaStream_t st, int out_dtype, float* ss_out) {
    if (rows <= 0) return;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(rows);
    cfg.blockDim = dim3(256);
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    const char* e = getenv("B200SHARD_PDL");
    cfg.numAttrs = (e != nullptr && e[0] == '0') ? 0 : 1;
    __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(out);
    const int normalize = eps >= 0.f ? 1 : 0;
    if (out_dtype == 2) cudaLaunchKernelEx(&cfg, rmsnorm_bf16_kernel<2>, x, norm_w, eps, n, o, normalize, ss_out);
    else if (out_dtype == 1) cudaLaunchKernelEx(&cfg, rmsnorm_bf16_kernel<1>, x, norm_w, eps, n, o, normalize, ss_out);
    else cudaLaunchKernelEx(&cfg, rmsnorm_bf16_kernel<0>, x, norm_w, eps, n, o, normalize, ss_out);
}
void launch_rmsnorm_bf16(const float* x, const __nv_bfloat16* norm_w, float eps, int rows, int n,
                         void* out, cudaStream_t st, int out_dtype) {
    launch_rmsnorm_any(x, norm_w, eps, rows, n, out, st, out_dtype, nullptr);
}
// x * norm_w -> fp16 without the 1 / rms factor (the consuming GEMM applies it to its finished dot products); ss_out (may be
// null): [rows] sums of squares for that consumer
void launch_scale_f16(const float* x, const __nv_bfloat16* norm_w, int rows, int n, __half* out, float* ss_out, cudaStream_t st) {
    launch_rmsnorm_any(x, norm_w, -1.f, rows, n, out, st, 2, ss_out);
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

__global__ void lastrow_to_f32_kernel(const void* __restrict__ h, int dtype, int S, int n, float* __restrict__ out) {
    const size_t r0 = ((size_t)blockIdx.x * S + (S - 1)) * n;
    for (int i = threadIdx.x; i < n; i += blockDim.x) out[(size_t)blockIdx.x * n + i] = load_as_float(h, dtype, (int64_t)(r0 + i));
}
void launch_lastrow_to_f32(const void* h, int dtype, int B, int S, int n, float* out, cudaStream_t st) {
    lastrow_to_f32_kernel<<<B, 256, 0, st>>>(h, dtype, S, n, out);
}

// torch.argmax(logits[:, -1, :], dim=-1) (utils/node_worker.py:264): first index of the maximum.
__global__ void __launch_bounds__(1024) argmax_kernel(const float* __restrict__ logits, int V, int64_t* tok,
                                                      float* margin) {
    __shared__ float sv[32], s2[32];
    __shared__ int si[32];
    asm volatile("griddepcontrol.wait;" ::: "memory");  // launched programmatically behind the lm_head kernel
    const float* row = logits + (size_t)blockIdx.x * V;
    float best = -INFINITY, second = -INFINITY;
    int bi = 0x7fffffff;
    auto take = [&](float v, int i) {  // a thread visits its indices in increasing order: '>' keeps the first maximum
        if (v > best) { second = best; best = v; bi = i; }
        else if (v > second) second = v;
    };
    if ((V & 3) == 0 && (reinterpret_cast<uintptr_t>(row) & 15) == 0) {
        // one decode step waits on this kernel: 128-bit loads, eight in flight per thread, instead of a chain of scalar ones
        const float4* row4 = reinterpret_cast<const float4*>(row);
        const int n4 = V >> 2;
        for (int i0 = threadIdx.x; i0 < n4; i0 += 8 * (int)blockDim.x) {
            float4 v[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                const int i = i0 + u * (int)blockDim.x;
                v[u] = (i < n4) ? __ldcg(row4 + i) : make_float4(-INFINITY, -INFINITY, -INFINITY, -INFINITY);
            }
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                const int i = 4 * (i0 + u * (int)blockDim.x);
                if (i < V) { take(v[u].x, i); take(v[u].y, i + 1); take(v[u].z, i + 2); take(v[u].w, i + 3); }
            }
        }
    } else {
        for (int i = threadIdx.x; i < V; i += blockDim.x) take(row[i], i);
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
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(B);
    cfg.blockDim = dim3(1024);
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    const char* e = getenv("B200SHARD_PDL");
    cfg.numAttrs = (e != nullptr && e[0] == '0') ? 0 : 1;
    cudaLaunchKernelEx(&cfg, argmax_kernel, logits, V, tok, margin);
}

// arg-max over a RANGE of the vocabulary (the vocabulary-split tail of the ring: the tail stage scores rows [0, split) of
// lm_head, its successor the rest).  logits: [B][n] scores of vocabulary rows offset .. offset + n - 1.  The range's best and
// second-best values and the best index are merged with an incoming partial result (in_*: another, LOWER range's, may be
// null) exactly as torch.argmax over the whole row would decide (first index of the maximum), and go out as a partial
// result (out_best / out_second / out_idx) and / or as the final token id and its top-1 margin.
__global__ void __launch_bounds__(1024) argmax_range_kernel(const float* __restrict__ logits, int n, int offset, const float* in_best,
                                                            const float* in_second, const int64_t* in_idx, float* out_best,
                                                            float* out_second, int64_t* out_idx, float* out_margin) {
    __shared__ float sv[32], s2[32];
    __shared__ int si[32];
    asm volatile("griddepcontrol.wait;" ::: "memory");
    const float* row = logits + (size_t)blockIdx.x * n;
    float best = -INFINITY, second = -INFINITY;
    int bi = 0x7fffffff;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {   // increasing indices per thread: '>' keeps the first maximum
        const float v = __ldcg(row + i);
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
            int64_t idx = (int64_t)bi + offset;
            if (in_best != nullptr) {   // the incoming range lies below this one: it wins ties
                const float ib = in_best[blockIdx.x], is = in_second[blockIdx.x];
                if (ib >= best) { second = fmaxf(is, best); best = ib; idx = in_idx[blockIdx.x]; }
                else second = fmaxf(second, ib);
            }
            if (out_best != nullptr) { out_best[blockIdx.x] = best; out_second[blockIdx.x] = second; }
            if (out_idx != nullptr) out_idx[blockIdx.x] = idx;
            if (out_margin != nullptr) out_margin[blockIdx.x] = best - second;
        }
    }
}
void launch_argmax_range(const float* logits, int B, int n, int offset, const float* in_best, const float* in_second, const int64_t* in_idx,
                         float* out_best, float* out_second, int64_t* out_idx, float* out_margin, cudaStream_t st) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(B);
    cfg.blockDim = dim3(1024);
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    const char* e = getenv("B200SHARD_PDL");
    cfg.numAttrs = (e != nullptr && e[0] == '0') ? 0 : 1;
    cudaLaunchKernelEx(&cfg, argmax_range_kernel, logits, n, offset, in_best, in_second, in_idx, out_best, out_second, out_idx, out_margin);
}

// embed_tokens(ids) (utils/node_worker.py:215,302): row gather, 128-bit copies.
__global__ void embed_any_kernel(const __nv_bfloat16* __restrict__ table, const int64_t* __restrict__ ids, int H, int V,
                                 void* __restrict__ out, int dtype) {
    int64_t id = ids[blockIdx.x];
    if (id < 0) id = 0;
    if (id >= V) id = V - 1;
    const __nv_bfloat16* src = table + (size_t)id * H;
    for (int i = threadIdx.x; i < H; i += blockDim.x) {
        const float v = __bfloat162float(src[i]);
        if (dtype == 1) reinterpret_cast<float*>(out)[(size_t)blockIdx.x * H + i] = v;
        else reinterpret_cast<__half*>(out)[(size_t)blockIdx.x * H + i] = __float2half_rn(fminf(fmaxf(v, -65504.f), 65504.f));
    }
}
__global__ void embed_kernel(const __nv_bfloat16* __restrict__ table, const int64_t* __restrict__ ids, int H, int V,
                             __nv_bfloat16* __restrict__ out) {
    int64_t id = ids[blockIdx.x];
    if (id < 0) id = 0;
    if (id >= V) id = V - 1;
    const uint4* src = reinterpret_cast<const uint4*>(table + (size_t)id * H);
    uint4* dst = reinterpret_cast<uint4*>(out + (size_t)blockIdx.x * H);
    for (int i = threadIdx.x; i < H / 8; i += blockDim.x) dst[i] = src[i];
}
void launch_embed(const __nv_bfloat16* table, const int64_t* ids, int n, int H, int V, void* out, int out_dtype,
                  cudaStream_t st) {
    if (n <= 0) return;
    if (out_dtype == 0) embed_kernel<<<n, 128, 0, st>>>(table, ids, H, V, reinterpret_cast<__nv_bfloat16*>(out));
    else embed_any_kernel<<<n, 128, 0, st>>>(table, ids, H, V, out, out_dtype);
}

// ---- weight ingest: .pth storage dtype -> bf16, and the paired gate/up layout
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
// eager_attention_forward (modeling_llama.py:199-221) for a handful of query rows.  For each (token m, head h) the
// cached keys [0, t_end[m]) are split into n_splits chunks, one CTA each.  Inside a CTA every half-warp owns one key
// at a time: its 16 lanes read the key's 256-byte K row and V row as one 16-byte chunk each, the dot product is reduced
// with four warp shuffles, and softmax runs online in fp32 (running max / sum / rescaled accumulator), so K and V are
// read once, together, with 8 x 16-byte loads in flight per lane.  The eight half-warp partials are merged in shared
// memory, the CTA's partial goes to scratch, and the last CTA of the (m, h) group to arrive (atomic ticket) merges the
// n_splits partials and writes the output — one launch, no separate combine pass.
// Scratch layout (floats): [rows*n_heads*n_splits][130] partials (acc[128], max, sum) followed by rows*n_heads
// tickets (uint32).  Tickets must be zero when the kernel starts; the last CTA puts its ticket back to zero.
constexpr int kAttnThreads = 128;
constexpr int kPartStride = kHeadDim + 2;

// Blackwell's packed fp32 pipe: one instruction, two independent fp32 FMAs (same rounding as two scalar FMAs).  A bf16
// pair unpacks straight into such a register pair: (v << 16, v & 0xffff0000).
__device__ __forceinline__ uint64_t f32x2(float lo, float hi) {
    uint64_t r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ uint64_t bf16x2_to_f32x2(uint32_t v) { return f32x2(bf_lo(v), bf_hi(v)); }
__device__ __forceinline__ uint64_t fma2(uint64_t a, uint64_t b, uint64_t c) {
    uint64_t d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}
__device__ __forceinline__ uint64_t mul2(uint64_t a, uint64_t b) {
    uint64_t d;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}
__device__ __forceinline__ float sum2(uint64_t v) {
    float lo, hi;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
    return lo + hi;
}
__device__ __forceinline__ float dot8(const uint4& k, const uint64_t (&q2)[4]) {
    uint64_t d = mul2(bf16x2_to_f32x2(k.x), q2[0]);
    d = fma2(bf16x2_to_f32x2(k.y), q2[1], d);
    d = fma2(bf16x2_to_f32x2(k.z), q2[2], d);
    d = fma2(bf16x2_to_f32x2(k.w), q2[3], d);
    return sum2(d);
}

constexpr int kAttnCluster = 8;
constexpr int kAttnClusterThreads = 256;

// Merge of the NG group partials of one CTA (g_acc / g_m / g_l, already published with __syncthreads) and of the n_splits
// CTAs of one (row, head): over distributed shared memory inside a cluster, through global scratch + arrival ticket
// otherwise, or not at all when the CTA saw every key.  Shared by the CUDA-core and the tensor-core kernels.
template <int NG, int CL>  // CL: CTAs of the cluster that shares one (row, head) — 2, 4 or 8; 0: no cluster (global scratch + ticket)
__device__ __forceinline__ void attn_decode_merge(const AttnDecodeParams& p, const float (*g_acc)[kHeadDim], const float* g_m,
                                                  const float* g_l, float (*c_part)[kPartStride], int& s_last, int split, int h,
                                                  int m, int tid) {
    float num = 0.f, den = 0.f, mx = -INFINITY;
    if (tid < kHeadDim) {
#pragma unroll
        for (int g = 0; g < NG; ++g) mx = fmaxf(mx, g_m[g]);
#pragma unroll
        for (int g = 0; g < NG; ++g) {
            const float w = (g_m[g] == -INFINITY) ? 0.f : exp2f(g_m[g] - mx);
            num = fmaf(w, g_acc[g][tid], num);
            den = fmaf(w, g_l[g], den);
        }
    }
    if (CL > 0) {
        uint32_t rank;
        asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(rank));
        // completes the arrive issued at kernel entry: every CTA of the cluster (rank 0 above all) has started, so its
        // shared memory exists and may be written from here
        asm volatile("barrier.cluster.wait.aligned;" ::: "memory");
        if (tid < kHeadDim) {
            uint32_t dst;
            asm volatile("mapa.shared::cluster.u32 %0, %1, 0;" : "=r"(dst) : "r"((uint32_t)__cvta_generic_to_shared(&c_part[rank][0])));
            asm volatile("st.shared::cluster.f32 [%0], %1;" ::"r"(dst + 4u * (uint32_t)tid), "f"(num) : "memory");
            if (tid == 0) {
                asm volatile("st.shared::cluster.f32 [%0], %1;" ::"r"(dst + 4u * kHeadDim), "f"(mx) : "memory");
                asm volatile("st.shared::cluster.f32 [%0], %1;" ::"r"(dst + 4u * (kHeadDim + 1)), "f"(den) : "memory");
            }
        }
        // every thread of every CTA of the cluster arrives (empty splits included); rank 0 outlives all remote stores
        asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
        asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
        if (rank != 0 || tid >= kHeadDim) return;
        float gmx = -INFINITY;
#pragma unroll
        for (int s = 0; s < (CL > 0 ? CL : 1); ++s) gmx = fmaxf(gmx, c_part[s][kHeadDim]);
        float n2 = 0.f, d2 = 0.f;
#pragma unroll
        for (int s = 0; s < (CL > 0 ? CL : 1); ++s) {
            const float pm = c_part[s][kHeadDim];
            const float w = (pm == -INFINITY) ? 0.f : exp2f(pm - gmx);
            n2 = fmaf(w, c_part[s][tid], n2);
            d2 = fmaf(w, c_part[s][kHeadDim + 1], d2);
        }
        p.out[((size_t)m * p.n_heads + h) * kHeadDim + tid] = n2 / d2;
        if (p.out_h != nullptr) p.out_h[((size_t)m * p.n_heads + h) * kHeadDim + tid] = __float2half_rn(n2 / d2);
        return;
    }
    if (p.n_splits == 1) {  // the CTA saw every key of its (row, head): nothing to merge (many rows: one CTA per row and head)
        if (tid < kHeadDim) {
            p.out[((size_t)m * p.n_heads + h) * kHeadDim + tid] = num / den;
            if (p.out_h != nullptr) p.out_h[((size_t)m * p.n_heads + h) * kHeadDim + tid] = __float2half_rn(num / den);
        }
        return;
    }
    float* part = p.part + (((size_t)m * p.n_heads + h) * p.n_splits + split) * kPartStride;
    if (tid < kHeadDim) {
        part[tid] = num;
        if (tid == 0) { part[kHeadDim] = mx; part[kHeadDim + 1] = den; }
    }
    // ---- the last CTA of this (m, h) to finish merges the splits
    unsigned int* ticket = p.tickets + ((size_t)m * p.n_heads + h);
    __threadfence();
    __syncthreads();
    if (tid == 0) {
        const unsigned int prev = atomicAdd(ticket, 1u);
        s_last = (prev == (unsigned int)p.n_splits - 1u);
        if (s_last) *ticket = 0u;  // ready for the next launch (graph replays included)
    }
    __syncthreads();
    if (!s_last || tid >= kHeadDim) return;
    __threadfence();
    const float* all = p.part + ((size_t)m * p.n_heads + h) * p.n_splits * kPartStride;
    float gmx = -INFINITY;
    for (int s = 0; s < p.n_splits; ++s) gmx = fmaxf(gmx, __ldcg(all + s * kPartStride + kHeadDim));
    float n2 = 0.f, d2 = 0.f;
    for (int s = 0; s < p.n_splits; ++s) {
        const float pm = __ldcg(all + s * kPartStride + kHeadDim);
        const float w = (pm == -INFINITY) ? 0.f : exp2f(pm - gmx);
        n2 = fmaf(w, __ldcg(all + s * kPartStride + tid), n2);
        d2 = fmaf(w, __ldcg(all + s * kPartStride + kHeadDim + 1), d2);
    }
    p.out[((size_t)m * p.n_heads + h) * kHeadDim + tid] = n2 / d2;
    if (p.out_h != nullptr) p.out_h[((size_t)m * p.n_heads + h) * kHeadDim + tid] = __float2half_rn(n2 / d2);
}

// CLUSTER = true: the n_splits CTAs of one (m, h) form a thread-block cluster (n_splits == kAttnCluster); the partials
// travel to cluster rank 0 through distributed shared memory and one hardware cluster barrier replaces the global
// scratch round trip (store, fence, atomic ticket, reload) — the longest links of this latency-bound kernel's chain.

template <int NT, int CL>
__global__ void __launch_bounds__(NT) attn_decode_kernel(const AttnDecodeParams p) {
    constexpr int NG = NT / 16;  // half-warps: keys processed side by side
    __shared__ float g_acc[NG][kHeadDim];
    __shared__ float g_m[NG], g_l[NG];
    __shared__ int s_last;
    __shared__ float c_part[CL > 0 ? CL : 1][kPartStride];  // rank 0's copy gathers the cluster's partials
    // distributed shared memory may only be touched once the target CTA is known to be running: arrive now (non-blocking),
    // wait right before the remote stores in attn_decode_merge — by then the wait is free
    if (CL > 0) asm volatile("barrier.cluster.arrive.relaxed.aligned;" ::: "memory");
    const int split = blockIdx.x, h = blockIdx.y, m = blockIdx.z;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int l16 = lane & 15, grp = warp * 2 + (lane >> 4);  // NG half-warps, 16 lanes x 8 dims each
    const int T = p.t_end[m];
    const int chunk = (T + p.n_splits - 1) / p.n_splits;
    const int t0 = split * chunk, t1 = min(T, t0 + chunk);
    const int b = m / p.S;
    const int slot = p.slots.v[b];
    const int kvh = h / (p.n_heads / p.n_kv);
    const int32_t* bt = p.block_table + (size_t)slot * p.max_pages_per_slot;
    const int64_t head_stride = (int64_t)p.page_size * kHeadDim;
    const int64_t page_stride = (int64_t)p.n_kv * head_stride;
    const float scale_log2 = 0.08838834764831845f * 1.4426950408889634f;  // 128^-0.5 * log2(e)
    // Keys cached before this forward call do not depend on the kernel that precedes us (which only appends the new
    // tokens' K/V and produces q): request the first window of them BEFORE griddepcontrol.wait, so that their DRAM
    // latency hides behind the predecessor's tail.  t_end / block_table / call_base were written by kernels that
    // completed before the predecessor could even start its own dependent part.
    const int t_old = (p.call_base != nullptr) ? min(t1, p.call_base[b]) : 0;
    constexpr int U = 8;  // keys in flight per half-warp
    // power-of-two page sizes (the usual 64) index with a shift and a mask instead of two integer divisions per key
    const int page_shift = (p.page_size & (p.page_size - 1)) == 0 ? 31 - __clz(p.page_size) : -1;
    const int64_t lane_base = kvh * head_stride + l16 * 8;
    auto kv_offset = [&](int t) -> int64_t {
        int pi, in_page;
        if (page_shift >= 0) { pi = t >> page_shift; in_page = t & (p.page_size - 1); }
        else { pi = t / p.page_size; in_page = t - pi * p.page_size; }
        return bt[pi] * page_stride + lane_base + (int64_t)in_page * kHeadDim;
    };
    uint4 kk[U], vv[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
        const int t = t0 + grp + NG * u;
        if (t < t_old) {
            const int64_t off = kv_offset(t);
            kk[u] = *reinterpret_cast<const uint4*>(p.k_pool + off);
            vv[u] = *reinterpret_cast<const uint4*>(p.v_pool + off);
        } else {
            kk[u] = make_uint4(0, 0, 0, 0);
            vv[u] = make_uint4(0, 0, 0, 0);
        }
    }
    asm volatile("griddepcontrol.wait;" ::: "memory");            // q and the new K/V come from the previous kernel
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    uint64_t q[4];  // the lane's 8 query values, pre-scaled, as four fp32 pairs
    {
        const float4* qp = reinterpret_cast<const float4*>(p.q + ((size_t)m * p.n_heads + h) * kHeadDim + l16 * 8);
        const float4 a = qp[0], c = qp[1];
        q[0] = f32x2(a.x * scale_log2, a.y * scale_log2); q[1] = f32x2(a.z * scale_log2, a.w * scale_log2);
        q[2] = f32x2(c.x * scale_log2, c.y * scale_log2); q[3] = f32x2(c.z * scale_log2, c.w * scale_log2);
    }
    float m_run = -INFINITY, l_run = 0.f;
    uint64_t acc2[4] = {0ull, 0ull, 0ull, 0ull};  // 8 fp32 accumulators as four pairs (bit pattern 0 = +0.0f)
    auto load_key = [&](int u, int t) {
        const int64_t off = kv_offset(t);
        kk[u] = *reinterpret_cast<const uint4*>(p.k_pool + off);
        vv[u] = *reinterpret_cast<const uint4*>(p.v_pool + off);
    };
    // the keys of the first window that could not be requested before the wait (this call's own tokens)
#pragma unroll
    for (int u = 0; u < U; ++u) {
        const int t = t0 + grp + NG * u;
        if (t >= t_old && t < t1) load_key(u, t);
    }
    // Keys are consumed in half-windows of HB = 4: the four dot products and their 16-lane reductions are independent
    // (instruction-level parallelism for a kernel that runs at low occupancy), the softmax state is updated once per
    // half-window (one rescale of the accumulator instead of four), and as soon as a half-window is consumed its
    // registers are refilled with the keys of the next pass, so loads stay in flight while the other half is computed.
    // Trip counts are uniform across the CTA (the shuffles need every lane of the warp); validity is per key.
    constexpr int HB = 4;
    for (int tb0 = t0; tb0 < t1; tb0 += NG * U) {
        const int tb = tb0 + grp;
#pragma unroll
        for (int hb = 0; hb < U; hb += HB) {
            float d[HB];
#pragma unroll
            for (int i = 0; i < HB; ++i) d[i] = dot8(kk[hb + i], q);
#pragma unroll
            for (int o = 8; o > 0; o >>= 1)
#pragma unroll
                for (int i = 0; i < HB; ++i) d[i] += __shfl_xor_sync(0xffffffffu, d[i], o);
            float bmax = -INFINITY;
#pragma unroll
            for (int i = 0; i < HB; ++i)
                if (tb + NG * (hb + i) < t1) bmax = fmaxf(bmax, d[i]);
            if (bmax > -INFINITY) {  // uniform within the half-warp; no shuffles inside
                const float m_new = fmaxf(m_run, bmax);
                const float alpha = exp2f(m_run - m_new);
                l_run *= alpha;
                const uint64_t alpha2 = f32x2(alpha, alpha);
#pragma unroll
                for (int e = 0; e < 4; ++e) acc2[e] = mul2(acc2[e], alpha2);
#pragma unroll
                for (int i = 0; i < HB; ++i) {
                    if (tb + NG * (hb + i) < t1) {
                        const float w = exp2f(d[i] - m_new);
                        l_run += w;
                        const uint4 v = vv[hb + i];
                        const uint64_t w2 = f32x2(w, w);
                        acc2[0] = fma2(w2, bf16x2_to_f32x2(v.x), acc2[0]);
                        acc2[1] = fma2(w2, bf16x2_to_f32x2(v.y), acc2[1]);
                        acc2[2] = fma2(w2, bf16x2_to_f32x2(v.z), acc2[2]);
                        acc2[3] = fma2(w2, bf16x2_to_f32x2(v.w), acc2[3]);
                    }
                }
                m_run = m_new;
            }
#pragma unroll
            for (int i = 0; i < HB; ++i) {  // refill: the same positions of the next pass
                const int t = tb + NG * (hb + i) + NG * U;
                if (t < t1) load_key(hb + i, t);
            }
        }
    }
    // ---- merge the half-warp partials of this CTA
    float acc[8];
#pragma unroll
    for (int e = 0; e < 4; ++e) asm("mov.b64 {%0, %1}, %2;" : "=f"(acc[2 * e]), "=f"(acc[2 * e + 1]) : "l"(acc2[e]));
#pragma unroll
    for (int e = 0; e < 8; ++e) g_acc[grp][l16 * 8 + e] = acc[e];
    if (l16 == 0) { g_m[grp] = m_run; g_l[grp] = l_run; }
    __syncthreads();
    attn_decode_merge<NG, CL>(p, g_acc, g_m, g_l, c_part, s_last, split, h, m, tid);
}

static bool attn_cluster_enabled() {
    static int v = -1;
    if (v < 0) {
        const char* e = getenv("B200SHARD_ATTN_CLUSTER");
        v = (e != nullptr && e[0] == '0') ? 0 : 1;
    }
    return v == 1;
}
int attn_decode_cluster_splits() { return attn_cluster_enabled() ? kAttnCluster : 0; }

void launch_attn_decode(const AttnDecodeParams& p_in, int M, cudaStream_t st) {
    AttnDecodeParams p = p_in;
    if (p.tickets == nullptr)  // default placement: right behind this call's partials
        p.tickets = reinterpret_cast<unsigned int*>(p.part + (size_t)M * p.n_heads * p.n_splits * kPartStride);
    // the CTAs that share one (row, head) form a thread-block cluster when there are 2, 4 or 8 of them
    const bool cluster = attn_cluster_enabled() && (p.n_splits == 2 || p.n_splits == 4 || p.n_splits == 8);
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(p.n_splits, p.n_heads, M);
    // 256-thread CTAs (128 keys per pass) while the whole grid is a single wave of at most two CTAs per SM; beyond that
    // 128-thread CTAs pack four to an SM and finish in fewer waves (4 rows x 8 splits: 16.3 us against 19.8 us)
    static int wide_env = -1;
    if (wide_env < 0) { const char* e = getenv("B200SHARD_ATTN_WIDE"); wide_env = e ? atoi(e) : 2; }
    const int64_t n_cta = (int64_t)p.n_splits * p.n_heads * M;
    const bool wide = cluster && (wide_env == 1 || (wide_env == 2 && n_cta <= 2 * 148));
    cfg.blockDim = dim3(wide ? kAttnClusterThreads : kAttnThreads);
    cfg.dynamicSmemBytes = 0;
    cfg.stream = st;
    cudaLaunchAttribute attr[2];
    int na = 0;
    if (cluster) {
        attr[na].id = cudaLaunchAttributeClusterDimension;
        attr[na].val.clusterDim.x = p.n_splits;
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
#define B200_ATTN_LAUNCH(NT_, CL_) cudaLaunchKernelEx(&cfg, attn_decode_kernel<NT_, CL_>, p)
    if (!cluster) B200_ATTN_LAUNCH(kAttnThreads, 0);
    else if (wide) {
        if (p.n_splits == 8) B200_ATTN_LAUNCH(kAttnClusterThreads, 8);
        else if (p.n_splits == 4) B200_ATTN_LAUNCH(kAttnClusterThreads, 4);
        else B200_ATTN_LAUNCH(kAttnClusterThreads, 2);
    } else {
        if (p.n_splits == 8) B200_ATTN_LAUNCH(kAttnThreads, 8);
        else if (p.n_splits == 4) B200_ATTN_LAUNCH(kAttnThreads, 4);
        else B200_ATTN_LAUNCH(kAttnThreads, 2);
    }
#undef B200_ATTN_LAUNCH
}
size_t attn_decode_scratch_floats(int rows, int n_heads, int n_splits) {
    return (size_t)rows * n_heads * n_splits * kPartStride + (size_t)rows * n_heads;
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
//    The partials need no clearing (each one is written before it is read, empty splits included), but the
//    arrival tickets behind them must be zero when the kernel starts: every call zeroes that region on the
//    caller's stream first (cudaMemsetAsync: no allocation), so a call that was aborted cannot poison the next.
//  * Because it is shared, two calls whose kernels can overlap on the GPU (different streams, or different
//    host threads on different streams) race on it and corrupt each other.  Issue them on one stream, or order
//    them with events.
//  * No cudaMalloc / cudaFree / allocating library routine is called, here or in the launcher.
constexpr int kStaticPartials = 4096;
constexpr size_t kStaticScratchFloats = (size_t)kStaticPartials * kPartStride + kStaticPartials;  // partials + tickets
constexpr size_t kStaticScratchBytes = kStaticScratchFloats * sizeof(float);  // 2,146,304 bytes
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

extern "C" B200_API int b200_attn_decode_static(
    const float* q, const void* k_pool, const void* v_pool, const int32_t* block_table, const int32_t* t_end, int rows,
    int n_heads, int n_kv_heads, int page_size, int max_pages_per_row, int n_splits, float* out, void* stream) {
    using namespace b200;
    char msg[256];
    if (q == nullptr || k_pool == nullptr || v_pool == nullptr || block_table == nullptr || t_end == nullptr || out == nullptr) {
        set_last_error("[ERROR] null argument");
        return B200_EINVAL;
    }
    if (rows < 1 || rows > kMaxBatch || n_heads < 1 || n_kv_heads < 1 || n_heads % n_kv_heads != 0 || page_size < 64 ||
        page_size % 64 != 0 || max_pages_per_row < 1 || n_splits < 1 || n_splits > kMaxSplits ||
        ((int64_t)max_pages_per_row * page_size + n_splits - 1) / n_splits > kMaxChunkKeys) {
        snprintf(msg, sizeof(msg), "[ERROR] attn_decode_static: need 1 <= rows <= %d, n_heads %% n_kv_heads == 0, page_size %% 64 == 0, "
                 "1 <= n_splits <= %d and at most %d keys per split", kMaxBatch, kMaxSplits, kMaxChunkKeys);
        set_last_error(msg);
        return B200_EINVAL;
    }
    const int64_t partials = (int64_t)rows * n_heads * n_splits;
    if (partials > kStaticPartials) {
        snprintf(msg, sizeof(msg), "[ERROR] attn_decode_static: rows*n_heads*n_splits = %lld exceeds the built-in scratch (%d partials, "
                 "%zu bytes); use b200_attn_decode with caller-provided scratch", (long long)partials, kStaticPartials, kStaticScratchBytes);
        set_last_error(msg);
        return B200_EINVAL;
    }
    float* scratch = nullptr;
    cudaError_t e = static_scratch_for_current_device(&scratch);
    if (e != cudaSuccess) {
        snprintf(msg, sizeof(msg), "[ERROR] CUDA: could not resolve the built-in scratch array on this device: %s", cudaGetErrorString(e));
        set_last_error(msg);
        return B200_ECUDA;
    }
    AttnDecodeParams a{};
    a.q = q; a.k_pool = reinterpret_cast<const __nv_bfloat16*>(k_pool); a.v_pool = reinterpret_cast<const __nv_bfloat16*>(v_pool);
    a.block_table = block_table; a.t_end = t_end;
    for (int i = 0; i < rows; ++i) a.slots.v[i] = i;
    a.S = 1; a.n_heads = n_heads; a.n_kv = n_kv_heads; a.page_size = page_size; a.max_pages_per_slot = max_pages_per_row;
    a.n_splits = n_splits; a.part = scratch; a.out = out;
    e = cudaMemsetAsync(scratch + (size_t)partials * kPartStride, 0, (size_t)rows * n_heads * sizeof(unsigned int), (cudaStream_t)stream);
    if (e == cudaSuccess) {
        launch_attn_decode(a, rows, (cudaStream_t)stream);
        e = cudaGetLastError();
    }
    if (e != cudaSuccess) {
        snprintf(msg, sizeof(msg), "[ERROR] CUDA: %s (attn_decode_static launch)", cudaGetErrorString(e));
        set_last_error(msg);
        return B200_ECUDA;
    }
    return B200_OK;
}
