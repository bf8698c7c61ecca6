// The weight-streaming skinny GEMM of the decode path: out[m][n] = sum_k W[n][k] * x[m][k] for M <= 4 token rows,
// 16-bit weights read exactly once at HBM speed.  Replaces the cuBLAS GEMV-shaped calls under q/k/v/o_proj and
// gate/up/down_proj (transformers modeling_llama.py:182-184,262-264,287-288) and lm_head
// (utils/node_worker.py:262) when a shard is asked for a handful of tokens, with the surrounding elementwise work
// fused in: RMSNorm prologue (:62-67), RoPE + paged K/V append (:146-168, cache_utils.py:102-121), residual add
// (:325,331), SiLU*mul (:183).
//
// Layout of the work: the whole CTA (512 threads) reads one weight row at a time — thread t owns the 16-byte chunk
// t (+512 j) of every row, so a row is one perfectly coalesced 8 KB burst and the thread's slice of x never changes.
// Rows are handed to CTAs in contiguous ranges that differ by at most one row (148 CTAs, one per SM: <1 % imbalance
// instead of the 23 % a warp-per-row split costs on 11008 rows).  R = 4 rows are in flight per thread, plus the next
// 4 prefetched: 128 KB of loads outstanding per SM with two CTAs resident.  Per group of rows: warp-shuffle reduction, one __syncthreads, and
// a handful of threads finish the sum and run the epilogue while the rest stream on.
//
// Programmatic dependent launch: the first group's weights are requested BEFORE griddepcontrol.wait (weights do not
// depend on the previous kernel), so a successor CTA that lands on an SM a predecessor CTA just left already has its
// first loads in flight while the predecessor's stragglers drain.
#include <stdlib.h>

#include "common.cuh"
#include "internal.h"

namespace b200 {

constexpr int kGemvThreads = 512;
constexpr int kGemvWarps = kGemvThreads / 32;
constexpr int kR = 4;  // weight rows per group

__device__ __forceinline__ float2 unpack2(uint32_t v) { return __half22float2(*reinterpret_cast<const __half2*>(&v)); }

// shared-memory layout of an activation row: the two float4 halves of chunk c = j*512 + t live at float4 index
// (j*2 + half)*512 + t, so that the CTA's reads of one step are contiguous (no bank conflicts)

template <int M, int EPI, bool NORM>
__global__ void __launch_bounds__(kGemvThreads, (M <= 2 ? 2 : 1)) gemv_kernel(const GemvParams p) {
    extern __shared__ __align__(16) float smem[];
    __shared__ float red[33];
    constexpr bool PAIRED = (EPI == EPI_QKV || EPI == EPI_SWIGLU);
    const int K = p.K, kchunks = K >> 3;
    const int NJ = (kchunks + kGemvThreads - 1) / kGemvThreads;
    const int Kpad = NJ * kGemvThreads * 8;
    float* xs = smem;                               // [M][Kpad]
    float* part = smem + (size_t)M * Kpad;          // [2][warps][kR*M]
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    // contiguous range of row units [ub, ue) for this CTA (paired epilogues: a unit pair = rows j and j+64 of a tile)
    int ub, ue;
    if (PAIRED) {
        const int64_t P = p.N >> 1;
        ub = 2 * (int)(P * blockIdx.x / gridDim.x);
        ue = 2 * (int)(P * (blockIdx.x + 1) / gridDim.x);
    } else {
        ub = (int)((int64_t)p.N * blockIdx.x / gridDim.x);
        ue = (int)((int64_t)p.N * (blockIdx.x + 1) / gridDim.x);
    }
    auto unit_row = [&](int u) -> int {
        if (PAIRED) { const int pr = u >> 1; return ((pr >> 6) << 7) + (pr & 63) + ((u & 1) << 6); }
        return u;
    };
    const int n_groups = (ue - ub + kR - 1) / kR;
    const int total_steps = n_groups * NJ;
    const uint4* W4 = reinterpret_cast<const uint4*>(p.W);
    const int64_t row_stride4 = K >> 3;

    auto load_step = [&](int g, int j, uint4 (&w)[kR]) {
        const int c = j * kGemvThreads + tid;
#pragma unroll
        for (int r = 0; r < kR; ++r) {
            const int u = ub + g * kR + r;
            w[r] = (u < ue && c < kchunks) ? ld_stream(W4 + (int64_t)unit_row(u) * row_stride4 + c) : make_uint4(0, 0, 0, 0);
        }
    };
    uint4 wc[kR], wn[kR];
    if (total_steps > 0) load_step(0, 0, wc);

    // everything above touched only weights; from here on we read what the previous kernel produced
    asm volatile("griddepcontrol.wait;" ::: "memory");
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");

    // ---- prologue: activations -> shared memory, fp32, in ONE pass.  Thread t stages exactly the chunks it will
    // consume itself (c = j*512 + t: 32 contiguous bytes per thread, 16 KB contiguous per CTA), so the main loop needs
    // no barrier after the staging.  With the RMSNorm prologue the row is staged as x * norm_weight and the scalar
    // rsqrt(mean(x^2) + eps) is applied to the finished dot products instead (it commutes with the sum), so the sum of
    // squares is accumulated on the fly; its block reduction is the only synchronisation before the first epilogue.
    __shared__ float rstd_s[4];
#pragma unroll
    for (int m = 0; m < M; ++m) {
        const float* xr = p.x + (size_t)m * p.x_stride;
        float4* xm4 = reinterpret_cast<float4*>(xs + (size_t)m * Kpad);
        float ss = 0.f;
        for (int jj = 0; jj < NJ; ++jj) {
            const int c = jj * kGemvThreads + tid;
            float4 lo = make_float4(0.f, 0.f, 0.f, 0.f), hi = lo;
            if (c < kchunks) {
                lo = *reinterpret_cast<const float4*>(xr + c * 8);
                hi = *reinterpret_cast<const float4*>(xr + c * 8 + 4);
                if (NORM) {
                    ss += lo.x * lo.x + lo.y * lo.y + lo.z * lo.z + lo.w * lo.w + hi.x * hi.x + hi.y * hi.y + hi.z * hi.z + hi.w * hi.w;
                    const uint4 wb = *reinterpret_cast<const uint4*>(p.norm_w + c * 8);
                    lo.x *= bf_lo(wb.x); lo.y *= bf_hi(wb.x); lo.z *= bf_lo(wb.y); lo.w *= bf_hi(wb.y);
                    hi.x *= bf_lo(wb.z); hi.y *= bf_hi(wb.z); hi.z *= bf_lo(wb.w); hi.w *= bf_hi(wb.w);
                }
            }
            xm4[(jj * 2) * kGemvThreads + tid] = lo;
            xm4[(jj * 2 + 1) * kGemvThreads + tid] = hi;
        }
        if (NORM) {
            ss = block_sum(ss, red);
            if (tid == 0) rstd_s[m] = rsqrtf(ss / (float)K + p.eps);
        }
    }
    // no __syncthreads: every thread reads back only what it wrote; rstd_s is published by the first group barrier

    const float4* xs4 = reinterpret_cast<const float4*>(xs);
    const int m_stride4 = Kpad >> 2;
    float acc[kR][M];
#pragma unroll
    for (int r = 0; r < kR; ++r)
#pragma unroll
        for (int m = 0; m < M; ++m) acc[r][m] = 0.f;

    int g = 0, j = 0, buf = 0;
    for (int step = 0; step < total_steps; ++step) {
        int gn = g, jn = j + 1;
        if (jn == NJ) { jn = 0; gn = g + 1; }
        if (step + 1 < total_steps) load_step(gn, jn, wn);
        const int i4 = (j * 2) * kGemvThreads + tid;
#pragma unroll
        for (int m = 0; m < M; ++m) {
            const float4 xl = xs4[m * m_stride4 + i4];
            const float4 xh = xs4[m * m_stride4 + i4 + kGemvThreads];
#pragma unroll
            for (int r = 0; r < kR; ++r) {
                const uint4 ww = wc[r];
                const float2 w0 = unpack2(ww.x), w1 = unpack2(ww.y), w2 = unpack2(ww.z), w3 = unpack2(ww.w);
                float a = acc[r][m];
                a = fmaf(w0.x, xl.x, a); a = fmaf(w0.y, xl.y, a);
                a = fmaf(w1.x, xl.z, a); a = fmaf(w1.y, xl.w, a);
                a = fmaf(w2.x, xh.x, a); a = fmaf(w2.y, xh.y, a);
                a = fmaf(w3.x, xh.z, a); a = fmaf(w3.y, xh.w, a);
                acc[r][m] = a;
            }
        }
        if (j == NJ - 1) {  // the group's rows are complete: reduce across the CTA and run the epilogue
            float* pb = part + (size_t)buf * kGemvWarps * kR * M;
#pragma unroll
            for (int r = 0; r < kR; ++r)
#pragma unroll
                for (int m = 0; m < M; ++m) {
                    const float v = warp_sum(acc[r][m]);
                    if (lane == 0) pb[warp * kR * M + r * M + m] = v;
                    acc[r][m] = 0.f;
                }
            __syncthreads();  // one barrier per group; `part` is double buffered so fast warps may run ahead
            if (PAIRED) {
                if (tid < (kR / 2) * M) {
                    const int prl = tid / M, m = tid % M;
                    const int u0 = ub + g * kR + 2 * prl;
                    if (u0 < ue) {
                        float a = 0.f, b = 0.f;
#pragma unroll
                        for (int w = 0; w < kGemvWarps; ++w) {
                            a += pb[w * kR * M + (2 * prl) * M + m];
                            b += pb[w * kR * M + (2 * prl + 1) * M + m];
                        }
                        if (NORM) { a *= rstd_s[m]; b *= rstd_s[m]; }
                        const int row0 = unit_row(u0);
                        const int tile = row0 >> 7, jj = row0 & 63;
                        if (EPI == EPI_SWIGLU) {
                            p.out[(size_t)m * (p.N >> 1) + tile * 64 + jj] = silu(a) * b;
                        } else {  // EPI_QKV: tile == head
                            const int head = tile;
                            if (head < p.n_heads + p.n_kv) {  // q or k head: rotate (rotate_half convention)
                                const float c = p.cos[m * kHalfDim + jj], s = p.sin[m * kHalfDim + jj];
                                const float lo = a * c - b * s, hi = b * c + a * s;
                                if (head < p.n_heads) {
                                    float* q = p.q_out + (size_t)m * p.n_heads * kHeadDim + head * kHeadDim + jj;
                                    q[0] = lo;
                                    q[kHalfDim] = hi;
                                } else {
                                    __nv_bfloat16* kd = p.k_pool + p.kv_off[m] + (int64_t)(head - p.n_heads) * p.head_stride + jj;
                                    kd[0] = __float2bfloat16_rn(lo);
                                    kd[kHalfDim] = __float2bfloat16_rn(hi);
                                }
                            } else {
                                __nv_bfloat16* vd = p.v_pool + p.kv_off[m] + (int64_t)(head - p.n_heads - p.n_kv) * p.head_stride + jj;
                                vd[0] = __float2bfloat16_rn(a);
                                vd[kHalfDim] = __float2bfloat16_rn(b);
                            }
                        }
                    }
                }
            } else {
                if (tid < kR * M) {
                    const int r = tid / M, m = tid % M;
                    const int u = ub + g * kR + r;
                    if (u < ue) {
                        float v = 0.f;
#pragma unroll
                        for (int w = 0; w < kGemvWarps; ++w) v += pb[w * kR * M + r * M + m];
                        if (NORM) v *= rstd_s[m];
                        float* o = p.out + (size_t)m * p.N + u;
                        *o = (EPI == EPI_RESID) ? (*o + v) : v;
                    }
                }
            }
            buf ^= 1;
        }
#pragma unroll
        for (int r = 0; r < kR; ++r) wc[r] = wn[r];
        g = gn;
        j = jn;
    }
}

// ---------------------------------------------------------------------------------------------------------------
// 2..8 token rows: the same CTA-wide streaming, but the multiply-accumulate runs on the tensor cores
// (mma.sync m16n8k16, fp16 x fp16 -> fp32): weights are the A operand straight from global memory, the (fp16) token
// rows are the 8-column B operand.  Measured motivation: on CUDA cores the fp16->fp32 conversions and FMAs scale with
// the row count and the SIMT kernel drops from 94 % of HBM bandwidth at 1 row to 40 % at 4 rows; here the row count is
// free up to 8.  A 16-byte global load gives a lane 8 consecutive k of ONE weight row, while an A fragment wants
// (row g, row g+8) x (k 2t..2t+1, 2t+8..2t+9): the k index is permuted consistently on both operands — lane (g, t)
// loads row g and row g+8 at physical k = k0 + 8t .. +7 and feeds two MMAs (first / second four elements), the B
// fragment is the token's 8 values at the same physical k.  A dot product does not care about the order of k.
// Work split: a block of 16 row units per CTA at a time, K split over the 16 warps, 8 loads in flight per lane.
constexpr int kMmaRows = 16;

__device__ __forceinline__ void mma_f16(float (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
                 : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t pack_h2(float a, float b) {
    a = fminf(fmaxf(a, -65504.f), 65504.f);
    b = fminf(fmaxf(b, -65504.f), 65504.f);
    __half2 t = __floats2half2_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&t);
}

template <int EPI, bool NORM>
__global__ void __launch_bounds__(kGemvThreads, 2) gemv_mma_kernel(const GemvParams p, const int M) {
    extern __shared__ __align__(16) float smem[];
    __shared__ float red[33];
    __shared__ float rstd_s[8];
    constexpr bool PAIRED = (EPI == EPI_QKV || EPI == EPI_SWIGLU);
    const int K = p.K, kchunks = K >> 3;
    const int KW = (((K + kGemvWarps - 1) / kGemvWarps) + 31) / 32 * 32;  // k range of one warp, multiple of 32
    const int steps = KW / 32;
    const int Kpad = KW * kGemvWarps + 32;  // +64 bytes per row: the 8 token rows of a B-fragment read land in different banks
    __half* xs = reinterpret_cast<__half*>(smem);                                   // [M][Kpad] fp16
    float* part = smem + ((size_t)M * Kpad / 2);                                   // [warps][16][8]
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, g = lane >> 2, t4 = lane & 3;

    const int n_blocks_total = (p.N + kMmaRows - 1) / kMmaRows;
    const int bb = (int)((int64_t)n_blocks_total * blockIdx.x / gridDim.x);
    const int be = (int)((int64_t)n_blocks_total * (blockIdx.x + 1) / gridDim.x);
    auto unit_row = [&](int u) -> int {
        if (PAIRED) { const int pr = u >> 1; return ((pr >> 6) << 7) + (pr & 63) + ((u & 1) << 6); }
        return u;
    };
    const uint4* W4 = reinterpret_cast<const uint4*>(p.W);
    const int64_t row_stride4 = K >> 3;
    const int kw0 = warp * KW;  // first k of this warp

    constexpr int U = 4;  // steps in flight: 2 x U 16-byte loads per lane
    auto load_chunk = [&](int blk, int s0, uint4 (&a0)[U], uint4 (&a1)[U]) {
        const int u0 = blk * kMmaRows + g, u1 = u0 + 8;
        const int64_t r0 = (u0 < p.N) ? (int64_t)unit_row(u0) * row_stride4 : -1, r1 = (u1 < p.N) ? (int64_t)unit_row(u1) * row_stride4 : -1;
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int c = ((kw0 + (s0 + u) * 32) >> 3) + t4;
            const bool ok = (s0 + u < steps) && (c < kchunks);
            a0[u] = (ok && r0 >= 0) ? ld_stream(W4 + r0 + c) : make_uint4(0, 0, 0, 0);
            a1[u] = (ok && r1 >= 0) ? ld_stream(W4 + r1 + c) : make_uint4(0, 0, 0, 0);
        }
    };
    uint4 a0[U], a1[U];
    if (bb < be) load_chunk(bb, 0, a0, a1);

    asm volatile("griddepcontrol.wait;" ::: "memory");
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");

    // ---- stage the token rows as fp16 (x * norm_weight when NORM; the 1/rms scalar is applied at the end)
    for (int m = 0; m < M; ++m) {
        const float* xr = p.x + (size_t)m * p.x_stride;
        __half* xm = xs + (size_t)m * Kpad;
        float ss = 0.f;
        for (int c = tid; c < (Kpad - 32) / 8; c += kGemvThreads) {
            uint4 o = make_uint4(0, 0, 0, 0);
            if (c < kchunks) {
                float4 lo = *reinterpret_cast<const float4*>(xr + c * 8), hi = *reinterpret_cast<const float4*>(xr + c * 8 + 4);
                if (NORM) {
                    ss += lo.x * lo.x + lo.y * lo.y + lo.z * lo.z + lo.w * lo.w + hi.x * hi.x + hi.y * hi.y + hi.z * hi.z + hi.w * hi.w;
                    const uint4 wb = *reinterpret_cast<const uint4*>(p.norm_w + c * 8);
                    lo.x *= bf_lo(wb.x); lo.y *= bf_hi(wb.x); lo.z *= bf_lo(wb.y); lo.w *= bf_hi(wb.y);
                    hi.x *= bf_lo(wb.z); hi.y *= bf_hi(wb.z); hi.z *= bf_lo(wb.w); hi.w *= bf_hi(wb.w);
                }
                o = make_uint4(pack_h2(lo.x, lo.y), pack_h2(lo.z, lo.w), pack_h2(hi.x, hi.y), pack_h2(hi.z, hi.w));
            }
            *reinterpret_cast<uint4*>(xm + c * 8) = o;
        }
        if (NORM) {
            ss = block_sum(ss, red);
            if (tid == 0) rstd_s[m] = rsqrtf(ss / (float)K + p.eps);
        }
    }
    __syncthreads();

    for (int blk = bb; blk < be; ++blk) {
        float acc[4] = {0.f, 0.f, 0.f, 0.f};
        for (int s0 = 0; s0 < steps; s0 += U) {
            if (!(blk == bb && s0 == 0)) load_chunk(blk, s0, a0, a1);
#pragma unroll
            for (int u = 0; u < U; ++u) {
                if (s0 + u < steps) {
                    uint4 xf = make_uint4(0, 0, 0, 0);
                    if (g < M) xf = *reinterpret_cast<const uint4*>(xs + (size_t)g * Kpad + kw0 + (s0 + u) * 32 + t4 * 8);
                    mma_f16(acc, a0[u].x, a1[u].x, a0[u].y, a1[u].y, xf.x, xf.y);
                    mma_f16(acc, a0[u].z, a1[u].z, a0[u].w, a1[u].w, xf.z, xf.w);
                }
            }
        }
        // cross-warp reduction of the 16 x 8 tile: c0,c1 = (row g, tokens 2t,2t+1), c2,c3 = (row g+8, same tokens)
        float* pw = part + warp * 128;
        pw[g * 8 + 2 * t4] = acc[0];
        pw[g * 8 + 2 * t4 + 1] = acc[1];
        pw[(g + 8) * 8 + 2 * t4] = acc[2];
        pw[(g + 8) * 8 + 2 * t4 + 1] = acc[3];
        __syncthreads();
        if (tid < 8 * M) {
            const int slot = tid / M, m = tid % M;  // paired: slot = pair index in the block; else slots 0..7 and 8..15
            auto total = [&](int r) {
                float v = 0.f;
#pragma unroll
                for (int w = 0; w < kGemvWarps; ++w) v += part[w * 128 + r * 8 + m];
                return NORM ? v * rstd_s[m] : v;
            };
            if (PAIRED) {
                const int u0 = blk * kMmaRows + 2 * slot;
                if (u0 < p.N) {
                    const float a = total(2 * slot), b = total(2 * slot + 1);
                    const int row0 = unit_row(u0);
                    const int tile = row0 >> 7, jj = row0 & 63;
                    if (EPI == EPI_SWIGLU) {
                        p.out[(size_t)m * (p.N >> 1) + tile * 64 + jj] = silu(a) * b;
                    } else {
                        const int head = tile;
                        if (head < p.n_heads + p.n_kv) {
                            const float c = p.cos[m * kHalfDim + jj], s = p.sin[m * kHalfDim + jj];
                            const float lo = a * c - b * s, hi = b * c + a * s;
                            if (head < p.n_heads) {
                                float* q = p.q_out + (size_t)m * p.n_heads * kHeadDim + head * kHeadDim + jj;
                                q[0] = lo;
                                q[kHalfDim] = hi;
                            } else {
                                __nv_bfloat16* kd = p.k_pool + p.kv_off[m] + (int64_t)(head - p.n_heads) * p.head_stride + jj;
                                kd[0] = __float2bfloat16_rn(lo);
                                kd[kHalfDim] = __float2bfloat16_rn(hi);
                            }
                        } else {
                            __nv_bfloat16* vd = p.v_pool + p.kv_off[m] + (int64_t)(head - p.n_heads - p.n_kv) * p.head_stride + jj;
                            vd[0] = __float2bfloat16_rn(a);
                            vd[kHalfDim] = __float2bfloat16_rn(b);
                        }
                    }
                }
            } else {
#pragma unroll
                for (int half = 0; half < 2; ++half) {
                    const int r = slot + 8 * half, u = blk * kMmaRows + r;
                    if (u < p.N) {
                        const float v = total(r);
                        float* o = p.out + (size_t)m * p.N + u;
                        *o = (EPI == EPI_RESID) ? (*o + v) : v;
                    }
                }
            }
        }
        __syncthreads();  // `part` is reused by the next block
    }
}

static bool pdl_enabled() {
    static int v = -1;
    if (v < 0) {
        const char* e = getenv("B200SHARD_PDL");
        v = (e != nullptr && e[0] == '0') ? 0 : 1;
    }
    return v == 1;
}

template <typename KernelT, typename... Args>
static void launch_pdl_args(KernelT kfn, dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args... args) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl_enabled() ? 1 : 0;
    cudaLaunchKernelEx(&cfg, kfn, args...);
}

size_t gemv_smem_bytes(int M, int K) {
    if (M == 1) {
        const int NJ = (K / 8 + kGemvThreads - 1) / kGemvThreads;
        return ((size_t)NJ * kGemvThreads * 8 + 2 * kGemvWarps * kR) * sizeof(float);
    }
    const int KW = (((K + kGemvWarps - 1) / kGemvWarps) + 31) / 32 * 32;
    return (size_t)M * (KW * kGemvWarps + 32) * 2 + (size_t)kGemvWarps * 128 * sizeof(float);
}

void launch_gemv(const GemvParams& p_in, int M, int epi, bool norm, int num_sms, cudaStream_t st) {
    GemvParams p = p_in;
    if (p.x_stride == 0) p.x_stride = p.K;
    const size_t smem = gemv_smem_bytes(M, p.K);
    const int max_ctas = (smem <= 110 * 1024) ? 2 * num_sms : num_sms;
    // every CTA walks its rows in groups: pick the CTA count so that all CTAs have the same number of groups
    // (e.g. 4096 rows -> 256 CTAs x 4 groups of 4 instead of 296 CTAs x 3.46 -> 4); the kernel is HBM-bound, a few idle
    // CTA slots cost nothing, an extra group on the critical path does
    const int group_rows = (M == 1) ? kR : kMmaRows;
    const int groups = (p.N + group_rows - 1) / group_rows;
    const int per_cta = (groups + max_ctas - 1) / max_ctas;
    int grid = (groups + per_cta - 1) / per_cta;
    if (grid < 1) grid = 1;
#define B200_GEMV_CASE(E, NRM)                                                                             \
    {                                                                                                      \
        if (M == 1) {                                                                                      \
            auto kfn = gemv_kernel<1, E, NRM>;                                                             \
            cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);             \
            launch_pdl_args(kfn, dim3(grid), dim3(kGemvThreads), smem, st, p);                             \
        } else {                                                                                           \
            auto kfn = gemv_mma_kernel<E, NRM>;                                                            \
            cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);             \
            launch_pdl_args(kfn, dim3(grid), dim3(kGemvThreads), smem, st, p, M);                          \
        }                                                                                                  \
    }
    if (epi == EPI_STORE) { if (norm) B200_GEMV_CASE(EPI_STORE, true) else B200_GEMV_CASE(EPI_STORE, false) }
    else if (epi == EPI_RESID) B200_GEMV_CASE(EPI_RESID, false)
    else if (epi == EPI_QKV) B200_GEMV_CASE(EPI_QKV, true)
    else B200_GEMV_CASE(EPI_SWIGLU, true)
#undef B200_GEMV_CASE
}

}  // namespace b200
