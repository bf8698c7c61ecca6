// Host-side launcher prototypes shared between api.cu and the kernel translation units.
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <stdint.h>

namespace b200 {

constexpr int kMaxBatch = 64;
constexpr int kMaxGemvRows = 8;   // token rows served by the weight-streaming kernels (1: SIMT fp32, 2..8: mma.sync fp16)
constexpr int kMaxSplits = 16;    // split-K factor bound of decode attention
constexpr int kMaxChunkKeys = 8192;  // keys one decode-attention CTA may hold scores for
struct SlotList { int32_t v[kMaxBatch]; };
struct InvFreq { float v[64]; };  // RoPE inverse frequencies, head_dim/2 entries

// ---- KV bookkeeping living in device memory (graph-replayable: nothing host-baked) -----
struct KvDev {
    int32_t* seq_lens;     // [max_slots] committed length per slot
    int32_t* block_table;  // [max_slots][max_pages_per_slot] page ids (-1 = none)
    int32_t* pages_held;   // [max_slots]
    int32_t* free_stack;   // [n_pages]
    int32_t* free_top;     // [1] number of free pages
    int32_t* error;        // [1] sticky: 1 = pool exhausted, 2 = slot overflow
    int32_t max_slots, max_pages_per_slot, n_pages, page_size;
};

// per-call token metadata (device arrays sized max_tokens / max_batch)
struct TokMeta {
    int32_t* call_base;  // [B] cached length of row b before this call
    int32_t* pos;        // [M] absolute position of token m
    int32_t* t_end;      // [M] number of keys token m attends to
    int64_t* kv_off;     // [M] element offset of (token, kv head 0) inside one layer's K (or V) pool
    float* cos;          // [M][64]
    float* sin;          // [M][64]
};

enum GemvEpi { EPI_STORE = 0, EPI_RESID = 1, EPI_QKV = 2, EPI_SWIGLU = 3 };

struct GemvParams {
    const void* W;           // [N,K] row-major, fp16
    const float* x;          // [M,K] fp32, row m at x + m * x_stride
    int64_t x_stride = 0;    // elements between token rows (0: K)
    const __nv_bfloat16* norm_w;  // [K] (NORM prologue) or nullptr
    float eps;
    int N, K;
    float* out;              // STORE: [M,N]; RESID: [M,N] += ; SWIGLU: [M,N/2]
    // QKV epilogue
    float* q_out;            // [M, n_heads*128]
    __nv_bfloat16* k_pool;   // this layer's K pool base
    __nv_bfloat16* v_pool;
    const int64_t* kv_off;
    const float* cos;
    const float* sin;
    int n_heads, n_kv;
    int64_t head_stride;     // page_size*128 elements between kv heads inside a page
};
void launch_gemv(const GemvParams& p, int M, int epi, bool norm, int num_sms, cudaStream_t st);
size_t gemv_smem_bytes(int M, int K);

void launch_kv_alloc(const KvDev& kv, const SlotList& slots, int B, int S, int32_t* call_base, cudaStream_t st);
void launch_kv_reset(const KvDev& kv, int slot, cudaStream_t st);
void launch_kv_init(const KvDev& kv, cudaStream_t st);
void launch_prep(const KvDev& kv, const SlotList& slots, int B, int S, int causal, const InvFreq& inv, int n_kv,
                 const TokMeta& tm, cudaStream_t st);

// `in` in any message dtype (B200_DT_* code) -> fp32 residual stream
void launch_load_residual(const void* in, int dtype, float* x, int64_t n, cudaStream_t st);
// kv_alloc + prep + (embedding gather | load_residual) of a single-token call for a few rows, one launch
void launch_decode_prologue(const KvDev& kv, const SlotList& slots, int B, int causal, const InvFreq& inv, int n_kv, const TokMeta& tm,
                            const void* in, int in_dtype, const int64_t* tok_in, const __nv_bfloat16* table, int V, int H, float* x,
                            cudaStream_t st);
// x fp32 [rows,n] -> out_dtype (B200_DT_* code: 0 bf16, 1 fp32, 2 fp16 saturating); norm_w == nullptr: plain convert
void launch_rmsnorm_bf16(const float* x, const __nv_bfloat16* norm_w, float eps, int rows, int n,
                         void* out, cudaStream_t st, int out_dtype = 0);
void launch_scale_f16(const float* x, const __nv_bfloat16* norm_w, int rows, int n, __half* out, float* ss_out, cudaStream_t st);
// last-row gather + optional norm, fp32 out [B,n] (input rows b*S+S-1)
void launch_lastrow_norm_f32(const float* x, const __nv_bfloat16* norm_w, float eps, int B, int S, int n,
                             float* out, cudaStream_t st);
void launch_lastrow_to_f32(const void* h, int dtype, int B, int S, int n, float* out, cudaStream_t st);
void launch_argmax(const float* logits, int B, int V, int64_t* tok, float* margin, cudaStream_t st);
void launch_argmax_range(const float* logits, int B, int n, int offset, const float* in_best, const float* in_second, const int64_t* in_idx,
                         float* out_best, float* out_second, int64_t* out_idx, float* out_margin, cudaStream_t st);
void launch_embed(const __nv_bfloat16* table, const int64_t* ids, int n, int H, int V, void* out, int out_dtype,
                  cudaStream_t st);
// src (.pth storage dtype) -> bf16, or -> fp16 (saturating) when f16_out
void launch_convert_to_bf16(const void* src, int dtype, void* dst, int64_t n, cudaStream_t st, bool f16_out = false);
// dst rows [(r/64)*128 + half*64 + r%64] <- src row r (gate: half 0, up: half 1); fp16 out
void launch_pack_gate_up(const void* src, int dtype, void* dst, int rows, int cols, int half, cudaStream_t st);

struct AttnDecodeParams {
    const float* q;            // [M, n_heads*128] fp32 (rope applied)
    const __nv_bfloat16* k_pool;
    const __nv_bfloat16* v_pool;
    const int32_t* block_table;  // [max_slots][max_pages_per_slot]
    const int32_t* t_end;      // [M]
    const int32_t* call_base;  // [B] keys cached before this call (may be prefetched early) or nullptr
    SlotList slots;            // row b -> slot
    int S;                     // tokens per row (m = b*S + s)
    int n_heads, n_kv, page_size, max_pages_per_slot, n_splits;
    float* part;               // scratch: [M][n_heads][n_splits][130] partials, then M*n_heads arrival tickets (zeroed)
    float* out;                // [M, n_heads*128] fp32
    __half* out_h = nullptr;   // optional fp16 copy of the output (operand of the tensor-core o_proj for 9..32 rows)
    unsigned int* tickets;     // [M*n_heads] arrival counters, zero before the launch (nullptr: right behind the partials)
};
void launch_attn_decode(const AttnDecodeParams& p, int M, cudaStream_t st);
size_t attn_decode_scratch_floats(int rows, int n_heads, int n_splits);
// tensor-core / TMA variant for many rows (attn_decode_tc.cu): one CTA per (row, kv head), no scratch.  max_keys = upper bound of
// t_end over the rows for as long as the launch may be replayed.  0 = launched, 1 = shape not served (use launch_attn_decode), -1 = error
int launch_attn_decode_tc(const AttnDecodeParams& p, int M, int max_keys, int64_t pool_rows, cudaStream_t st);
int attn_decode_cluster_splits();  // split count served by the cluster variant (0: disabled by B200SHARD_ATTN_CLUSTER=0)

struct AttnPrefillParams {
    const __nv_bfloat16* q;    // [B*S, n_heads*128] bf16 (rope applied): high part
    const __nv_bfloat16* q_lo; // bf16(q - q_hi) or nullptr: second MMA pass over K so q keeps ~16 mantissa bits
    int out_f16;               // output element type: 0 bf16, 1 fp16
    const __nv_bfloat16* k_pool;
    const __nv_bfloat16* v_pool;
    const int32_t* block_table;
    const int32_t* call_base;  // [B]
    SlotList slots;
    int B, S, causal;
    int n_heads, n_kv, page_size, max_pages_per_slot;
    void* out;                 // [B*S, n_heads*128] 16-bit
};
void launch_attn_prefill(const AttnPrefillParams& p, cudaStream_t st);
// tcgen05 / TMA version (needs q as a hi/lo pair); pool_rows = n_pages * n_kv * page_size rows of 128 in one layer's pool
int launch_attn_prefill_tc(const AttnPrefillParams& p, int64_t pool_rows, cudaStream_t st);
int tma_map_2d(void* map_out, const void* ptr, int rows, int cols, int box_rows);

// tcgen05 GEMM: C[M,N] = A[M,K] * W[N,K]^T with fused epilogues
// Weight-streaming tensor-core GEMM for 9..32 token rows (gemm_stream_tc.cu): operands fp16, epilogues = GemvEpi
constexpr int kMaxStreamRows = 32;
constexpr int kGemvRowsDefault = 3;  // measured crossover: 4 rows and more run faster on the tcgen05 streaming GEMM (105 vs 111 us per 7B layer)
struct StreamGemmParams {
    const void* W;            // [N,K] fp16 (paired layouts as for the GEMV)
    const void* X;            // [M,K] fp16
    int M, N, K;
    int split = 0;            // cluster size of the split-K (1, 2, 4, 8); 0: chosen by the launcher
    float* out = nullptr;     // EPI_STORE / EPI_RESID: fp32 [M,N]
    __half* act_h = nullptr;  // EPI_SWIGLU: fp16 [M,N/2]
    float* q_out = nullptr;   // EPI_QKV: fp32 [M, n_heads*128]
    __nv_bfloat16* k_pool = nullptr;
    __nv_bfloat16* v_pool = nullptr;
    const int64_t* kv_off = nullptr;
    const float* cos = nullptr;
    const float* sin = nullptr;
    int n_heads = 0, n_kv = 0;
    int64_t head_stride = 0;
    // RMSNorm folded into the GEMMs (batched-decode path).  Producer side, EPI_RESID: besides x += ..., also write
    // xw_out[m][n] = fp16(x_new * next_norm_w[n]), the un-normalised operand of the NEXT projection.  Consumer side, EPI_QKV /
    // EPI_SWIGLU with x_res != nullptr: every CTA recomputes rstd[m] = rsqrt(mean(x_res[m]^2) + eps) from the fp32 residual
    // while its MMAs run, and scales the finished dot products by it (LlamaRMSNorm, modeling_llama.py:62-67).
    const __nv_bfloat16* next_norm_w = nullptr;
    __half* xw_out = nullptr;
    const float* x_res = nullptr;   // [M,K] fp32
    float eps = 0.f;
    // Row statistics handed from producer to consumer instead of being recomputed by every consuming CTA: the producer's CTA
    // number c (tile * split + rank, `gemm_stream_parts` of them) writes row_ss_out[m * parts + c] = sum of x_new[m][n]^2 over
    // the columns n it owns, in a fixed order; the consumer adds the `row_ss_parts` partials of a row in a fixed order too
    // (bit-reproducible, no atomics) and needs M * parts floats instead of the M x K residual block.
    float* row_ss_out = nullptr;
    const float* row_ss = nullptr;
    int row_ss_parts = 0;
    // stream-K kernel (gemm_streamk_tc.cu): partial-sum slots [grid][MP][128] floats and one flag per CTA (zero between launches)
    float* ws = nullptr;
    unsigned int* ws_flags = nullptr;
};
int launch_gemm_streamk(const StreamGemmParams& p, int epi, int num_sms, cudaStream_t st);
const char* gemm_streamk_last_error();
size_t gemm_streamk_ws_floats(int num_sms);
int launch_gemm_stream(const StreamGemmParams& p, int epi, int num_sms, cudaStream_t st);
int gemm_stream_split(int N, int K, int num_sms);
int gemm_stream_parts(int N, int K, int num_sms);   // CTAs (= row_ss partials per token row) of the default launch of an N x K GEMM
bool gemm_streamk_active();                         // B200SHARD_STREAMK=1: that kernel neither writes nor reads row_ss
const char* gemm_stream_last_error();

// Persistent single-token decode kernel (decode_mega.cu): every layer of the shard in one cooperative launch.
struct MegaLayer { const __half *wqkv, *wo, *wgu, *wdown; const __nv_bfloat16 *norm1, *norm2; };
struct MegaParams {
    const MegaLayer* layers;   // [n_layers] in device memory
    int n_layers, H, I, n_heads, n_kv, qkv_rows;
    float eps;
    float* x;                  // [H] fp32 residual stream (in / out)
    float* qkv_raw;            // [qkv_rows] q | k | v before RoPE
    float* attn_out;           // [n_heads * 128]
    float* gu_raw;             // [2 I] gate / up in the packed row order of wgu
    __nv_bfloat16* k_pool;     // layer 0's pools; layer l at + l * layer_stride
    __nv_bfloat16* v_pool;
    int64_t layer_stride, head_stride;
    const int32_t* block_table;
    int slot, max_pages_per_slot, page_size;
    const int32_t* t_end;      // TokMeta of the call (row 0)
    const int64_t* kv_off;
    const float* cos;
    const float* sin;
    float* part;               // [n_heads][n_splits][130] split-K partials
    unsigned int* tickets;     // [n_heads], zero between launches
    unsigned int* bar;         // [8] words 6-7: the grid barrier's 64-bit arrival counter; 2..5: failure record of a wait that ran out
    int n_splits = 0, n_slots = 0, xs_floats = 0, profile = 0;   // filled in by the launcher
};
int launch_decode_mega(const MegaParams& p, int num_sms, cudaStream_t st);   // 0 launched, 1 shape not served, -1 error
const char* decode_mega_last_error();

enum GemmEpi { GEMM_STORE_BF16 = 0, GEMM_RESID_F32 = 1, GEMM_SWIGLU_BF16 = 2, GEMM_STORE_F32 = 3, GEMM_QKV_ROPE = 4 };
struct GemmParams {
    const void* A;           // [M,K] bf16, or fp16 when ab_f16
    const void* W;           // [N,K] same element type as A
    int M, N, K;
    void* out;               // STORE_BF16: bf16 [M,N]; RESID_F32: fp32 [M,N] +=; SWIGLU: 16-bit [M,N/2]; STORE_F32
    int epi;
    int ab_f16 = 0;          // operand element type of A and W: 0 bf16, 1 fp16
    int out_f16 = 0;         // SWIGLU output element type
    // QKV_ROPE epilogue
    __nv_bfloat16* q_hi = nullptr;  // [M, n_heads*128] bf16(q)
    __nv_bfloat16* q_lo = nullptr;  // [M, n_heads*128] bf16(q - q_hi)
    __nv_bfloat16* k_pool = nullptr;
    __nv_bfloat16* v_pool = nullptr;
    const int64_t* kv_off = nullptr;
    const float* cos = nullptr;
    const float* sin = nullptr;
    int n_heads = 0, n_kv = 0;
    int64_t head_stride = 0;
};
// returns cudaError-like int (0 ok); message via get_gemm_error()
int launch_gemm(const GemmParams& p, cudaStream_t st);
const char* gemm_error();

// records the message b200_last_error() returns for the calling thread (defined in api.cu)
void set_last_error(const char* msg);

// hop
constexpr int kHopHeaderBytes = 256;
struct HopHeader { unsigned char b[kHopHeaderBytes]; };
void launch_hop_send_msg(const HopHeader& hdr, const void* payload, int64_t bytes, void* peer_slot, uint64_t* peer_flag,
                         uint64_t seq, unsigned int* counter, cudaStream_t st);
// payload == nullptr: the payload was already written in place; only publish the flag
void launch_hop_send(const void* payload, int64_t bytes, void* peer_slot, uint64_t* peer_flag, uint64_t seq,
                     unsigned int* counter, cudaStream_t st);
void launch_hop_recv(void* out, int64_t bytes, const void* local_slot, const uint64_t* local_flag, uint64_t seq,
                     int64_t max_wait_cycles, int* err, cudaStream_t st);

}  // namespace b200
