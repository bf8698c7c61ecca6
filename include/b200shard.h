/*
 * b200shard.h — C ABI of libb200shard.so: the shard-execution hot path of
 * seanbonjean/llm-sharding (NodeWorker.pass_through_shard) as hand-written sm_100a CUDA.
 *
 * The reference has no FFI on this path: the boundary it sits behind is a Python class,
 * LlamaShardPart (utils/shard_loader.py:8-78), driven by NodeWorker
 * (utils/node_worker.py:127-185 load_shards, :226-272 pass_through_shard, :319-355
 * clear_KV_cache) and by Communicator (utils/node_worker.py:13-67) for the hop.  Each entry
 * point below names the reference interface it replaces.  INTEGRATION.md shows the ctypes
 * binding a maintainer of the reference would add.
 *
 * Conventions: every function returns 0 on success or a negative B200_E* code (no C++
 * exception crosses the ABI); b200_last_error() gives the message for the calling thread.
 * Pointers named d_* are DEVICE pointers owned by the caller; h_* are host pointers.
 * `stream` is a cudaStream_t passed as void*.  Activations crossing the ABI are bf16 unless
 * b200_shard_set_io_dtype chose fp16 / fp32 ("model dtype" of the reference's messages,
 * utils/node_worker.py:252,267-271); token ids are int64 like torch.long (utils/node_worker.py:208,264).
 */
#ifndef B200SHARD_H
#define B200SHARD_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define B200_API __attribute__((visibility("default")))
#else
#define B200_API
#endif

#define B200_OK 0
#define B200_EINVAL (-1)   /* -> ValueError  ("[ERROR] ..." texts of the reference)      */
#define B200_ESTATE (-2)   /* -> RuntimeError (protocol / placement violations)           */
#define B200_ECUDA (-3)    /* -> RuntimeError (CUDA runtime / driver failure)             */
#define B200_ENOMEM (-4)   /* -> torch.cuda.OutOfMemoryError analogue (KV pages, HBM)     */
#define B200_ETIMEOUT (-5) /* hop wait exceeded its bound                                 */

/* element type of weight tensors handed to b200_shard_set_* (the .pth storage dtype,
 * utils/model_sharder.py:24,72-88); the library converts to bf16 while repacking */
#define B200_DT_BF16 0
#define B200_DT_F32 1
#define B200_DT_F16 2

typedef struct b200_shard b200_shard_t; /* replaces LlamaShardPart (+ NodeWorker's lm_head / embed_tokens) */
typedef struct b200_kv b200_kv_t;       /* replaces transformers DynamicCache held in NodeWorker.past_key_value */
typedef struct b200_hop b200_hop_t;     /* replaces Communicator's data sockets on one NVSwitch box */

typedef struct {
    int32_t hidden_size;       /* LlamaConfig fields read at utils/shard_loader.py:35, utils/node_worker.py:88 */
    int32_t intermediate_size;
    int32_t n_heads;
    int32_t n_kv_heads;
    int32_t head_dim;          /* must be 128 */
    int32_t vocab_size;
    float rms_eps;
    float rope_theta;
    int32_t n_layers;          /* end - start of this shard (utils/node_worker.py:127-137) */
    int32_t has_final_norm;    /* end == layer_num (utils/node_worker.py:155-157, utils/shard_loader.py:49-55) */
    int32_t has_lm_head;       /* end == layer_num (utils/node_worker.py:158-164) */
    int32_t has_embedding;     /* can_receive_user_request (utils/node_worker.py:105-107,119-124) */
    int32_t max_tokens;        /* largest B*S of one forward call; sizes the workspace */
    int32_t max_batch;         /* largest B of one forward call */
} b200_shard_config_t;

/* ---- shard lifecycle: LlamaShardPart.__init__ / NodeWorker.load_shards ---------------- */
B200_API int b200_shard_create(const b200_shard_config_t* cfg, int device, b200_shard_t** out);
/* One block_{i}.pth (utils/shard_loader.py:42-45): the 9 tensors in state_dict key order
 * q,k,v,o,gate,up,down [out,in] row-major, input_layernorm, post_attention_layernorm.
 * The library repacks into its own fused-QKV / paired gate-up layout and owns that copy;
 * the caller's tensors may be freed when the call returns (it synchronises `stream`). */
B200_API int b200_shard_set_layer(b200_shard_t* s, int rel_layer, const void* d_q, const void* d_k, const void* d_v,
                         const void* d_o, const void* d_gate, const void* d_up, const void* d_down,
                         const void* d_input_norm, const void* d_post_norm, int dtype, void* stream);
/* final_norm.pth / lm_head.pth / embedding.pth (utils/shard_loader.py:49-55, utils/node_worker.py:160-163,121-124) */
B200_API int b200_shard_set_final_norm(b200_shard_t* s, const void* d_weight, int dtype, void* stream);
B200_API int b200_shard_set_lm_head(b200_shard_t* s, const void* d_weight, int dtype, void* stream);
B200_API int b200_shard_set_embedding(b200_shard_t* s, const void* d_weight, int dtype, void* stream);
/* Replace the RoPE inverse-frequency table (head_dim/2 = 64 floats, host memory).  The default is
 * 1 / rope_theta^(2j/128) (compute_default_rope_parameters, modeling_llama.py:93-120); scaled variants such as the
 * Llama-3.x "llama3" rope_type (transformers modeling_rope_utils.py:550-626) are computed by the host and set here. */
B200_API int b200_shard_set_rope_inv_freq(b200_shard_t* s, const float* h_inv_freq, int n);
/* Element type of the hidden states that cross the ABI (d_hidden_in / d_hidden_out of the forward calls, b200_shard_embed's
 * output, b200_shard_lm_head_argmax's input): B200_DT_BF16 (default), B200_DT_F16 or B200_DT_F32 — the dtype the node was
 * started with (start_node.py:16-17 uses float16; utils/node_worker.py:252 casts incoming messages to it).  Only the
 * boundary changes: inside, the residual stream stays fp32 and the operands stay as described in DESIGN.md. */
B200_API int b200_shard_set_io_dtype(b200_shard_t* s, int dtype);
/* `del self.shard` + empty_cache (utils/node_worker.py:139-147) */
B200_API int b200_shard_destroy(b200_shard_t* s);

/* ---- KV lifecycle: DynamicCache() / clear_KV_cache / get_seq_length -------------------- */
/* One slot = one sequence (row of a batch).  Pages of `page_size` tokens (multiple of 64)
 * are handed to slots on demand from a pool of `n_pages` shared by all layers of the shard. */
B200_API int b200_kv_create(const b200_shard_t* s, int max_slots, int page_size, int n_pages, b200_kv_t** out);
B200_API int b200_kv_reset(b200_kv_t* kv, int slot /* -1: every slot */, void* stream); /* clear_KV_cache, utils/node_worker.py:319-355 */
B200_API int b200_kv_seq_len(const b200_kv_t* kv, int slot); /* past_key_value.get_seq_length(), utils/forwarding_utils.py:15-16 */
/* Sticky device-side error of the page allocator: 0 = none, 1 = page pool exhausted, 2 = slot capacity exceeded.  The host
 * checks capacity before every call, so this fires only when device state ran ahead of the host mirror (a captured decode
 * graph replayed more often than the pool allows).  Copies one int from the device (synchronises).  Stands where
 * torch.cuda.OutOfMemoryError would surface inside DynamicCache.update (transformers cache_utils.py:102-121). */
B200_API int b200_kv_error(b200_kv_t* kv);
/* Also destroys every CUDA graph the shard captured against this cache (their nodes point into its pools). */
B200_API int b200_kv_destroy(b200_kv_t* kv);

/* ---- the hot path: LlamaShardPart.forward (utils/shard_loader.py:57-78) ---------------- */
/* d_hidden_in / d_hidden_out: bf16 [B,S,H].  h_slots[b] is the KV slot of row b; every row
 * must have the same cached length (the reference has no per-row lengths).
 * Positions are kv_seq_len(slot) .. +S-1 (build_position_ids, utils/forwarding_utils.py:24),
 * cos/sin are recomputed on device from them (LlamaRotaryEmbedding, called at
 * utils/node_worker.py:241) so they need not travel with the message.
 * causal = 0 reproduces the reference (no mask is passed: utils/node_worker.py:253-258);
 * causal = 1 is real-Llama masking.  K,V of the S new tokens are appended to the slots.
 * The final RMSNorm is applied when the shard has one (utils/shard_loader.py:76-77). */
B200_API int b200_shard_forward(b200_shard_t* s, b200_kv_t* kv, const int32_t* h_slots, const void* d_hidden_in,
                       int B, int S, int causal, void* d_hidden_out, void* stream);
/* Tail: lm_head + argmax of the last position (utils/node_worker.py:260-265), computed for
 * row S-1 only.  d_hidden is the forward output (after the final norm).  d_token_out int64 [B];
 * d_margin_out (optional, may be NULL) float [B] = top-1 minus top-2 logit. */
B200_API int b200_shard_lm_head_argmax(b200_shard_t* s, const void* d_hidden, int B, int S, int64_t* d_token_out,
                              float* d_margin_out, void* stream);
/* Fused forward + tail for tail shards: same as the two calls above but the fp32 residual
 * never round-trips through bf16.  d_hidden_out may be NULL. */
B200_API int b200_shard_forward_argmax(b200_shard_t* s, b200_kv_t* kv, const int32_t* h_slots, const void* d_hidden_in,
                              int B, int S, int causal, void* d_hidden_out, int64_t* d_token_out,
                              float* d_margin_out, void* stream);
/* Entry node that is also the head (and possibly the tail) of the chain: embed_tokens (utils/node_worker.py:215,302)
 * + forward (+ lm_head/argmax when the shard holds it) with token ids in and token ids (tail) or hidden states out, so
 * that one greedy step of a single-GPU model — NodeWorker.receive_next_token's re-embed (utils/node_worker.py:302)
 * followed by pass_through_shard (:226-272) — is one call and, for S == 1, one CUDA-graph replay.
 * d_tok_in int64 [B*S]; d_hidden_out may be NULL on a tail shard; d_tok_out / d_margin_out as above. */
B200_API int b200_shard_step_tokens(b200_shard_t* s, b200_kv_t* kv, const int32_t* h_slots, const int64_t* d_tok_in,
                                    int B, int S, int causal, void* d_hidden_out, int64_t* d_tok_out,
                                    float* d_margin_out, void* stream);
/* Vocabulary-split tail for a layer-sharded ring of single-token rows.  The reference puts final norm + lm_head + argmax on
 * the last worker (utils/node_worker.py:155-164,260-264), which makes that stage the slowest of the ring by one 262 MB
 * matrix per token.  Here the tail may score only lm_head rows [0, split) and hand the rest to the stage that follows it
 * in the ring (the head, which receives the token anyway): no counterpart in the reference, same logits bit for bit, same
 * arg-max (first index of the maximum).
 *   b200_shard_forward_partial   (tail)  forward over the shard's layers, then d_msg_out receives
 *        float x[B][H]        the last fp32 residual row of every batch row (before the final norm)
 *        float best[B], second[B]; int64 idx[B]   best / second-best score and best row over lm_head rows [0, split)
 *        = B200_PARTIAL_MSG_BYTES(B, H) bytes.  Up to three rows per call both halves run the fused-norm GEMV of the unsplit
 *        tail (bit-identical tokens and margins); more rows: both halves normalise into fp16 and use the streaming
 *        tensor-core GEMM (scores comparable across the halves to the last bit, equal to the unsplit call's within fp16 rounding).
 *   b200_shard_set_vocab_tail    (next stage)  rows [split, V) of lm_head ([V - split, H], row-major) and the final norm
 *        weight (NULL: the model has none), device pointers of element type `dtype`; the split of a shard is fixed.
 *   b200_shard_token_from_partial (next stage)  scores its rows against x (same fused final norm), merges with the
 *        partial result and writes the token ids (int64 [B]) and, if asked, the top-1 margins. */
#define B200_PARTIAL_MSG_BYTES(B, H) ((size_t)(B) * ((size_t)(H) * 4 + 16))
B200_API int b200_shard_forward_partial(b200_shard_t* s, b200_kv_t* kv, const int32_t* h_slots, const void* d_hidden_in, int B, int S,
                                        int causal, int split, void* d_msg_out, void* stream);
B200_API int b200_shard_set_vocab_tail(b200_shard_t* s, const void* d_lm_head_rows, const void* d_final_norm, int split, int dtype,
                                       void* stream);
B200_API int b200_shard_token_from_partial(b200_shard_t* s, const void* d_msg_in, int B, int64_t* d_tok_out, float* d_margin_out,
                                           void* stream);
/* Entry node: embed_tokens(ids) (utils/node_worker.py:215,302). d_ids int64 [n] -> bf16 [n,H] */
B200_API int b200_shard_embed(b200_shard_t* s, const int64_t* d_ids, int n, void* d_hidden_out, void* stream);

/* ---- the hop: Communicator.transfer_data / receive_data (utils/node_worker.py:44-67) --- */
/* One receive mailbox per process, in device memory, exported over CUDA IPC.  A sender maps
 * the peer's mailbox, stores the payload into it with 128-bit NVLink stores from a kernel and
 * then releases a sequence flag; the receiver's kernel acquires the flag and copies out. */
#define B200_HOP_HANDLE_BYTES 80
B200_API int b200_hop_create(int device, int n_ring_slots, int64_t slot_bytes, b200_hop_t** out);
B200_API int b200_hop_export(b200_hop_t* h, void* h_handle_out /* B200_HOP_HANDLE_BYTES */);
B200_API int b200_hop_connect(b200_hop_t* h, const void* h_peer_handle, int peer_device, int same_process);
/* d_payload -> peer mailbox ring slot (seq % the PEER's n_ring_slots — the ring geometry travels inside the exported
 * handle); `seq` must increase by 1 per message.  The ring has no back-pressure: at most n_ring_slots messages may be
 * un-consumed at any time (the chain runtime guarantees it by giving every request stream its own slot; a token only
 * re-enters a slot after the previous message of that stream came back around the ring). */
B200_API int b200_hop_send(b200_hop_t* h, const void* d_payload, int64_t bytes, uint64_t seq, void* stream);
/* wait (on the GPU) for message `seq` in the local mailbox, copy it to d_out; max_wait_us bounds the spin */
B200_API int b200_hop_recv(b200_hop_t* h, void* d_out, int64_t bytes, uint64_t seq, int64_t max_wait_us, void* stream);
/* Zero-copy variant: the producer's last kernel writes its output straight into b200_hop_peer_slot(h, seq) (a mapped
 * peer pointer: the stores cross NVLink) and then calls b200_hop_send with d_payload == NULL (publish only); the
 * consumer calls b200_hop_recv with d_out == NULL (wait only) and lets its first kernel read b200_hop_local_slot. */
B200_API void* b200_hop_peer_slot(b200_hop_t* h, uint64_t seq);
B200_API void* b200_hop_local_slot(b200_hop_t* h, uint64_t seq);
/* Whole messages for a host-driven loop (NodeController.run_worker_loop, utils/node_worker.py:493-559, which receives,
 * classifies and dispatches on the host): a header of at most B200_HOP_HEADER_BYTES describing the message (kind, shapes,
 * dtype, pos_offset, slots: whatever the caller's wire format needs, opaque to the library) is stored at the start of
 * the peer's ring slot by the same kernel that copies the payload behind it and releases the flag, so the tensor never
 * leaves device memory and nothing is pickled (the reference's transfer_data does torch.save -> file -> socket,
 * utils/node_worker.py:44-54).  slot_bytes of the receiver must hold header + payload. */
#define B200_HOP_HEADER_BYTES 256
B200_API int b200_hop_send_msg(b200_hop_t* h, const void* h_header, int header_bytes, const void* d_payload, int64_t bytes,
                               uint64_t seq, void* stream);
/* Non-blocking receive_data(no_block=True) (utils/node_worker.py:56-67): 1 and the header in h_header_out
 * (B200_HOP_HEADER_BYTES) when message `seq` has arrived, 0 when it has not (zmq.Again in the reference), < 0 on error.
 * The payload then sits at b200_hop_local_slot(h, seq) + B200_HOP_HEADER_BYTES in device memory.  Uses a private stream:
 * never waits for work queued on the caller's streams. */
B200_API int b200_hop_poll(b200_hop_t* h, uint64_t seq, void* h_header_out);
B200_API int b200_hop_error(b200_hop_t* h); /* nonzero once a recv kernel timed out */
B200_API int b200_hop_destroy(b200_hop_t* h);

/* ---- introspection -------------------------------------------------------------------- */
B200_API const char* b200_last_error(void);
B200_API int b200_version(void);
/* counters since creation: [0] kernel launches, [1] algorithmic HBM bytes (weights + KV read
 * + KV write), [2] algorithmic flops, [3] forward calls */
B200_API int b200_shard_counters(const b200_shard_t* s, uint64_t h_out[4]);
B200_API int b200_shard_reset_counters(b200_shard_t* s);
/* 1: capture single-token (S == 1) forwards into CUDA graphs keyed by (B, slots, pointers) */
B200_API int b200_shard_set_graphs(b200_shard_t* s, int enabled);

/* ---- single-kernel entry points (unit parity tests and micro-benchmarks) -------------- */
/* Launch ONE of the decode-path kernels of layer `layer` exactly as b200_shard_forward would (same template
 * instantiation, same buffers; inputs are whatever the workspace holds): kind 0 = RMSNorm+QKV+RoPE+KV-append,
 * 1 = decode attention over the slots' current context, 2 = o_proj+residual, 3 = RMSNorm+gate/up+SwiGLU,
 * 4 = down_proj+residual, 5 = final RMSNorm + lm_head (layer ignored).  `rows` = token rows (1..8).
 * Used by bench.py to time the dominant kernel on its own stream for the roofline figure. */
B200_API int b200_shard_run_kernel(b200_shard_t* s, b200_kv_t* kv, int kind, int layer, int rows, void* stream);
/* LlamaRMSNorm.forward (transformers modeling_llama.py:62-67): fp32 in -> bf16 out, [rows, n] */
B200_API int b200_rmsnorm(const float* d_x, const void* d_weight_bf16, float eps, int rows, int n, void* d_out_bf16,
                 void* stream);
/* Split-K decode attention over a paged K/V pool (eager_attention_forward, transformers modeling_llama.py:199-221, for
 * single-token queries): row r attends to keys [0, d_t_end[r]) found through d_block_table[r][*] (page ids into the
 * pools, layout [page][kv_head][page_size][128] bf16).  d_q / d_out fp32 [rows, n_heads*128].  d_scratch: caller-owned
 * global memory for the per-split partials and
 * arrival tickets, rows*n_heads*(n_splits*130 + 1) floats; its contents need no initialisation. */
B200_API int b200_attn_decode(const float* d_q, const void* d_k_pool, const void* d_v_pool, const int32_t* d_block_table,
                              const int32_t* d_t_end, int rows, int n_heads, int n_kv_heads, int page_size,
                              int max_pages_per_row, int n_splits, float* d_scratch, float* d_out, void* stream);
/* The same attention for MANY rows at once (batched decode: one new token of each of a few dozen requests) on the tensor
 * cores: K/V pages go HBM -> shared memory by TMA and are consumed in place by tcgen05.mma (one CTA per (row, K/V head), all
 * query heads of the group together); no scratch memory.  max_keys >= every d_t_end[r]; n_pool_pages = pages in the pools.
 * Up to 8 query heads per K/V head.  Values stored past a row's length inside its last page must be finite. */
B200_API int b200_attn_decode_tc(const float* d_q, const void* d_k_pool, const void* d_v_pool, const int32_t* d_block_table,
                                 const int32_t* d_t_end, int rows, int n_heads, int n_kv_heads, int page_size,
                                 int max_pages_per_row, int n_pool_pages, int max_keys, float* d_out, void* stream);
/* No counterpart in the reference.  b200_attn_decode without any device-memory allocation, for callers that cannot
 * allocate: a process whose HBM is already fully owned by another pool (serving runtime, torch caching allocator), a
 * per-token loop that must not hit the device-wide synchronisation of cudaMalloc / cudaFree, or a tool working on K/V
 * pages it owns without a shard handle.  The split-K partials go to an array built into the library (about 2 MiB per
 * GPU, held until the process exits) instead of d_scratch; inputs and outputs are still the caller's.
 * Limit: rows * n_heads * n_splits <= 4096, otherwise B200_EINVAL.  The first call on a device looks up the array's
 * address — make it before starting any stream capture on that device.  All calls share the array: calls that can
 * overlap on the GPU (different streams) race; keep them on one stream or order them with events. */
B200_API int b200_attn_decode_static(const float* d_q, const void* d_k_pool, const void* d_v_pool,
                                     const int32_t* d_block_table, const int32_t* d_t_end, int rows, int n_heads,
                                     int n_kv_heads, int page_size, int max_pages_per_row, int n_splits, float* d_out,
                                     void* stream);
/* Prefill attention alone (eager_attention_forward, transformers modeling_llama.py:199-221, for S query tokens per row on
 * top of d_base_len[b] cached ones): the tcgen05 / TMA kernel b200_shard_forward launches for more than 32 token rows.
 * Queries arrive as the (hi, lo) bf16 pair the QKV epilogue produces (hi = bf16(q), lo = bf16(q - hi)), [B*S, n_heads*128];
 * K/V of all base_len[b] + S positions must already be in the pools (layout [page][kv_head][page_size][128] bf16,
 * n_pool_pages pages), reached through d_block_table[b][*].  causal = 0: every query sees all base_len + S keys (what the
 * reference does, utils/node_worker.py:253-258); causal = 1: query s sees keys 0 .. base_len + s.  d_out [B*S, n_heads*128]
 * bf16 (out_f16 = 0) or fp16 (1). */
B200_API int b200_attn_prefill(const void* d_q_hi, const void* d_q_lo, const void* d_k_pool, const void* d_v_pool,
                               const int32_t* d_block_table, const int32_t* d_base_len, int B, int S, int n_heads,
                               int n_kv_heads, int page_size, int max_pages_per_row, int n_pool_pages, int causal,
                               int out_f16, void* d_out, void* stream);
/* out[M,N] (fp32) = A[M,K] (bf16) * W[N,K]^T (bf16): the tcgen05/TMA GEMM used for prefill */
B200_API int b200_gemm_bf16(const void* d_a, const void* d_w, int M, int N, int K, float* d_out, void* stream);
/* same kernel with fp16 operands (A[M,K] and W[N,K] both fp16): what the shard path runs — it holds the bf16 model
 * weights as fp16 (exact in fp16's normal range) so that activations carry 10 mantissa bits into the tensor cores */
B200_API int b200_gemm_f16(const void* d_a, const void* d_w, int M, int N, int K, float* d_out, void* stream);
/* out[M,N] (fp32) = X[M,K] (fp32) * W[N,K]^T (fp16), M <= 8: the weight-streaming decode kernels (M == 1: CUDA cores,
 * fp32 activations; 2 <= M <= 8: mma.sync tensor cores, activations rounded to fp16) */
B200_API int b200_gemv_f16(const float* d_x, const void* d_w, int M, int N, int K, float* d_out, void* stream);
/* out[M,N] (fp32) = X[M,K] (fp16) * W[N,K]^T (fp16), M <= 32: the weight-streaming tensor-core GEMM of batched decode
 * (tcgen05, weight tile as the 128-row operand, split-K over a cluster of `split` = 1, 2, 4 or 8 CTAs reduced through
 * distributed shared memory; split = 0 lets the library choose).  Needs K % 8 == 0 and K >= 64.  Stands where the
 * reference's nn.Linear calls (transformers modeling_llama.py:182-184,262-264,287-288) run for a 9..32-row batch. */
B200_API int b200_gemm_stream_f16(const void* d_x, const void* d_w, int M, int N, int K, int split, float* d_out, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200SHARD_H */
