// C ABI of libb200shard.so (include/b200shard.h): shard / KV lifecycle and the orchestration of
// one LlamaShardPart.forward (utils/shard_loader.py:57-78) as a fixed sequence of sm_100a kernels.
#include <cuda_runtime.h>
#include <math.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <map>
#include <string>
#include <vector>

#include "../../include/b200shard.h"
#include "internal.h"

using namespace b200;

static thread_local std::string g_err;
static int fail(int code, const char* fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    g_err = buf;
    return code;
}
namespace b200 {
void set_last_error(const char* msg) { g_err = msg ? msg : ""; }
}  // namespace b200
#define CU_TRY(expr)                                                                                      \
    do {                                                                                                  \
        cudaError_t e_ = (expr);                                                                          \
        if (e_ != cudaSuccess)                                                                            \
            return fail(e_ == cudaErrorMemoryAllocation ? B200_ENOMEM : B200_ECUDA, "[ERROR] CUDA: %s at %s:%d (%s)", \
                        cudaGetErrorString(e_), __FILE__, __LINE__, #expr);                               \
    } while (0)

struct LayerW {
    __nv_bfloat16 *wqkv = nullptr, *wo = nullptr, *wgu = nullptr, *wdown = nullptr, *norm1 = nullptr, *norm2 = nullptr;
    void* slab = nullptr;
    bool set = false;
};

// A captured graph has the KV pools, block table and length arrays of one b200_kv baked in.  The key therefore names
// the cache by a process-wide generation number handed out by b200_kv_create — never by its host address, which the heap
// may hand to the next b200_kv — and b200_kv_destroy drops every graph of that generation.
struct GraphKey {
    int B, causal, with_argmax, pad_;
    const void *in, *out, *tok, *margin, *tok_in;
    uint64_t kv_gen;
    int32_t slots[kMaxBatch];
    bool operator<(const GraphKey& o) const { return memcmp(this, &o, sizeof(GraphKey)) < 0; }
};
struct GraphEntry {
    cudaGraphExec_t exec = nullptr;
    uint64_t last_use = 0;
    uint64_t launches = 0;   // kernels inside (counted while capturing)
};
constexpr size_t kMaxCachedGraphs = 64;  // least-recently-used graphs beyond this are destroyed (callers that pass fresh pointers every step)

constexpr int kRowSsParts = 512;   // most CTAs a residual-producing streaming GEMM launches (per-CTA partial sums of squares per token row)
struct b200_shard {
    b200_shard_config_t cfg;
    int device = 0, num_sms = 148;
    int nq = 0, nkvd = 0, qkv_rows = 0;
    std::vector<LayerW> layers;
    __nv_bfloat16 *final_norm = nullptr, *lm_head = nullptr, *embedding = nullptr;
    // workspace
    float* x = nullptr;                 // fp32 residual stream [max_tokens, H]
    __nv_bfloat16 *xn = nullptr, *qrot = nullptr, *qrot_lo = nullptr, *attn_b = nullptr, *act_b = nullptr;  // 16-bit staging (xn / attn_b / act_b hold fp16)
    float *q_f = nullptr, *attn_f = nullptr, *act_f = nullptr;  // GEMV path (<= 4 rows), fp32
    float *lastrow = nullptr, *logits = nullptr, *attn_part = nullptr;
    float* sk_ws = nullptr;             // stream-K GEMM: partial-sum slots, one per SM
    unsigned int* sk_flags = nullptr;   // and their flags (zero between launches)
    // vocabulary-split tail: rows [aux_split, V) of lm_head and the final norm, held by the stage that FOLLOWS the tail in a ring
    __nv_bfloat16 *aux_lm_head = nullptr, *aux_final_norm = nullptr;
    int aux_split = 0;
    MegaLayer* mega_layers = nullptr;   // persistent decode kernel: the layers' weight pointers, in device memory
    unsigned int* mega_bar = nullptr;   // and its grid barrier (arrivals, generation), then the kernel's failure record
    float* row_ss = nullptr;            // folded RMSNorm: two buffers of [kMaxStreamRows][kRowSsParts] partial sums of squares
    __nv_bfloat16* lastrow_b = nullptr;
    TokMeta tm{};
    InvFreq inv{};
    uint64_t n_launch = 0, n_bytes = 0, n_flops = 0, n_calls = 0;
    bool graphs = false;
    int tail_ss_parts = 0;        // > 0: that down_proj also left the rows' partial sums of squares in row_ss (buffer A)
    bool tail_xw_ready = false;   // set by enqueue_forward when the last layer wrote the lm_head operand (folded final norm)
    int io_dtype = B200_DT_BF16;  // element type of hidden states crossing the ABI (the node's dtype, utils/node_worker.py:252)
    std::map<GraphKey, GraphEntry> graph_cache;
    uint64_t graph_clock = 0;
    std::vector<void*> allocs;
};
static void drop_graphs(b200_shard* s, uint64_t kv_gen /* 0: all */) {
    for (auto it = s->graph_cache.begin(); it != s->graph_cache.end();) {
        if (kv_gen == 0 || it->first.kv_gen == kv_gen) {
            cudaGraphExecDestroy(it->second.exec);
            it = s->graph_cache.erase(it);
        } else {
            ++it;
        }
    }
}

struct b200_kv {
    b200_shard* shard = nullptr;
    uint64_t gen = 0;  // unique per b200_kv_create (graph-cache key)
    KvDev dev{};
    __nv_bfloat16 *k_pool = nullptr, *v_pool = nullptr;
    int64_t layer_stride = 0;  // elements per layer in a pool
    std::vector<int> h_len, h_pages;
    int h_free = 0;
    std::vector<void*> allocs;
};


template <typename T>
static int dev_alloc(std::vector<void*>& keep, T** out, size_t n_elems) {
    void* p = nullptr;
    size_t bytes = n_elems * sizeof(T);
    if (bytes == 0) bytes = 16;
    cudaError_t e = cudaMalloc(&p, bytes);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return fail(B200_ENOMEM, "[ERROR] CUDA out of memory allocating %zu bytes: %s", bytes, cudaGetErrorString(e));
    }
    keep.push_back(p);
    *out = reinterpret_cast<T*>(p);
    return 0;
}
#define TRY(expr) do { int r_ = (expr); if (r_ != 0) return r_; } while (0)

extern "C" B200_API const char* b200_last_error(void) { return g_err.c_str(); }
extern "C" B200_API int b200_version(void) { return 100; }

extern "C" B200_API int b200_shard_create(const b200_shard_config_t* c, int device, b200_shard_t** out) {
    if (c == nullptr || out == nullptr) return fail(B200_EINVAL, "[ERROR] null argument");
    if (c->head_dim != 128) return fail(B200_EINVAL, "[ERROR] head_dim must be 128 (got %d)", c->head_dim);
    if (c->n_layers <= 0) return fail(B200_EINVAL, "[ERROR] start or end is invalid");
    if (c->n_heads <= 0 || c->n_kv_heads <= 0 || c->n_heads % c->n_kv_heads != 0)
        return fail(B200_EINVAL, "[ERROR] n_heads must be a positive multiple of n_kv_heads");
    if (c->hidden_size % 8 != 0 || c->intermediate_size % 64 != 0 || c->vocab_size % 8 != 0)
        return fail(B200_EINVAL, "[ERROR] hidden_size %% 8, intermediate_size %% 64 and vocab_size %% 8 must be 0");
    if (c->max_batch < 1 || c->max_batch > kMaxBatch || c->max_tokens < c->max_batch)
        return fail(B200_EINVAL, "[ERROR] need 1 <= max_batch <= %d and max_tokens >= max_batch", kMaxBatch);
    CU_TRY(cudaSetDevice(device));
    b200_shard* s = new b200_shard();
    s->cfg = *c;
    s->device = device;
    cudaDeviceProp prop;
    CU_TRY(cudaGetDeviceProperties(&prop, device));
    s->num_sms = prop.multiProcessorCount;
    s->nq = c->n_heads * 128;
    s->nkvd = c->n_kv_heads * 128;
    s->qkv_rows = s->nq + 2 * s->nkvd;
    s->layers.resize(c->n_layers);
    for (int j = 0; j < 64; ++j)  // compute_default_rope_parameters, modeling_llama.py:116-118 (fp32 like torch)
        s->inv.v[j] = 1.0f / powf(c->rope_theta, (float)(2 * j) / 128.0f);
    const size_t MT = (size_t)c->max_tokens, H = c->hidden_size, I = c->intermediate_size;
    const size_t V = c->vocab_size;
    int r = 0;
    auto& A = s->allocs;
    r |= dev_alloc(A, &s->x, MT * H);
    r |= dev_alloc(A, &s->xn, MT * H);
    r |= dev_alloc(A, &s->qrot, MT * s->nq);
    r |= dev_alloc(A, &s->qrot_lo, MT * s->nq);
    r |= dev_alloc(A, &s->attn_b, MT * s->nq);
    r |= dev_alloc(A, &s->act_b, MT * I);
    r |= dev_alloc(A, &s->q_f, (size_t)kMaxStreamRows * s->nq);
    r |= dev_alloc(A, &s->attn_f, (size_t)kMaxStreamRows * s->nq);
    r |= dev_alloc(A, &s->act_f, (size_t)kMaxGemvRows * I);
    // partials for rows x splits <= 8 x 16 per head, followed by one arrival ticket per (row, head) for up to 32 rows
    const size_t part_floats = (size_t)kMaxGemvRows * c->n_heads * kMaxSplits * 130 + (size_t)kMaxStreamRows * c->n_heads;
    r |= dev_alloc(A, &s->attn_part, part_floats);
    r |= dev_alloc(A, &s->sk_ws, gemm_streamk_ws_floats(s->num_sms));
    r |= dev_alloc(A, &s->sk_flags, (size_t)2 * s->num_sms + 8);
    r |= dev_alloc(A, &s->row_ss, (size_t)2 * kMaxStreamRows * kRowSsParts);
    r |= dev_alloc(A, &s->mega_layers, (size_t)c->n_layers);
    r |= dev_alloc(A, &s->mega_bar, (size_t)8);
    r |= dev_alloc(A, &s->lastrow, (size_t)c->max_batch * H);
    r |= dev_alloc(A, &s->lastrow_b, (size_t)c->max_batch * H);
    r |= dev_alloc(A, &s->logits, (size_t)c->max_batch * V);
    r |= dev_alloc(A, &s->tm.call_base, (size_t)c->max_batch);
    r |= dev_alloc(A, &s->tm.pos, MT);
    r |= dev_alloc(A, &s->tm.t_end, MT);
    r |= dev_alloc(A, &s->tm.kv_off, MT);
    r |= dev_alloc(A, &s->tm.cos, MT * 64);
    r |= dev_alloc(A, &s->tm.sin, MT * 64);
    if (r != 0) { b200_shard_destroy(s); return B200_ENOMEM; }
    // arrival tickets of decode attention start at zero (the kernel re-zeroes them itself afterwards)
    CU_TRY(cudaMemset(s->attn_part, 0, part_floats * sizeof(float)));
    CU_TRY(cudaMemset(s->sk_flags, 0, ((size_t)2 * s->num_sms + 8) * sizeof(unsigned int)));
    CU_TRY(cudaMemset(s->mega_bar, 0, 8 * sizeof(unsigned int)));
    *out = s;
    return 0;
}

extern "C" B200_API int b200_shard_destroy(b200_shard_t* s) {
    if (s == nullptr) return 0;
    cudaSetDevice(s->device);
    cudaDeviceSynchronize();
    drop_graphs(s, 0);
    for (void* p : s->allocs) cudaFree(p);
    for (auto& l : s->layers) if (l.slab) cudaFree(l.slab);
    delete s;
    return 0;
}

static size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

extern "C" B200_API int b200_shard_set_layer(b200_shard_t* s, int rel, const void* q, const void* k, const void* v, const void* o,
                                    const void* gate, const void* up, const void* down, const void* n1, const void* n2,
                                    int dtype, void* stream) {
    if (s == nullptr) return fail(B200_EINVAL, "[ERROR] null shard");
    if (rel < 0 || rel >= s->cfg.n_layers)
        return fail(B200_EINVAL, "[ERROR] String list: shard_weights length must be equal to (end - start)");
    if (dtype < 0 || dtype > 2) return fail(B200_EINVAL, "[ERROR] unsupported weight dtype %d", dtype);
    CU_TRY(cudaSetDevice(s->device));
    cudaStream_t st = (cudaStream_t)stream;
    LayerW& L = s->layers[rel];
    const size_t H = s->cfg.hidden_size, I = s->cfg.intermediate_size;
    const size_t n_qkv = (size_t)s->qkv_rows * H, n_o = H * s->nq, n_gu = 2 * I * H, n_dn = H * I;
    if (L.slab == nullptr) {
        size_t off = 0, o_qkv = off; off = align_up(off + n_qkv * 2, 1024);
        size_t o_o = off; off = align_up(off + n_o * 2, 1024);
        size_t o_gu = off; off = align_up(off + n_gu * 2, 1024);
        size_t o_dn = off; off = align_up(off + n_dn * 2, 1024);
        size_t o_n1 = off; off = align_up(off + H * 2, 1024);
        size_t o_n2 = off; off = align_up(off + H * 2, 1024);
        cudaError_t e = cudaMalloc(&L.slab, off);
        if (e != cudaSuccess) { cudaGetLastError(); return fail(B200_ENOMEM, "[ERROR] CUDA out of memory for layer weights (%zu bytes)", off); }
        char* b = reinterpret_cast<char*>(L.slab);
        L.wqkv = reinterpret_cast<__nv_bfloat16*>(b + o_qkv);
        L.wo = reinterpret_cast<__nv_bfloat16*>(b + o_o);
        L.wgu = reinterpret_cast<__nv_bfloat16*>(b + o_gu);
        L.wdown = reinterpret_cast<__nv_bfloat16*>(b + o_dn);
        L.norm1 = reinterpret_cast<__nv_bfloat16*>(b + o_n1);
        L.norm2 = reinterpret_cast<__nv_bfloat16*>(b + o_n2);
    }
    launch_convert_to_bf16(q, dtype, L.wqkv, (int64_t)s->nq * H, st, true);
    launch_convert_to_bf16(k, dtype, L.wqkv + (size_t)s->nq * H, (int64_t)s->nkvd * H, st, true);
    launch_convert_to_bf16(v, dtype, L.wqkv + (size_t)(s->nq + s->nkvd) * H, (int64_t)s->nkvd * H, st, true);
    launch_convert_to_bf16(o, dtype, L.wo, (int64_t)n_o, st, true);
    launch_pack_gate_up(gate, dtype, L.wgu, (int)I, (int)H, 0, st);
    launch_pack_gate_up(up, dtype, L.wgu, (int)I, (int)H, 1, st);
    launch_convert_to_bf16(down, dtype, L.wdown, (int64_t)n_dn, st, true);
    launch_convert_to_bf16(n1, dtype, L.norm1, (int64_t)H, st);
    launch_convert_to_bf16(n2, dtype, L.norm2, (int64_t)H, st);
    CU_TRY(cudaGetLastError());
    CU_TRY(cudaStreamSynchronize(st));
    MegaLayer ml;
    ml.wqkv = reinterpret_cast<const __half*>(L.wqkv); ml.wo = reinterpret_cast<const __half*>(L.wo);
    ml.wgu = reinterpret_cast<const __half*>(L.wgu); ml.wdown = reinterpret_cast<const __half*>(L.wdown);
    ml.norm1 = L.norm1; ml.norm2 = L.norm2;
    CU_TRY(cudaMemcpy(s->mega_layers + rel, &ml, sizeof(ml), cudaMemcpyHostToDevice));
    L.set = true;
    return 0;
}

static int set_single(b200_shard_t* s, __nv_bfloat16** dst, const void* w, size_t n, int dtype, void* stream, int present,
                      const char* what, bool f16_out = false) {
    if (s == nullptr || w == nullptr) return fail(B200_EINVAL, "[ERROR] null argument");
    if (!present) return fail(B200_ESTATE, "[ERROR] this shard was created without %s", what);
    CU_TRY(cudaSetDevice(s->device));
    if (*dst == nullptr) TRY(dev_alloc(s->allocs, dst, n));
    launch_convert_to_bf16(w, dtype, *dst, (int64_t)n, (cudaStream_t)stream, f16_out);
    CU_TRY(cudaGetLastError());
    CU_TRY(cudaStreamSynchronize((cudaStream_t)stream));
    return 0;
}
extern "C" B200_API int b200_shard_set_final_norm(b200_shard_t* s, const void* w, int dtype, void* stream) {
    return set_single(s, &s->final_norm, w, s ? s->cfg.hidden_size : 0, dtype, stream, s ? s->cfg.has_final_norm : 0, "a final norm");
}
extern "C" B200_API int b200_shard_set_lm_head(b200_shard_t* s, const void* w, int dtype, void* stream) {
    return set_single(s, &s->lm_head, w, s ? (size_t)s->cfg.vocab_size * s->cfg.hidden_size : 0, dtype, stream,
                      s ? s->cfg.has_lm_head : 0, "an lm_head", true);
}
extern "C" B200_API int b200_shard_set_embedding(b200_shard_t* s, const void* w, int dtype, void* stream) {
    return set_single(s, &s->embedding, w, s ? (size_t)s->cfg.vocab_size * s->cfg.hidden_size : 0, dtype, stream,
                      s ? s->cfg.has_embedding : 0, "an embedding layer");
}

// ------------------------------------------------------------------------------------- KV
extern "C" B200_API int b200_kv_create(const b200_shard_t* s, int max_slots, int page_size, int n_pages, b200_kv_t** out) {
    if (s == nullptr || out == nullptr) return fail(B200_EINVAL, "[ERROR] null argument");
    if (max_slots < 1 || page_size < 64 || page_size % 64 != 0 || n_pages < 1)
        return fail(B200_EINVAL, "[ERROR] need max_slots >= 1, page_size a multiple of 64, n_pages >= 1");
    CU_TRY(cudaSetDevice(s->device));
    static uint64_t next_gen = 0;
    b200_kv* kv = new b200_kv();
    kv->shard = const_cast<b200_shard*>(s);
    kv->gen = __atomic_add_fetch(&next_gen, 1, __ATOMIC_RELAXED);
    KvDev& d = kv->dev;
    d.max_slots = max_slots;
    d.page_size = page_size;
    d.n_pages = n_pages;
    d.max_pages_per_slot = n_pages;  // a single slot may take the whole pool, up to the attention kernels' reach
    if ((int64_t)d.max_pages_per_slot * page_size > (int64_t)kMaxSplits * kMaxChunkKeys)
        d.max_pages_per_slot = (int)((int64_t)kMaxSplits * kMaxChunkKeys / page_size);
    kv->layer_stride = (int64_t)n_pages * s->cfg.n_kv_heads * page_size * 128;
    int r = 0;
    r |= dev_alloc(kv->allocs, &d.seq_lens, (size_t)max_slots);
    r |= dev_alloc(kv->allocs, &d.block_table, (size_t)max_slots * d.max_pages_per_slot);
    r |= dev_alloc(kv->allocs, &d.pages_held, (size_t)max_slots);
    r |= dev_alloc(kv->allocs, &d.free_stack, (size_t)n_pages);
    r |= dev_alloc(kv->allocs, &d.free_top, 1);
    r |= dev_alloc(kv->allocs, &d.error, 1);
    r |= dev_alloc(kv->allocs, &kv->k_pool, (size_t)kv->layer_stride * s->cfg.n_layers);
    r |= dev_alloc(kv->allocs, &kv->v_pool, (size_t)kv->layer_stride * s->cfg.n_layers);
    if (r != 0) { b200_kv_destroy(kv); return B200_ENOMEM; }
    // pages past a sequence's end are read (and masked) by the attention kernels: start from zeros
    CU_TRY(cudaMemset(kv->k_pool, 0, (size_t)kv->layer_stride * s->cfg.n_layers * 2));
    CU_TRY(cudaMemset(kv->v_pool, 0, (size_t)kv->layer_stride * s->cfg.n_layers * 2));
    launch_kv_init(d, 0);
    CU_TRY(cudaGetLastError());
    CU_TRY(cudaDeviceSynchronize());
    kv->h_len.assign(max_slots, 0);
    kv->h_pages.assign(max_slots, 0);
    kv->h_free = n_pages;
    *out = kv;
    return 0;
}
extern "C" B200_API int b200_kv_destroy(b200_kv_t* kv) {
    if (kv == nullptr) return 0;
    cudaSetDevice(kv->shard->device);
    cudaDeviceSynchronize();
    drop_graphs(kv->shard, kv->gen);  // their nodes point into the pools freed below
    for (void* p : kv->allocs) cudaFree(p);
    delete kv;
    return 0;
}
extern "C" B200_API int b200_kv_reset(b200_kv_t* kv, int slot, void* stream) {
    if (kv == nullptr) return fail(B200_EINVAL, "[ERROR] null kv");
    if (slot < -1 || slot >= kv->dev.max_slots) return fail(B200_EINVAL, "[ERROR] kv slot %d out of range", slot);
    CU_TRY(cudaSetDevice(kv->shard->device));
    launch_kv_reset(kv->dev, slot, (cudaStream_t)stream);
    CU_TRY(cudaGetLastError());
    const int lo = slot < 0 ? 0 : slot, hi = slot < 0 ? kv->dev.max_slots : slot + 1;
    for (int i = lo; i < hi; ++i) {
        kv->h_free += kv->h_pages[i];
        kv->h_pages[i] = 0;
        kv->h_len[i] = 0;
    }
    return 0;
}
extern "C" B200_API int b200_kv_seq_len(const b200_kv_t* kv, int slot) {
    if (kv == nullptr || slot < 0 || slot >= kv->dev.max_slots) return fail(B200_EINVAL, "[ERROR] kv slot out of range");
    return kv->h_len[slot];
}

extern "C" B200_API int b200_kv_error(b200_kv_t* kv) {
    if (kv == nullptr) return fail(B200_EINVAL, "[ERROR] null kv");
    int v = 0;
    CU_TRY(cudaSetDevice(kv->shard->device));
    CU_TRY(cudaMemcpy(&v, kv->dev.error, sizeof(int), cudaMemcpyDeviceToHost));  // synchronises with the kernels that may set it
    if (v == 1) g_err = "[ERROR] KV page pool exhausted on the device (a captured graph was replayed past the pool's capacity)";
    if (v == 2) g_err = "[ERROR] KV slot capacity exceeded on the device";
    return v;
}

// ------------------------------------------------------------------------------------- forward
// split-K factor of decode attention: about four CTAs per SM, at least what the per-CTA key bound needs
static int attn_splits(const b200_shard* s, const b200_kv* kv, int M) {
    static int forced = -1;
    if (forced < 0) {
        const char* e = getenv("B200SHARD_ATTN_SPLITS");
        forced = e ? atoi(e) : 0;
    }
    // latency-bound kernel; measured sweet spot on B200 (32 heads): 16 splits for one row, 8 for 2..8 rows
    int ns = (M == 1) ? (4 * s->num_sms + s->cfg.n_heads - 1) / s->cfg.n_heads : 8;
    // one cluster of 8 / 4 / 2 CTAs per (row, head), partials over DSMEM: about 256 CTAs in all (measured in situ at 128- and
    // 1024-token contexts: 4 rows 100.7 -> 95.5 us per layer, 8 rows 110.6 -> 103.1 with 2 instead of 8 CTAs per head)
    if (attn_decode_cluster_splits() > 0) ns = M == 1 ? 8 : (M < 4 ? 4 : 2);
    if (M > kMaxGemvRows) {
        // batched decode: rows x heads already gives hundreds of CTAs; split only as far as ~4 CTAs per SM needs,
        // and never beyond what the scratch for the partials (sized for 8 rows x 16 splits) holds
        ns = (4 * s->num_sms + M * s->cfg.n_heads - 1) / (M * s->cfg.n_heads);
        const int cap = (kMaxGemvRows * kMaxSplits) / M;
        if (ns > cap) ns = cap;
        if (ns >= 8) ns = attn_decode_cluster_splits() > 0 ? 8 : ns;
        if (attn_decode_cluster_splits() > 0 && ns > 1 && ns < 8) ns = ns >= 4 ? 4 : 2;   // cluster sizes are powers of two
    }
    if (forced > 0 && M <= kMaxGemvRows) ns = forced;
    // (the online-softmax kernel streams any number of keys per CTA: no minimum split count follows from the context length)
    (void)kv;
    return ns < 1 ? 1 : (ns > kMaxSplits ? kMaxSplits : ns);
}
// does the split-K scratch (8 rows x 16 splits of partials per head) hold this call?  The cluster variant uses none.
static bool attn_scratch_fits(int M, int ns) {
    const bool cluster = attn_decode_cluster_splits() > 0 && (ns == 2 || ns == 4 || ns == 8);   // the cluster variant uses no scratch
    return cluster || ns == 1 || (int64_t)ns * M <= kMaxGemvRows * kMaxSplits;
}

// Rows from which single-token attention runs on the tensor cores with TMA-staged K/V (attn_decode_tc.cu).
// B200SHARD_ATTN_DECODE_TC=n moves the threshold, 0 turns it off (A/B against the CUDA-core kernel).
static int attn_decode_tc_min_rows() {
    static int v = -1;
    if (v < 0) {
        const char* e = getenv("B200SHARD_ATTN_DECODE_TC");
        v = e ? atoi(e) : 8;
        if (v < 0) v = 0;
    }
    return v;
}
// Measured in the layer (profiles/attn_decode_tc_r2.md): from 16 rows the tensor-core kernel wins at every context length
// (16 rows x 640 keys: 143 -> 124 us per layer), at 8..15 rows up to about a thousand keys (8 x 640: 113 -> 104); below 8
// rows, and for 8..15 rows on longer contexts, the CUDA-core kernel with its 2-CTA clusters per head is as fast or faster.
static bool use_attn_decode_tc(int M, int max_keys) {
    const int lo = attn_decode_tc_min_rows();
    if (lo == 0 || M < lo) return false;
    (void)max_keys;
    return true;
}

static bool stream_gemm_enabled() {  // B200SHARD_STREAM_GEMM=0: 9..32-row decode takes the 128 x 128 tile GEMMs instead (A/B)
    static int v = -1;
    if (v < 0) {
        const char* e = getenv("B200SHARD_STREAM_GEMM");
        v = (e != nullptr && e[0] == '0') ? 0 : 1;
    }
    return v == 1;
}

static int gemv_row_limit(const b200_shard* s) {
    const int kmax = s->cfg.intermediate_size > s->cfg.hidden_size ? s->cfg.intermediate_size : s->cfg.hidden_size;
    static int cap = -1;  // B200SHARD_GEMV_ROWS: rows up to which the CUDA-core / mma.sync GEMVs are used (A/B against the tcgen05 path)
    if (cap < 0) {
        const char* e = getenv("B200SHARD_GEMV_ROWS");
        cap = e ? atoi(e) : kGemvRowsDefault;
        if (cap < 0 || cap > kMaxGemvRows) cap = kGemvRowsDefault;   // 0: even a single row takes the tensor-core streaming GEMM (experiment)
    }
    int lim = cap;
    while (lim > 0 && gemv_smem_bytes(lim, kmax) > 220 * 1024) --lim;
    return lim;
}

static bool mega_enabled() {
    static int v = -1;
    if (v < 0) { const char* e = getenv("B200SHARD_MEGA"); v = (e != nullptr && e[0] == '1') ? 1 : 0; }
    return v == 1;
}

static int check_call(b200_shard_t* s, b200_kv_t* kv, const int32_t* h_slots, const void* in, int B, int S, SlotList* sl) {
    if (s == nullptr || kv == nullptr || h_slots == nullptr || in == nullptr) return fail(B200_EINVAL, "[ERROR] null argument");
    if (kv->shard != s) return fail(B200_ESTATE, "[ERROR] this KV cache belongs to another shard");
    if (B < 1 || B > s->cfg.max_batch || S < 1 || (int64_t)B * S > s->cfg.max_tokens)
        return fail(B200_EINVAL, "[ERROR] batch %d x seq_len %d exceeds the shard's limits (max_batch %d, max_tokens %d)", B, S,
                    s->cfg.max_batch, s->cfg.max_tokens);
    for (auto& l : s->layers)
        if (!l.set) return fail(B200_ESTATE, "[ERROR] a layer of this shard has no weights loaded");
    int need_total = 0;
    for (int b = 0; b < B; ++b) {  // argument errors first (ValueError), capacity second (out-of-memory)
        const int slot = h_slots[b];
        if (slot < 0 || slot >= kv->dev.max_slots) return fail(B200_EINVAL, "[ERROR] kv slot %d out of range", slot);
        for (int c = 0; c < b; ++c)
            if (h_slots[c] == slot) return fail(B200_EINVAL, "[ERROR] kv slot %d used by two rows of one batch", slot);
        if (kv->h_len[slot] != kv->h_len[h_slots[0]])
            return fail(B200_EINVAL, "[ERROR] rows of one batch must have equal cached lengths");
    }
    for (int b = 0; b < B; ++b) {
        const int slot = h_slots[b];
        const int need = (kv->h_len[slot] + S + kv->dev.page_size - 1) / kv->dev.page_size;
        if (need > kv->dev.max_pages_per_slot) return fail(B200_ENOMEM, "[ERROR] KV slot capacity exceeded");
        need_total += need > kv->h_pages[slot] ? need - kv->h_pages[slot] : 0;
        sl->v[b] = slot;
    }
    if (need_total > kv->h_free) return fail(B200_ENOMEM, "[ERROR] KV page pool exhausted (%d pages needed, %d free)", need_total, kv->h_free);
    return 0;
}

// RMSNorm folded into the streaming GEMMs: up to B200SHARD_FOLD_ROWS rows (default: all 32 of the path).  The consumer's
// row statistics either come from re-reading the residual block in every consuming CTA, or as per-CTA partial sums of
// squares left by the producer (row_ss).  B200SHARD_ROW_SS: 0 never, 1 from 17 rows (default: below that the re-read hides
// behind the weight stream and the partials cost 0.3-0.5 us per layer), 2 always.
static int row_ss_mode() {
    static int v = -1;
    if (v < 0) { const char* e = getenv("B200SHARD_ROW_SS"); v = e != nullptr ? atoi(e) : 1; }
    return gemm_streamk_active() ? 0 : v;
}
static bool row_ss_for(int M) { return row_ss_mode() == 2 || (row_ss_mode() == 1 && M > 16); }
static int fold_row_limit() {
    static int v = -1;
    if (v < 0) { const char* e = getenv("B200SHARD_FOLD_ROWS"); v = e != nullptr ? atoi(e) : kMaxStreamRows; }
    return row_ss_mode() == 0 && v > 16 ? 16 : v;   // without the partials, folding beyond 16 rows is a net loss (136 -> 138 us)
}
// Batched single-token tail: the last down_proj writes x * w_final as the fp16 operand of lm_head and the lm_head GEMM
// applies 1 / rms in its epilogue — no separate final-norm or cast launch (same fold as between the layers).
static bool tail_fold(const b200_shard* s, int B, int S, bool stream_path) {
    return stream_path && S == 1 && B <= fold_row_limit() && s->cfg.has_final_norm && s->final_norm != nullptr && s->cfg.hidden_size % 64 == 0;
}

// enqueue every kernel of one forward (+ optional tail); no host synchronisation
static int enqueue_forward(b200_shard_t* s, b200_kv_t* kv, const SlotList& sl, const void* in, const int64_t* tok_in,
                           int B, int S, int causal, void* out, int64_t* tok, float* margin, bool with_tail,
                           cudaStream_t st) {
    const b200_shard_config_t& c = s->cfg;
    const int M = B * S, H = c.hidden_size, I = c.intermediate_size;
    const int64_t head_stride = (int64_t)kv->dev.page_size * 128;
    static int fused_prologue = -1;   // B200SHARD_FUSED_PROLOGUE=0: the four separate launches (A/B)
    if (fused_prologue < 0) { const char* e = getenv("B200SHARD_FUSED_PROLOGUE"); fused_prologue = (e != nullptr && e[0] == '0') ? 0 : 1; }
    if (fused_prologue == 1 && S == 1 && B <= kMaxGemvRows && H % 8 == 0) {
        launch_decode_prologue(kv->dev, sl, B, causal, s->inv, c.n_kv_heads, s->tm, in, s->io_dtype, tok_in, s->embedding, c.vocab_size,
                               H, s->x, st);
        s->n_launch += 1;
    } else {
        if (tok_in != nullptr) {  // entry == head node: embed_tokens (utils/node_worker.py:215,302) straight into the workspace
            launch_embed(s->embedding, tok_in, M, H, c.vocab_size, s->xn, B200_DT_BF16, st);
            s->n_launch += 1;
            in = s->xn;
        }
        launch_kv_alloc(kv->dev, sl, B, S, s->tm.call_base, st);
        launch_prep(kv->dev, sl, B, S, causal, s->inv, c.n_kv_heads, s->tm, st);
        launch_load_residual(in, tok_in != nullptr ? B200_DT_BF16 : s->io_dtype, s->x, (int64_t)M * H, st);
        s->n_launch += 3;
    }
    const bool gemv_path = M <= gemv_row_limit(s);
    s->tail_xw_ready = false;
    int max_keys = 0;   // longest context any row of this call attends to (the host mirrors the device lengths)
    for (int b = 0; b < B; ++b) max_keys = kv->h_len[sl.v[b]] + S > max_keys ? kv->h_len[sl.v[b]] + S : max_keys;
    const bool stream_path = !gemv_path && M <= kMaxStreamRows && stream_gemm_enabled() && H % 64 == 0 && I % 64 == 0 &&
                             attn_scratch_fits(M, attn_splits(s, kv, M));
    // One row, one token: every layer in ONE persistent kernel (decode_mega.cu) whose weight stream runs through the phase
    // boundaries.  B200SHARD_MEGA=0 keeps the five launches per layer (A/B).
    bool mega_done = false;
    if (gemv_path && M == 1 && mega_enabled()) {
        MegaParams mp{};
        mp.layers = s->mega_layers; mp.n_layers = c.n_layers; mp.H = H; mp.I = I; mp.n_heads = c.n_heads; mp.n_kv = c.n_kv_heads;
        mp.qkv_rows = s->qkv_rows; mp.eps = c.rms_eps; mp.x = s->x; mp.qkv_raw = s->q_f; mp.attn_out = s->attn_f; mp.gu_raw = s->act_f;
        mp.k_pool = kv->k_pool; mp.v_pool = kv->v_pool; mp.layer_stride = kv->layer_stride; mp.head_stride = head_stride;
        mp.block_table = kv->dev.block_table; mp.slot = sl.v[0]; mp.max_pages_per_slot = kv->dev.max_pages_per_slot;
        mp.page_size = kv->dev.page_size; mp.t_end = s->tm.t_end; mp.kv_off = s->tm.kv_off; mp.cos = s->tm.cos; mp.sin = s->tm.sin;
        mp.part = s->attn_part;
        mp.tickets = reinterpret_cast<unsigned int*>(s->attn_part + (size_t)kMaxGemvRows * c.n_heads * kMaxSplits * 130);
        mp.bar = s->mega_bar;
        const int rc = launch_decode_mega(mp, s->num_sms, st);
        if (rc < 0) return fail(B200_ECUDA, "[ERROR] %s", decode_mega_last_error());
        if (rc == 0) { mega_done = true; s->n_launch += 1; }
    }
    for (int l = 0; l < c.n_layers && !mega_done; ++l) {
        const LayerW& L = s->layers[l];
        __nv_bfloat16* kp = kv->k_pool + (int64_t)l * kv->layer_stride;
        __nv_bfloat16* vp = kv->v_pool + (int64_t)l * kv->layer_stride;
        if (gemv_path) {
            GemvParams g{};
            g.W = L.wqkv; g.x = s->x; g.norm_w = L.norm1; g.eps = c.rms_eps; g.N = s->qkv_rows; g.K = H;
            g.q_out = s->q_f; g.k_pool = kp; g.v_pool = vp; g.kv_off = s->tm.kv_off; g.cos = s->tm.cos; g.sin = s->tm.sin;
            g.n_heads = c.n_heads; g.n_kv = c.n_kv_heads; g.head_stride = head_stride;
            launch_gemv(g, M, EPI_QKV, true, s->num_sms, st);
            AttnDecodeParams a{};
            a.q = s->q_f; a.k_pool = kp; a.v_pool = vp; a.block_table = kv->dev.block_table; a.t_end = s->tm.t_end; a.call_base = s->tm.call_base;
            a.slots = sl; a.S = S; a.n_heads = c.n_heads; a.n_kv = c.n_kv_heads; a.page_size = kv->dev.page_size;
            a.max_pages_per_slot = kv->dev.max_pages_per_slot;
            a.n_splits = attn_splits(s, kv, M);
            a.part = s->attn_part; a.out = s->attn_f;
            a.tickets = reinterpret_cast<unsigned int*>(s->attn_part + (size_t)kMaxGemvRows * c.n_heads * kMaxSplits * 130);
            launch_attn_decode(a, M, st);
            GemvParams o{};
            o.W = L.wo; o.x = s->attn_f; o.N = H; o.K = s->nq; o.out = s->x;
            launch_gemv(o, M, EPI_RESID, false, s->num_sms, st);
            GemvParams gu{};
            gu.W = L.wgu; gu.x = s->x; gu.norm_w = L.norm2; gu.eps = c.rms_eps; gu.N = 2 * I; gu.K = H; gu.out = s->act_f;
            launch_gemv(gu, M, EPI_SWIGLU, true, s->num_sms, st);
            GemvParams d{};
            d.W = L.wdown; d.x = s->act_f; d.N = H; d.K = I; d.out = s->x;
            launch_gemv(d, M, EPI_RESID, false, s->num_sms, st);
            s->n_launch += 5;
        } else if (stream_path) {
            // 4..32 token rows (batched decode, short prompts): still bound by reading the weights once, now on the tensor cores
            // (gemm_stream_tc.cu); operands fp16, residual stream and q fp32, K/V rounded to bf16 once
            // RMSNorm is folded into the GEMMs: the projection that updates the residual stream also writes x * w_norm of the
            // NEXT norm as the fp16 operand (s->xn), the consuming projection applies 1 / rms in its epilogue.  Only the very
            // first norm of the call has no producer and runs as its own (un-normalised: weight only) pass.
            // Every consuming CTA re-reads the M x H residual block for the statistics: free up to 16 rows (138 -> 111 us per
            // layer at 16, 100 -> 89 at 4), a net loss at 32 (136 -> 138), where the two separate norm passes stay.
            const bool fold = M <= fold_row_limit();
            // partial sums of squares: buffer A travels down_proj (or the scale pass) -> QKV / lm_head, buffer B o_proj -> gate/up
            const int parts_o = gemm_stream_parts(H, s->nq, s->num_sms), parts_d = gemm_stream_parts(H, I, s->num_sms);
            const bool rss = fold && row_ss_for(M) && parts_o <= kRowSsParts && parts_d <= kRowSsParts;
            float* ss_a = s->row_ss;
            float* ss_b = s->row_ss + (size_t)kMaxStreamRows * kRowSsParts;
            if (!fold) launch_rmsnorm_bf16(s->x, L.norm1, c.rms_eps, M, H, s->xn, st, B200_DT_F16);
            else if (l == 0) launch_scale_f16(s->x, L.norm1, M, H, reinterpret_cast<__half*>(s->xn), rss ? ss_a : nullptr, st);
            StreamGemmParams g{};
            g.W = L.wqkv; g.X = s->xn; g.M = M; g.N = s->qkv_rows; g.K = H;
            g.q_out = s->q_f; g.k_pool = kp; g.v_pool = vp; g.kv_off = s->tm.kv_off; g.cos = s->tm.cos; g.sin = s->tm.sin;
            g.n_heads = c.n_heads; g.n_kv = c.n_kv_heads; g.head_stride = head_stride; g.ws = s->sk_ws; g.ws_flags = s->sk_flags;
            if (fold) { g.x_res = s->x; g.eps = c.rms_eps; }
            if (rss) { g.row_ss = ss_a; g.row_ss_parts = l == 0 ? 1 : parts_d; }
            if (launch_gemm_stream(g, EPI_QKV, s->num_sms, st) != 0) return fail(B200_ECUDA, "[ERROR] %s", gemm_stream_last_error());
            AttnDecodeParams a{};
            a.q = s->q_f; a.k_pool = kp; a.v_pool = vp; a.block_table = kv->dev.block_table; a.t_end = s->tm.t_end; a.call_base = s->tm.call_base;
            a.slots = sl; a.S = S; a.n_heads = c.n_heads; a.n_kv = c.n_kv_heads; a.page_size = kv->dev.page_size;
            a.max_pages_per_slot = kv->dev.max_pages_per_slot;
            a.n_splits = attn_splits(s, kv, M);
            a.part = s->attn_part; a.out = s->attn_f; a.out_h = reinterpret_cast<__half*>(s->attn_b);
            a.tickets = reinterpret_cast<unsigned int*>(s->attn_part + (size_t)kMaxGemvRows * c.n_heads * kMaxSplits * 130);
            int tc_rc = 1;
            if (use_attn_decode_tc(M, max_keys))
                tc_rc = launch_attn_decode_tc(a, M, max_keys, (int64_t)kv->dev.n_pages * c.n_kv_heads * kv->dev.page_size, st);
            if (tc_rc < 0) return fail(B200_ECUDA, "[ERROR] %s", gemm_error());
            if (tc_rc == 1) launch_attn_decode(a, M, st);
            StreamGemmParams o{};
            o.W = L.wo; o.X = s->attn_b; o.M = M; o.N = H; o.K = s->nq; o.out = s->x; o.ws = s->sk_ws; o.ws_flags = s->sk_flags;
            if (fold) { o.next_norm_w = L.norm2; o.xw_out = reinterpret_cast<__half*>(s->xn); }
            if (rss) o.row_ss_out = ss_b;
            if (launch_gemm_stream(o, EPI_RESID, s->num_sms, st) != 0) return fail(B200_ECUDA, "[ERROR] %s", gemm_stream_last_error());
            if (!fold) launch_rmsnorm_bf16(s->x, L.norm2, c.rms_eps, M, H, s->xn, st, B200_DT_F16);
            StreamGemmParams gu{};
            gu.W = L.wgu; gu.X = s->xn; gu.M = M; gu.N = 2 * I; gu.K = H; gu.act_h = reinterpret_cast<__half*>(s->act_b); gu.ws = s->sk_ws; gu.ws_flags = s->sk_flags;
            if (fold) { gu.x_res = s->x; gu.eps = c.rms_eps; }
            if (rss) { gu.row_ss = ss_b; gu.row_ss_parts = parts_o; }
            if (launch_gemm_stream(gu, EPI_SWIGLU, s->num_sms, st) != 0) return fail(B200_ECUDA, "[ERROR] %s", gemm_stream_last_error());
            StreamGemmParams d{};
            d.W = L.wdown; d.X = s->act_b; d.M = M; d.N = H; d.K = I; d.out = s->x; d.ws = s->sk_ws; d.ws_flags = s->sk_flags;
            if (fold && l + 1 < c.n_layers) { d.next_norm_w = s->layers[l + 1].norm1; d.xw_out = reinterpret_cast<__half*>(s->xn); }
            else if (with_tail && tail_fold(s, B, S, true)) {
                d.next_norm_w = s->final_norm; d.xw_out = reinterpret_cast<__half*>(s->xn);
                s->tail_xw_ready = true;
            }
            if (rss && d.xw_out != nullptr) { d.row_ss_out = ss_a; s->tail_ss_parts = parts_d; }
            else s->tail_ss_parts = 0;
            if (launch_gemm_stream(d, EPI_RESID, s->num_sms, st) != 0) return fail(B200_ECUDA, "[ERROR] %s", gemm_stream_last_error());
            s->n_launch += fold ? 5 + (l == 0 ? 1 : 0) : 7;
        } else {
            // activations feeding the tensor cores are fp16 (10 mantissa bits) against exact bf16 weights; q reaches
            // attention as a (hi, lo) bf16 pair; K/V are rounded to bf16 exactly once, inside the QKV epilogue
            launch_rmsnorm_bf16(s->x, L.norm1, c.rms_eps, M, H, s->xn, st, B200_DT_F16);
            GemmParams g{};
            g.A = s->xn; g.W = L.wqkv; g.M = M; g.N = s->qkv_rows; g.K = H; g.out = nullptr; g.epi = GEMM_QKV_ROPE; g.ab_f16 = 1;
            g.q_hi = s->qrot; g.q_lo = s->qrot_lo; g.k_pool = kp; g.v_pool = vp; g.kv_off = s->tm.kv_off;
            g.cos = s->tm.cos; g.sin = s->tm.sin; g.n_heads = c.n_heads; g.n_kv = c.n_kv_heads; g.head_stride = head_stride;
            if (launch_gemm(g, st) != 0) return fail(B200_ECUDA, "[ERROR] %s", gemm_error());
            AttnPrefillParams ap{};
            ap.q = s->qrot; ap.q_lo = s->qrot_lo; ap.out_f16 = 1; ap.k_pool = kp; ap.v_pool = vp; ap.block_table = kv->dev.block_table; ap.call_base = s->tm.call_base;
            ap.slots = sl; ap.B = B; ap.S = S; ap.causal = causal; ap.n_heads = c.n_heads; ap.n_kv = c.n_kv_heads;
            ap.page_size = kv->dev.page_size; ap.max_pages_per_slot = kv->dev.max_pages_per_slot; ap.out = s->attn_b;
            static int attn_tc = -1;
            if (attn_tc < 0) { const char* e = getenv("B200SHARD_ATTN_TC"); attn_tc = (e != nullptr && e[0] == '0') ? 0 : 1; }
            if (attn_tc == 1) {  // tcgen05 + TMA kernel; B200SHARD_ATTN_TC=0 keeps the mma.sync kernel for A/B checks
                if (launch_attn_prefill_tc(ap, (int64_t)kv->dev.n_pages * c.n_kv_heads * kv->dev.page_size, st) != 0)
                    return fail(B200_ECUDA, "[ERROR] %s", gemm_error());
            } else {
                launch_attn_prefill(ap, st);
            }
            GemmParams o{};
            o.A = s->attn_b; o.W = L.wo; o.M = M; o.N = H; o.K = s->nq; o.out = s->x; o.epi = GEMM_RESID_F32; o.ab_f16 = 1;
            if (launch_gemm(o, st) != 0) return fail(B200_ECUDA, "[ERROR] %s", gemm_error());
            launch_rmsnorm_bf16(s->x, L.norm2, c.rms_eps, M, H, s->xn, st, B200_DT_F16);
            GemmParams gu{};
            gu.A = s->xn; gu.W = L.wgu; gu.M = M; gu.N = 2 * I; gu.K = H; gu.out = s->act_b; gu.epi = GEMM_SWIGLU_BF16;
            gu.ab_f16 = 1; gu.out_f16 = 1;
            if (launch_gemm(gu, st) != 0) return fail(B200_ECUDA, "[ERROR] %s", gemm_error());
            GemmParams d{};
            d.A = s->act_b; d.W = L.wdown; d.M = M; d.N = H; d.K = I; d.out = s->x; d.epi = GEMM_RESID_F32; d.ab_f16 = 1;
            if (launch_gemm(d, st) != 0) return fail(B200_ECUDA, "[ERROR] %s", gemm_error());
            s->n_launch += 7;
        }
    }
    if (out != nullptr) {  // final RMSNorm when this is the tail shard (utils/shard_loader.py:76-77), else a plain cast
        launch_rmsnorm_bf16(s->x, c.has_final_norm ? s->final_norm : nullptr, c.rms_eps, M, H, out, st, s->io_dtype);
        s->n_launch += 1;
    }
    if (with_tail && B > gemv_row_limit(s) && !s->tail_xw_ready) {  // batched tail goes through the GEMM; small batches fuse the norm into the GEMV
        launch_lastrow_norm_f32(s->x, c.has_final_norm ? s->final_norm : nullptr, c.rms_eps, B, S, H, s->lastrow, st);
        s->n_launch += 1;
    }
    (void)tok; (void)margin;
    return 0;
}

// from_residual: the last rows are still in the fp32 residual stream s->x ([B,S,H]); otherwise they are in s->lastrow
static int enqueue_tail(b200_shard_t* s, int B, int S, bool from_residual, int64_t* tok, float* margin, cudaStream_t st) {
    const b200_shard_config_t& c = s->cfg;
    if (B <= gemv_row_limit(s)) {
        GemvParams g{};
        g.W = s->lm_head; g.N = c.vocab_size; g.K = c.hidden_size; g.out = s->logits;
        if (from_residual) {  // final RMSNorm (utils/shard_loader.py:76-77) as the GEMV's prologue, row S-1 of every batch row
            g.x = s->x + (size_t)(S - 1) * c.hidden_size; g.x_stride = (int64_t)S * c.hidden_size;
            g.norm_w = c.has_final_norm ? s->final_norm : nullptr; g.eps = c.rms_eps;
            launch_gemv(g, B, EPI_STORE, c.has_final_norm != 0, s->num_sms, st);
        } else {
            g.x = s->lastrow;
            launch_gemv(g, B, EPI_STORE, false, s->num_sms, st);
        }
    } else if (from_residual && s->tail_xw_ready) {   // the last down_proj of this very call left x * w_final in s->xn
        StreamGemmParams g{};
        g.W = s->lm_head; g.X = s->xn; g.M = B; g.N = c.vocab_size; g.K = c.hidden_size; g.out = s->logits; g.ws = s->sk_ws; g.ws_flags = s->sk_flags;
        g.x_res = s->x; g.eps = c.rms_eps;
        if (s->tail_ss_parts > 0) { g.row_ss = s->row_ss; g.row_ss_parts = s->tail_ss_parts; }
        if (launch_gemm_stream(g, EPI_STORE, s->num_sms, st) != 0) return fail(B200_ECUDA, "[ERROR] %s", gemm_stream_last_error());
    } else {
        launch_rmsnorm_bf16(s->lastrow, nullptr, 0.f, B, c.hidden_size, s->lastrow_b, st, B200_DT_F16);
        if (B <= kMaxStreamRows && stream_gemm_enabled() && c.hidden_size % 64 == 0) {
            StreamGemmParams g{};
            g.W = s->lm_head; g.X = s->lastrow_b; g.M = B; g.N = c.vocab_size; g.K = c.hidden_size; g.out = s->logits; g.ws = s->sk_ws; g.ws_flags = s->sk_flags;
            if (launch_gemm_stream(g, EPI_STORE, s->num_sms, st) != 0) return fail(B200_ECUDA, "[ERROR] %s", gemm_stream_last_error());
        } else {
            GemmParams g{};
            g.A = s->lastrow_b; g.W = s->lm_head; g.M = B; g.N = c.vocab_size; g.K = c.hidden_size; g.out = s->logits;
            g.epi = GEMM_STORE_F32; g.ab_f16 = 1;
            if (launch_gemm(g, st) != 0) return fail(B200_ECUDA, "[ERROR] %s", gemm_error());
        }
        s->n_launch += 1;
    }
    launch_argmax(s->logits, B, c.vocab_size, tok, margin, st);
    s->n_launch += 2;
    return 0;
}

// Vocabulary-split tail, first half (the tail shard): the message for the next stage = the last fp32 residual row of every
// batch row, then best / second-best score and best index over lm_head rows [0, split).  Same GEMV, same fused final norm as
// enqueue_tail, so every logit has the bits it would have in the unsplit call.
// Scores of lm_head rows W [n_rows, H] for B token rows -> s->logits [B][n_rows].  Up to the GEMV row limit: the GEMV with the
// final norm fused (x: fp32 rows, x_stride apart), the very kernel of the unsplit tail.  More rows: normalise into fp16 and run
// the streaming tensor-core GEMM (x must then be contiguous [B][H]); both halves of a split vocabulary take the same route, so
// their scores are comparable to the last bit.
static int score_rows(b200_shard_t* s, const __nv_bfloat16* W, int n_rows, const __nv_bfloat16* norm_w, const float* x, int64_t x_stride,
                      int B, cudaStream_t st) {
    const b200_shard_config_t& c = s->cfg;
    const int H = c.hidden_size;
    if (B <= gemv_row_limit(s)) {
        GemvParams g{};
        g.W = W; g.N = n_rows; g.K = H; g.out = s->logits; g.x = x; g.x_stride = x_stride; g.norm_w = norm_w; g.eps = c.rms_eps;
        launch_gemv(g, B, EPI_STORE, norm_w != nullptr, s->num_sms, st);
        s->n_launch += 1;
        return 0;
    }
    if (x_stride != H) return fail(B200_EINVAL, "[ERROR] batched scoring needs contiguous rows");
    launch_rmsnorm_bf16(x, norm_w, c.rms_eps, B, H, s->lastrow_b, st, B200_DT_F16);
    if (B <= kMaxStreamRows && stream_gemm_enabled() && H % 64 == 0) {
        StreamGemmParams g{};
        g.W = W; g.X = s->lastrow_b; g.M = B; g.N = n_rows; g.K = H; g.out = s->logits; g.ws = s->sk_ws; g.ws_flags = s->sk_flags;
        if (launch_gemm_stream(g, EPI_STORE, s->num_sms, st) != 0) return fail(B200_ECUDA, "[ERROR] %s", gemm_stream_last_error());
    } else {
        GemmParams g{};
        g.A = s->lastrow_b; g.W = W; g.M = B; g.N = n_rows; g.K = H; g.out = s->logits; g.epi = GEMM_STORE_F32; g.ab_f16 = 1;
        if (launch_gemm(g, st) != 0) return fail(B200_ECUDA, "[ERROR] %s", gemm_error());
    }
    s->n_launch += 2;
    return 0;
}

static int enqueue_tail_partial(b200_shard_t* s, int B, int S, int split, void* msg, cudaStream_t st) {
    const b200_shard_config_t& c = s->cfg;
    const int H = c.hidden_size;
    if (B > c.max_batch) return fail(B200_EINVAL, "[ERROR] more rows than max_batch");
    if (split < 1 || split >= c.vocab_size) return fail(B200_EINVAL, "[ERROR] split must lie inside the vocabulary");
    float* mx = reinterpret_cast<float*>(msg);
    float* best = mx + (size_t)B * H;
    float* second = best + B;
    int64_t* idx = reinterpret_cast<int64_t*>(second + B);
    launch_lastrow_norm_f32(s->x, nullptr, 0.f, B, S, H, mx, st);
    const __nv_bfloat16* nw = c.has_final_norm ? s->final_norm : nullptr;
    if (B <= gemv_row_limit(s)) { TRY(score_rows(s, s->lm_head, split, nw, s->x + (size_t)(S - 1) * H, (int64_t)S * H, B, st)); }
    else { TRY(score_rows(s, s->lm_head, split, nw, mx, H, B, st)); }
    launch_argmax_range(s->logits, B, split, 0, nullptr, nullptr, nullptr, best, second, idx, nullptr, st);
    s->n_launch += 2;
    return 0;
}

static void account(b200_shard_t* s, b200_kv_t* kv, const SlotList& sl, int B, int S, int causal, bool tail) {
    const b200_shard_config_t& c = s->cfg;
    const uint64_t H = c.hidden_size, I = c.intermediate_size;
    const uint64_t params = (uint64_t)s->qkv_rows * H + H * s->nq + 3 * I * H;
    uint64_t bytes = (uint64_t)c.n_layers * (params * 2 + 2 * H * 2);
    uint64_t flops = (uint64_t)c.n_layers * 2 * params * (uint64_t)B * S;
    for (int b = 0; b < B; ++b) {
        const uint64_t T = (uint64_t)kv->h_len[sl.v[b]] + S;
        bytes += (uint64_t)c.n_layers * 2 * T * s->nkvd * 2;         // K,V read once per call
        bytes += (uint64_t)c.n_layers * 2 * (uint64_t)S * s->nkvd * 2;  // K,V append
        const uint64_t pairs = causal ? ((uint64_t)S * (2 * T - S + 1)) / 2 : (uint64_t)S * T;
        flops += (uint64_t)c.n_layers * 4 * pairs * s->nq;
    }
    if (tail) {
        bytes += (uint64_t)c.vocab_size * H * 2 + H * 2;
        flops += 2ull * c.vocab_size * H * B;
    }
    s->n_bytes += bytes;
    s->n_flops += flops;
    s->n_calls += 1;
}

// partial_split > 0: vocabulary-split tail — instead of the token, `msg` receives the message of enqueue_tail_partial
static int forward_impl(b200_shard_t* s, b200_kv_t* kv, const int32_t* h_slots, const void* in, const int64_t* tok_in,
                        int B, int S, int causal, void* out, int64_t* tok, float* margin, bool with_tail, void* stream,
                        int partial_split = 0, void* msg = nullptr) {
    SlotList sl{};
    TRY(check_call(s, kv, h_slots, tok_in != nullptr ? (const void*)tok_in : in, B, S, &sl));
    if (tok_in != nullptr && (!s->cfg.has_embedding || s->embedding == nullptr))
        return fail(B200_ESTATE, "[ERROR] this node does not have embedding layer while receiving user request.");
    if (with_tail) {
        if (!s->cfg.has_lm_head || s->lm_head == nullptr) return fail(B200_ESTATE, "[ERROR] this shard does not hold the lm_head");
        if (tok == nullptr && partial_split == 0) return fail(B200_EINVAL, "[ERROR] null token output");
        if (partial_split > 0 && msg == nullptr) return fail(B200_EINVAL, "[ERROR] null message output");
    }
    if (s->cfg.has_final_norm && s->final_norm == nullptr) return fail(B200_ESTATE, "[ERROR] final norm weight not loaded");
    CU_TRY(cudaSetDevice(s->device));
    cudaStream_t st = (cudaStream_t)stream;
    account(s, kv, sl, B, S, causal, with_tail);
    if (partial_split > 0) {   // only rows [0, split) of lm_head are read here
        s->n_bytes -= (uint64_t)(s->cfg.vocab_size - partial_split) * s->cfg.hidden_size * 2;
        s->n_flops -= 2ull * (uint64_t)(s->cfg.vocab_size - partial_split) * s->cfg.hidden_size * B;
    }
    bool done = false;
    if (s->graphs && S == 1) {
        GraphKey key;
        memset(&key, 0, sizeof(key));
        key.B = B; key.causal = causal; key.with_argmax = with_tail ? (partial_split > 0 ? 2 + partial_split : 1) : 0;
        // the tensor-core decode attention sizes its score buffer in whole kilo-keys: one graph per bucket of the context length
        if (attn_decode_tc_min_rows() > 0 && B >= attn_decode_tc_min_rows()) key.pad_ = (kv->h_len[sl.v[0]] + S + 1023) / 1024;
        key.in = in; key.out = partial_split > 0 ? msg : out; key.tok = tok; key.margin = margin; key.kv_gen = kv->gen; key.tok_in = tok_in;
        for (int b = 0; b < B; ++b) key.slots[b] = sl.v[b];
        auto it = s->graph_cache.find(key);
        if (it == s->graph_cache.end()) {
            // capture on a side stream so that the caller's stream state is untouched
            cudaStream_t cs;
            CU_TRY(cudaStreamCreateWithFlags(&cs, cudaStreamNonBlocking));
            const uint64_t launches_before = s->n_launch;
            CU_TRY(cudaStreamBeginCapture(cs, cudaStreamCaptureModeThreadLocal));
            int r = enqueue_forward(s, kv, sl, in, tok_in, B, S, causal, out, tok, margin, with_tail, cs);
            if (r == 0 && with_tail) r = partial_split > 0 ? enqueue_tail_partial(s, B, S, partial_split, msg, cs) : enqueue_tail(s, B, S, true, tok, margin, cs);
            cudaGraph_t graph = nullptr;
            cudaError_t ce = cudaStreamEndCapture(cs, &graph);
            const uint64_t captured = s->n_launch - launches_before;
            s->n_launch = launches_before;
            if (r != 0 || ce != cudaSuccess || graph == nullptr) {
                cudaGetLastError();
                cudaStreamDestroy(cs);
                if (r != 0) return r;
                return fail(B200_ECUDA, "[ERROR] CUDA graph capture failed: %s", cudaGetErrorString(ce));
            }
            cudaGraphExec_t exec = nullptr;
            ce = cudaGraphInstantiate(&exec, graph, 0);
            cudaGraphDestroy(graph);
            cudaStreamDestroy(cs);
            if (ce != cudaSuccess) return fail(B200_ECUDA, "[ERROR] cudaGraphInstantiate failed: %s", cudaGetErrorString(ce));
            if (s->graph_cache.size() >= kMaxCachedGraphs) {  // evict the least recently used graph
                auto victim = s->graph_cache.begin();
                for (auto j = s->graph_cache.begin(); j != s->graph_cache.end(); ++j)
                    if (j->second.last_use < victim->second.last_use) victim = j;
                // a replay of the victim may still be running on some stream; destroying an exec waits for nothing, so drain first
                cudaDeviceSynchronize();
                cudaGraphExecDestroy(victim->second.exec);
                s->graph_cache.erase(victim);
            }
            GraphEntry ge;
            ge.exec = exec;
            ge.launches = captured;
            it = s->graph_cache.emplace(key, ge).first;
        }
        it->second.last_use = ++s->graph_clock;
        CU_TRY(cudaGraphLaunch(it->second.exec, st));
        s->n_launch += it->second.launches;
        done = true;
    }
    if (!done) {
        TRY(enqueue_forward(s, kv, sl, in, tok_in, B, S, causal, out, tok, margin, with_tail, st));
        if (with_tail) {
            if (partial_split > 0) { TRY(enqueue_tail_partial(s, B, S, partial_split, msg, st)); }
            else { TRY(enqueue_tail(s, B, S, true, tok, margin, st)); }
        }
        CU_TRY(cudaGetLastError());
    }
    for (int b = 0; b < B; ++b) {  // host mirror of what kv_alloc_kernel does on the device
        const int slot = sl.v[b];
        const int need = (kv->h_len[slot] + S + kv->dev.page_size - 1) / kv->dev.page_size;
        if (need > kv->h_pages[slot]) { kv->h_free -= need - kv->h_pages[slot]; kv->h_pages[slot] = need; }
        kv->h_len[slot] += S;
    }
    return 0;
}

extern "C" B200_API int b200_shard_forward(b200_shard_t* s, b200_kv_t* kv, const int32_t* h_slots, const void* in, int B, int S,
                                  int causal, void* out, void* stream) {
    if (out == nullptr) return fail(B200_EINVAL, "[ERROR] null output");
    return forward_impl(s, kv, h_slots, in, nullptr, B, S, causal, out, nullptr, nullptr, false, stream);
}
extern "C" B200_API int b200_shard_forward_argmax(b200_shard_t* s, b200_kv_t* kv, const int32_t* h_slots, const void* in, int B, int S,
                                         int causal, void* out, int64_t* tok, float* margin, void* stream) {
    return forward_impl(s, kv, h_slots, in, nullptr, B, S, causal, out, tok, margin, true, stream);
}
extern "C" B200_API int b200_shard_step_tokens(b200_shard_t* s, b200_kv_t* kv, const int32_t* h_slots, const int64_t* tok_in, int B,
                                               int S, int causal, void* out, int64_t* tok, float* margin, void* stream) {
    if (tok_in == nullptr) return fail(B200_EINVAL, "[ERROR] null token input");
    const bool tail = s != nullptr && s->cfg.has_lm_head;
    if (!tail && out == nullptr) return fail(B200_EINVAL, "[ERROR] null output");
    return forward_impl(s, kv, h_slots, nullptr, tok_in, B, S, causal, out, tail ? tok : nullptr, tail ? margin : nullptr, tail, stream);
}

extern "C" B200_API int b200_shard_forward_partial(b200_shard_t* s, b200_kv_t* kv, const int32_t* h_slots, const void* in, int B, int S,
                                                   int causal, int split, void* msg_out, void* stream) {
    if (split < 1) return fail(B200_EINVAL, "[ERROR] split must be positive");
    return forward_impl(s, kv, h_slots, in, nullptr, B, S, causal, nullptr, nullptr, nullptr, true, stream, split, msg_out);
}
extern "C" B200_API int b200_shard_set_vocab_tail(b200_shard_t* s, const void* lm_head_rows, const void* final_norm, int split, int dtype,
                                                  void* stream) {
    if (s == nullptr || lm_head_rows == nullptr) return fail(B200_EINVAL, "[ERROR] null argument");
    if (split < 1 || split >= s->cfg.vocab_size) return fail(B200_EINVAL, "[ERROR] split must lie inside the vocabulary");
    if (dtype < 0 || dtype > 2) return fail(B200_EINVAL, "[ERROR] unsupported weight dtype %d", dtype);
    CU_TRY(cudaSetDevice(s->device));
    const size_t H = s->cfg.hidden_size, rows = (size_t)(s->cfg.vocab_size - split);
    if (s->aux_lm_head != nullptr && s->aux_split != split) return fail(B200_ESTATE, "[ERROR] the vocabulary split of a shard cannot change");
    if (s->aux_lm_head == nullptr) TRY(dev_alloc(s->allocs, &s->aux_lm_head, rows * H));
    launch_convert_to_bf16(lm_head_rows, dtype, s->aux_lm_head, (int64_t)(rows * H), (cudaStream_t)stream, true);
    if (final_norm != nullptr) {
        if (s->aux_final_norm == nullptr) TRY(dev_alloc(s->allocs, &s->aux_final_norm, H));
        launch_convert_to_bf16(final_norm, dtype, s->aux_final_norm, (int64_t)H, (cudaStream_t)stream);
    }
    CU_TRY(cudaGetLastError());
    CU_TRY(cudaStreamSynchronize((cudaStream_t)stream));
    s->aux_split = split;
    return 0;
}
extern "C" B200_API int b200_shard_token_from_partial(b200_shard_t* s, const void* msg_in, int B, int64_t* tok_out, float* margin_out,
                                                      void* stream) {
    if (s == nullptr || msg_in == nullptr || tok_out == nullptr) return fail(B200_EINVAL, "[ERROR] null argument");
    if (s->aux_lm_head == nullptr) return fail(B200_ESTATE, "[ERROR] this shard holds no vocabulary tail (b200_shard_set_vocab_tail)");
    if (B < 1 || B > s->cfg.max_batch) return fail(B200_EINVAL, "[ERROR] bad row count %d", B);
    CU_TRY(cudaSetDevice(s->device));
    cudaStream_t st = (cudaStream_t)stream;
    const int H = s->cfg.hidden_size, rows = s->cfg.vocab_size - s->aux_split;
    const float* mx = reinterpret_cast<const float*>(msg_in);
    const float* best = mx + (size_t)B * H;
    const float* second = best + B;
    const int64_t* idx = reinterpret_cast<const int64_t*>(second + B);
    TRY(score_rows(s, s->aux_lm_head, rows, s->aux_final_norm, mx, H, B, st));
    launch_argmax_range(s->logits, B, rows, s->aux_split, best, second, idx, nullptr, nullptr, tok_out, margin_out, st);
    s->n_launch += 1;
    s->n_bytes += (uint64_t)rows * H * 2;
    s->n_flops += 2ull * rows * H * B;
    CU_TRY(cudaGetLastError());
    return 0;
}

extern "C" B200_API int b200_shard_run_kernel(b200_shard_t* s, b200_kv_t* kv, int kind, int layer, int rows, void* stream) {
    if (s == nullptr || kv == nullptr) return fail(B200_EINVAL, "[ERROR] null argument");
    if (layer < 0 || layer >= s->cfg.n_layers || rows < 1 || rows > gemv_row_limit(s)) return fail(B200_EINVAL, "[ERROR] bad layer / rows");
    CU_TRY(cudaSetDevice(s->device));
    cudaStream_t st = (cudaStream_t)stream;
    const b200_shard_config_t& c = s->cfg;
    const LayerW& L = s->layers[layer];
    const int H = c.hidden_size, I = c.intermediate_size;
    GemvParams g{};
    switch (kind) {
        case 0:
            g.W = L.wqkv; g.x = s->x; g.norm_w = L.norm1; g.eps = c.rms_eps; g.N = s->qkv_rows; g.K = H;
            g.q_out = s->q_f; g.k_pool = kv->k_pool + (int64_t)layer * kv->layer_stride; g.v_pool = kv->v_pool + (int64_t)layer * kv->layer_stride;
            g.kv_off = s->tm.kv_off; g.cos = s->tm.cos; g.sin = s->tm.sin; g.n_heads = c.n_heads; g.n_kv = c.n_kv_heads;
            g.head_stride = (int64_t)kv->dev.page_size * 128;
            launch_gemv(g, rows, EPI_QKV, true, s->num_sms, st);
            break;
        case 1: {
            AttnDecodeParams a{};
            a.q = s->q_f; a.k_pool = kv->k_pool + (int64_t)layer * kv->layer_stride; a.v_pool = kv->v_pool + (int64_t)layer * kv->layer_stride;
            a.block_table = kv->dev.block_table; a.t_end = s->tm.t_end; a.call_base = s->tm.call_base;
            for (int i = 0; i < rows; ++i) a.slots.v[i] = i;
            a.S = 1; a.n_heads = c.n_heads; a.n_kv = c.n_kv_heads; a.page_size = kv->dev.page_size;
            a.max_pages_per_slot = kv->dev.max_pages_per_slot;
            a.n_splits = attn_splits(s, kv, rows);
            a.part = s->attn_part; a.out = s->attn_f;
            a.tickets = reinterpret_cast<unsigned int*>(s->attn_part + (size_t)kMaxGemvRows * c.n_heads * kMaxSplits * 130);
            launch_attn_decode(a, rows, st);
            break;
        }
        case 5:
            if (s->lm_head == nullptr) return fail(B200_ESTATE, "[ERROR] this shard does not hold the lm_head");
            g.W = s->lm_head; g.N = c.vocab_size; g.K = H; g.out = s->logits; g.x = s->x; g.x_stride = H;
            g.norm_w = c.has_final_norm ? s->final_norm : nullptr; g.eps = c.rms_eps;
            launch_gemv(g, rows, EPI_STORE, c.has_final_norm != 0, s->num_sms, st);
            break;
        case 2:
            g.W = L.wo; g.x = s->attn_f; g.N = H; g.K = s->nq; g.out = s->x;
            launch_gemv(g, rows, EPI_RESID, false, s->num_sms, st);
            break;
        case 3:
            g.W = L.wgu; g.x = s->x; g.norm_w = L.norm2; g.eps = c.rms_eps; g.N = 2 * I; g.K = H; g.out = s->act_f;
            launch_gemv(g, rows, EPI_SWIGLU, true, s->num_sms, st);
            break;
        case 4:
            g.W = L.wdown; g.x = s->act_f; g.N = H; g.K = I; g.out = s->x;
            launch_gemv(g, rows, EPI_RESID, false, s->num_sms, st);
            break;
        default:
            return fail(B200_EINVAL, "[ERROR] unknown kernel kind %d", kind);
    }
    CU_TRY(cudaGetLastError());
    s->n_launch += 1;
    return 0;
}

extern "C" B200_API int b200_shard_lm_head_argmax(b200_shard_t* s, const void* hidden, int B, int S, int64_t* tok, float* margin,
                                         void* stream) {
    if (s == nullptr || hidden == nullptr || tok == nullptr) return fail(B200_EINVAL, "[ERROR] null argument");
    if (!s->cfg.has_lm_head || s->lm_head == nullptr) return fail(B200_ESTATE, "[ERROR] this shard does not hold the lm_head");
    if (B < 1 || B > s->cfg.max_batch || S < 1) return fail(B200_EINVAL, "[ERROR] bad batch / seq_len");
    CU_TRY(cudaSetDevice(s->device));
    cudaStream_t st = (cudaStream_t)stream;
    launch_lastrow_to_f32(hidden, s->io_dtype, B, S, s->cfg.hidden_size, s->lastrow, st);
    s->n_launch += 1;
    TRY(enqueue_tail(s, B, S, false, tok, margin, st));
    CU_TRY(cudaGetLastError());
    s->n_bytes += (uint64_t)s->cfg.vocab_size * s->cfg.hidden_size * 2;
    return 0;
}
extern "C" B200_API int b200_shard_embed(b200_shard_t* s, const int64_t* ids, int n, void* out, void* stream) {
    if (s == nullptr || ids == nullptr || out == nullptr) return fail(B200_EINVAL, "[ERROR] null argument");
    if (!s->cfg.has_embedding || s->embedding == nullptr)
        return fail(B200_ESTATE, "[ERROR] this node does not have embedding layer while receiving user request.");
    CU_TRY(cudaSetDevice(s->device));
    launch_embed(s->embedding, ids, n, s->cfg.hidden_size, s->cfg.vocab_size, out, s->io_dtype, (cudaStream_t)stream);
    CU_TRY(cudaGetLastError());
    s->n_launch += 1;
    return 0;
}

extern "C" B200_API int b200_shard_counters(const b200_shard_t* s, uint64_t h_out[4]) {
    if (s == nullptr || h_out == nullptr) return fail(B200_EINVAL, "[ERROR] null argument");
    h_out[0] = s->n_launch; h_out[1] = s->n_bytes; h_out[2] = s->n_flops; h_out[3] = s->n_calls;
    return 0;
}
extern "C" B200_API int b200_shard_reset_counters(b200_shard_t* s) {
    if (s == nullptr) return fail(B200_EINVAL, "[ERROR] null shard");
    s->n_launch = s->n_bytes = s->n_flops = s->n_calls = 0;
    return 0;
}
extern "C" B200_API int b200_shard_set_rope_inv_freq(b200_shard_t* s, const float* h_inv_freq, int n) {
    if (s == nullptr || h_inv_freq == nullptr) return fail(B200_EINVAL, "[ERROR] null argument");
    if (n != 64) return fail(B200_EINVAL, "[ERROR] inv_freq must have head_dim/2 = 64 entries (got %d)", n);
    for (int j = 0; j < 64; ++j) s->inv.v[j] = h_inv_freq[j];
    drop_graphs(s, 0);  // captured graphs hold the old table by value
    return 0;
}
extern "C" B200_API int b200_shard_set_io_dtype(b200_shard_t* s, int dtype) {
    if (s == nullptr) return fail(B200_EINVAL, "[ERROR] null shard");
    if (dtype != B200_DT_BF16 && dtype != B200_DT_F32 && dtype != B200_DT_F16) return fail(B200_EINVAL, "[ERROR] unsupported message dtype %d", dtype);
    if (dtype != s->io_dtype) drop_graphs(s, 0);  // captured graphs have the conversion kernels of the old dtype baked in
    s->io_dtype = dtype;
    return 0;
}
extern "C" B200_API int b200_shard_set_graphs(b200_shard_t* s, int enabled) {
    if (s == nullptr) return fail(B200_EINVAL, "[ERROR] null shard");
    s->graphs = enabled != 0;
    return 0;
}

// ------------------------------------------------------------------------------------- hop
struct b200_hop {
    int device = 0, n_slots = 0;
    int64_t slot_bytes = 0;
    char* local = nullptr;       // [n_slots][slot_bytes] then flags[n_slots] (uint64)
    char* peer = nullptr;        // mapped mailbox of the next node
    bool peer_ipc = false;
    unsigned int* counter = nullptr;
    int* err = nullptr;
    int n_slots_peer = 0;
    int64_t slot_bytes_peer = 0;
    cudaStream_t poll_stream = nullptr;  // host polls of the flag / header must not queue behind the compute stream
    uint64_t* h_poll = nullptr;          // pinned: [0] flag, then the 256-byte header
};
struct HopHandle {  // exactly B200_HOP_HANDLE_BYTES
    cudaIpcMemHandle_t ipc;  // 64 bytes, valid across processes
    char* raw;               // valid inside the exporting process
    uint64_t geom;           // bit 0: ipc handle valid; bits 8..31: ring slots; bits 32..63: slot bytes / 256
};
static uint64_t hop_pack_geom(bool has_ipc, int n_slots, int64_t slot_bytes) {
    return (has_ipc ? 1ull : 0ull) | ((uint64_t)(n_slots & 0xffffff) << 8) | ((uint64_t)(slot_bytes / 256) << 32);
}
static_assert(sizeof(HopHandle) == B200_HOP_HANDLE_BYTES, "handle size");

extern "C" B200_API int b200_hop_create(int device, int n_ring_slots, int64_t slot_bytes, b200_hop_t** out) {
    if (out == nullptr || n_ring_slots < 1 || n_ring_slots > 0xffffff || slot_bytes < 16 || slot_bytes > (1ll << 39))
        return fail(B200_EINVAL, "[ERROR] bad hop arguments");
    CU_TRY(cudaSetDevice(device));
    b200_hop* h = new b200_hop();
    h->device = device;
    h->n_slots = n_ring_slots;
    h->slot_bytes = (slot_bytes + 255) / 256 * 256;
    const size_t total = (size_t)h->n_slots * h->slot_bytes + (size_t)h->n_slots * 8 + 256;
    cudaError_t e = cudaMalloc(&h->local, total);
    if (e != cudaSuccess) { cudaGetLastError(); delete h; return fail(B200_ENOMEM, "[ERROR] CUDA out of memory for the hop mailbox"); }
    e = cudaMalloc(&h->counter, 256);
    if (e != cudaSuccess) { cudaGetLastError(); cudaFree(h->local); delete h; return fail(B200_ENOMEM, "[ERROR] CUDA out of memory for the hop counters"); }
    e = cudaMemset(h->local, 0, total);
    if (e == cudaSuccess) e = cudaMemset(h->counter, 0, 256);
    if (e != cudaSuccess) { cudaGetLastError(); cudaFree(h->local); cudaFree(h->counter); delete h; return fail(B200_ECUDA, "[ERROR] CUDA: %s (hop mailbox init)", cudaGetErrorString(e)); }
    h->err = reinterpret_cast<int*>(reinterpret_cast<char*>(h->counter) + 128);
    CU_TRY(cudaDeviceSynchronize());
    *out = h;
    return 0;
}
extern "C" B200_API int b200_hop_export(b200_hop_t* h, void* handle_out) {
    if (h == nullptr || handle_out == nullptr) return fail(B200_EINVAL, "[ERROR] null argument");
    CU_TRY(cudaSetDevice(h->device));
    HopHandle hh;
    memset(&hh, 0, sizeof(hh));
    cudaError_t e = cudaIpcGetMemHandle(&hh.ipc, h->local);
    hh.geom = hop_pack_geom(e == cudaSuccess, h->n_slots, h->slot_bytes);
    if (e != cudaSuccess) cudaGetLastError();  // still usable inside one process through `raw`
    hh.raw = h->local;
    memcpy(handle_out, &hh, sizeof(hh));
    return 0;
}
extern "C" B200_API int b200_hop_connect(b200_hop_t* h, const void* peer_handle, int peer_device, int same_process) {
    if (h == nullptr || peer_handle == nullptr) return fail(B200_EINVAL, "[ERROR] null argument");
    CU_TRY(cudaSetDevice(h->device));
    if (h->peer != nullptr && h->peer_ipc) { cudaIpcCloseMemHandle(h->peer); }
    h->peer = nullptr;
    {   // the peer's ring geometry travels in the handle: slot and flag addresses are computed from ITS sizes
        HopHandle g;
        memcpy(&g, peer_handle, sizeof(g));
        const int pn = (int)((g.geom >> 8) & 0xffffff);
        const int64_t pb = (int64_t)(g.geom >> 32) * 256;
        if (pn < 1 || pb < 256) return fail(B200_EINVAL, "[ERROR] malformed hop handle (ring geometry missing)");
        h->n_slots_peer = pn;
        h->slot_bytes_peer = pb;
    }
    if (same_process) {
        HopHandle hh;
        memcpy(&hh, peer_handle, sizeof(hh));
        char* raw = hh.raw;
        if (peer_device != h->device) {
            int can = 0;
            CU_TRY(cudaDeviceCanAccessPeer(&can, h->device, peer_device));
            if (!can) return fail(B200_ECUDA, "[ERROR] GPU %d cannot map GPU %d's memory", h->device, peer_device);
            cudaError_t e = cudaDeviceEnablePeerAccess(peer_device, 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) CU_TRY(e);
            cudaGetLastError();
        }
        h->peer = raw;
        h->peer_ipc = false;
    } else {
        HopHandle hh;
        memcpy(&hh, peer_handle, sizeof(hh));
        if (!(hh.geom & 1ull)) return fail(B200_ECUDA, "[ERROR] the peer could not export a CUDA IPC handle");
        void* p = nullptr;
        CU_TRY(cudaIpcOpenMemHandle(&p, hh.ipc, cudaIpcMemLazyEnablePeerAccess));
        h->peer = reinterpret_cast<char*>(p);
        h->peer_ipc = true;
    }
    return 0;
}
extern "C" B200_API int b200_hop_send(b200_hop_t* h, const void* payload, int64_t bytes, uint64_t seq, void* stream) {
    if (h == nullptr || h->peer == nullptr) return fail(B200_ESTATE, "[ERROR] hop has no connected peer");
    if (bytes < 0 || bytes > h->slot_bytes_peer) return fail(B200_EINVAL, "[ERROR] hop payload of %lld bytes exceeds the peer's slot (%lld)", (long long)bytes, (long long)h->slot_bytes_peer);
    CU_TRY(cudaSetDevice(h->device));
    const int slot = (int)(seq % (uint64_t)h->n_slots_peer);
    char* pslot = h->peer + (size_t)slot * h->slot_bytes_peer;
    uint64_t* pflag = reinterpret_cast<uint64_t*>(h->peer + (size_t)h->n_slots_peer * h->slot_bytes_peer) + slot;
    launch_hop_send(payload, bytes, pslot, pflag, seq, h->counter, (cudaStream_t)stream);
    CU_TRY(cudaGetLastError());
    return 0;
}
extern "C" B200_API void* b200_hop_peer_slot(b200_hop_t* h, uint64_t seq) {
    if (h == nullptr || h->peer == nullptr) return nullptr;
    return h->peer + (size_t)(seq % (uint64_t)h->n_slots_peer) * h->slot_bytes_peer;
}
extern "C" B200_API void* b200_hop_local_slot(b200_hop_t* h, uint64_t seq) {
    if (h == nullptr) return nullptr;
    return h->local + (size_t)(seq % (uint64_t)h->n_slots) * h->slot_bytes;
}
extern "C" B200_API int b200_hop_recv(b200_hop_t* h, void* out, int64_t bytes, uint64_t seq, int64_t max_wait_us, void* stream) {
    if (h == nullptr) return fail(B200_EINVAL, "[ERROR] null argument");
    if (out == nullptr) bytes = 0;  // wait only: the consumer reads the mailbox slot in place
    if (bytes < 0 || bytes > h->slot_bytes) return fail(B200_EINVAL, "[ERROR] hop payload exceeds the slot");
    CU_TRY(cudaSetDevice(h->device));
    const int slot = (int)(seq % (uint64_t)h->n_slots);
    const char* lslot = h->local + (size_t)slot * h->slot_bytes;
    const uint64_t* lflag = reinterpret_cast<const uint64_t*>(h->local + (size_t)h->n_slots * h->slot_bytes) + slot;
    launch_hop_recv(out, bytes, lslot, lflag, seq, max_wait_us * 2000 /* ~2 GHz */, h->err, (cudaStream_t)stream);
    CU_TRY(cudaGetLastError());
    return 0;
}
extern "C" B200_API int b200_hop_send_msg(b200_hop_t* h, const void* h_header, int header_bytes, const void* d_payload,
                                          int64_t bytes, uint64_t seq, void* stream) {
    if (h == nullptr || h->peer == nullptr) return fail(B200_ESTATE, "[ERROR] hop has no connected peer");
    if (header_bytes < 0 || header_bytes > kHopHeaderBytes || (header_bytes > 0 && h_header == nullptr))
        return fail(B200_EINVAL, "[ERROR] hop message header must be at most %d bytes", kHopHeaderBytes);
    if (bytes < 0 || bytes + kHopHeaderBytes > h->slot_bytes_peer || (bytes > 0 && d_payload == nullptr))
        return fail(B200_EINVAL, "[ERROR] hop message of %lld + %d bytes exceeds the peer's slot (%lld)", (long long)bytes,
                    kHopHeaderBytes, (long long)h->slot_bytes_peer);
    CU_TRY(cudaSetDevice(h->device));
    HopHeader hdr;
    memset(hdr.b, 0, sizeof(hdr.b));
    if (header_bytes > 0) memcpy(hdr.b, h_header, (size_t)header_bytes);
    const int slot = (int)(seq % (uint64_t)h->n_slots_peer);
    char* pslot = h->peer + (size_t)slot * h->slot_bytes_peer;
    uint64_t* pflag = reinterpret_cast<uint64_t*>(h->peer + (size_t)h->n_slots_peer * h->slot_bytes_peer) + slot;
    launch_hop_send_msg(hdr, d_payload, bytes, pslot, pflag, seq, h->counter, (cudaStream_t)stream);
    CU_TRY(cudaGetLastError());
    return 0;
}
extern "C" B200_API int b200_hop_poll(b200_hop_t* h, uint64_t seq, void* h_header_out) {
    if (h == nullptr) return fail(B200_EINVAL, "[ERROR] null argument");
    CU_TRY(cudaSetDevice(h->device));
    if (h->poll_stream == nullptr) {
        CU_TRY(cudaStreamCreateWithFlags(&h->poll_stream, cudaStreamNonBlocking));
        CU_TRY(cudaHostAlloc(reinterpret_cast<void**>(&h->h_poll), 8 + kHopHeaderBytes, cudaHostAllocDefault));
    }
    const int slot = (int)(seq % (uint64_t)h->n_slots);
    const uint64_t* lflag = reinterpret_cast<const uint64_t*>(h->local + (size_t)h->n_slots * h->slot_bytes) + slot;
    CU_TRY(cudaMemcpyAsync(h->h_poll, lflag, 8, cudaMemcpyDeviceToHost, h->poll_stream));
    CU_TRY(cudaStreamSynchronize(h->poll_stream));
    if (h->h_poll[0] < seq + 1) return 0;  // not there yet
    if (h_header_out != nullptr) {
        // the flag was released after the header and the payload: a copy ordered after the flag read sees both
        CU_TRY(cudaMemcpyAsync(h->h_poll + 1, h->local + (size_t)slot * h->slot_bytes, kHopHeaderBytes, cudaMemcpyDeviceToHost, h->poll_stream));
        CU_TRY(cudaStreamSynchronize(h->poll_stream));
        memcpy(h_header_out, h->h_poll + 1, kHopHeaderBytes);
    }
    return 1;
}
extern "C" B200_API int b200_hop_error(b200_hop_t* h) {
    if (h == nullptr) return 0;
    int v = 0;
    cudaSetDevice(h->device);
    cudaMemcpy(&v, h->err, sizeof(int), cudaMemcpyDeviceToHost);
    return v;
}
extern "C" B200_API int b200_hop_destroy(b200_hop_t* h) {
    if (h == nullptr) return 0;
    cudaSetDevice(h->device);
    cudaDeviceSynchronize();
    if (h->peer != nullptr && h->peer_ipc) cudaIpcCloseMemHandle(h->peer);
    if (h->poll_stream != nullptr) cudaStreamDestroy(h->poll_stream);
    if (h->h_poll != nullptr) cudaFreeHost(h->h_poll);
    cudaFree(h->local);
    cudaFree(h->counter);
    delete h;
    return 0;
}

// ------------------------------------------------------------------------------------- single kernels
extern "C" B200_API int b200_rmsnorm(const float* x, const void* w, float eps, int rows, int n, void* out, void* stream) {
    if (x == nullptr || out == nullptr || rows < 1 || n < 8 || n % 8 != 0) return fail(B200_EINVAL, "[ERROR] bad rmsnorm arguments");
    launch_rmsnorm_bf16(x, reinterpret_cast<const __nv_bfloat16*>(w), eps, rows, n, reinterpret_cast<__nv_bfloat16*>(out), (cudaStream_t)stream);
    CU_TRY(cudaGetLastError());
    return 0;
}
extern "C" B200_API int b200_attn_decode(const float* q, const void* k_pool, const void* v_pool, const int32_t* block_table,
                                         const int32_t* t_end, int rows, int n_heads, int n_kv_heads, int page_size,
                                         int max_pages_per_row, int n_splits, float* scratch, float* out, void* stream) {
    if (q == nullptr || k_pool == nullptr || v_pool == nullptr || block_table == nullptr || t_end == nullptr ||
        scratch == nullptr || out == nullptr)
        return fail(B200_EINVAL, "[ERROR] null argument");
    if (rows < 1 || rows > kMaxBatch || n_heads < 1 || n_kv_heads < 1 || n_heads % n_kv_heads != 0 || page_size < 64 ||
        page_size % 64 != 0 || max_pages_per_row < 1 || n_splits < 1 || n_splits > kMaxSplits ||
        ((int64_t)max_pages_per_row * page_size + n_splits - 1) / n_splits > kMaxChunkKeys)
        return fail(B200_EINVAL, "[ERROR] attn_decode: need 1 <= rows <= %d, n_heads %% n_kv_heads == 0, page_size %% 64 == 0, "
                    "1 <= n_splits <= %d and at most %d keys per split", kMaxBatch, kMaxSplits, kMaxChunkKeys);
    AttnDecodeParams a{};
    a.q = q; a.k_pool = reinterpret_cast<const __nv_bfloat16*>(k_pool); a.v_pool = reinterpret_cast<const __nv_bfloat16*>(v_pool);
    a.block_table = block_table; a.t_end = t_end;
    for (int i = 0; i < rows; ++i) a.slots.v[i] = i;
    a.S = 1; a.n_heads = n_heads; a.n_kv = n_kv_heads; a.page_size = page_size; a.max_pages_per_slot = max_pages_per_row;
    a.n_splits = n_splits; a.part = scratch; a.out = out;
    CU_TRY(cudaMemsetAsync(scratch + (size_t)rows * n_heads * n_splits * 130, 0, (size_t)rows * n_heads * sizeof(unsigned int),
                           (cudaStream_t)stream));
    launch_attn_decode(a, rows, (cudaStream_t)stream);
    CU_TRY(cudaGetLastError());
    return 0;
}
extern "C" B200_API int b200_attn_decode_tc(const float* q, const void* k_pool, const void* v_pool, const int32_t* block_table,
                                            const int32_t* t_end, int rows, int n_heads, int n_kv_heads, int page_size,
                                            int max_pages_per_row, int n_pool_pages, int max_keys, float* out, void* stream) {
    if (q == nullptr || k_pool == nullptr || v_pool == nullptr || block_table == nullptr || t_end == nullptr || out == nullptr)
        return fail(B200_EINVAL, "[ERROR] null argument");
    if (rows < 1 || rows > kMaxBatch || n_heads < 1 || n_kv_heads < 1 || n_heads % n_kv_heads != 0 || page_size < 64 ||
        page_size % 64 != 0 || max_pages_per_row < 1 || n_pool_pages < 1 || max_keys < 1)
        return fail(B200_EINVAL, "[ERROR] attn_decode_tc: need 1 <= rows <= %d, n_heads %% n_kv_heads == 0, page_size %% 64 == 0", kMaxBatch);
    AttnDecodeParams a{};
    a.q = q; a.k_pool = reinterpret_cast<const __nv_bfloat16*>(k_pool); a.v_pool = reinterpret_cast<const __nv_bfloat16*>(v_pool);
    a.block_table = block_table; a.t_end = t_end;
    for (int i = 0; i < rows; ++i) a.slots.v[i] = i;
    a.S = 1; a.n_heads = n_heads; a.n_kv = n_kv_heads; a.page_size = page_size; a.max_pages_per_slot = max_pages_per_row;
    a.n_splits = 1; a.out = out;
    const int rc = launch_attn_decode_tc(a, rows, max_keys, (int64_t)n_pool_pages * n_kv_heads * page_size, (cudaStream_t)stream);
    if (rc == 1) return fail(B200_EINVAL, "[ERROR] attn_decode_tc: more than 8 query heads per K/V head, or a context too long for its score buffer");
    if (rc < 0) return fail(B200_ECUDA, "[ERROR] %s", gemm_error());
    CU_TRY(cudaGetLastError());
    return 0;
}
extern "C" B200_API int b200_attn_prefill(const void* q_hi, const void* q_lo, const void* k_pool, const void* v_pool,
                                          const int32_t* block_table, const int32_t* base_len, int B, int S, int n_heads,
                                          int n_kv_heads, int page_size, int max_pages_per_row, int n_pool_pages, int causal,
                                          int out_f16, void* out, void* stream) {
    if (q_hi == nullptr || q_lo == nullptr || k_pool == nullptr || v_pool == nullptr || block_table == nullptr ||
        base_len == nullptr || out == nullptr)
        return fail(B200_EINVAL, "[ERROR] null argument");
    if (B < 1 || B > kMaxBatch || S < 1 || n_heads < 1 || n_kv_heads < 1 || n_heads % n_kv_heads != 0 || page_size < 64 ||
        page_size % 64 != 0 || max_pages_per_row < 1 || n_pool_pages < 1)
        return fail(B200_EINVAL, "[ERROR] attn_prefill: need 1 <= B <= %d, S >= 1, n_heads %% n_kv_heads == 0, page_size %% 64 == 0", kMaxBatch);
    AttnPrefillParams a{};
    a.q = reinterpret_cast<const __nv_bfloat16*>(q_hi); a.q_lo = reinterpret_cast<const __nv_bfloat16*>(q_lo);
    a.out_f16 = out_f16 ? 1 : 0;
    a.k_pool = reinterpret_cast<const __nv_bfloat16*>(k_pool); a.v_pool = reinterpret_cast<const __nv_bfloat16*>(v_pool);
    a.block_table = block_table; a.call_base = base_len;
    for (int i = 0; i < B; ++i) a.slots.v[i] = i;
    a.B = B; a.S = S; a.causal = causal ? 1 : 0; a.n_heads = n_heads; a.n_kv = n_kv_heads; a.page_size = page_size;
    a.max_pages_per_slot = max_pages_per_row; a.out = out;
    if (launch_attn_prefill_tc(a, (int64_t)n_pool_pages * n_kv_heads * page_size, (cudaStream_t)stream) != 0)
        return fail(B200_ECUDA, "[ERROR] %s", gemm_error());
    CU_TRY(cudaGetLastError());
    return 0;
}
extern "C" B200_API int b200_gemm_bf16(const void* a, const void* w, int M, int N, int K, float* out, void* stream) {
    if (a == nullptr || w == nullptr || out == nullptr) return fail(B200_EINVAL, "[ERROR] null argument");
    GemmParams g{};
    g.A = a; g.W = w; g.M = M; g.N = N; g.K = K; g.out = out; g.epi = GEMM_STORE_F32;
    if (launch_gemm(g, (cudaStream_t)stream) != 0) return fail(B200_EINVAL, "[ERROR] %s", gemm_error());
    CU_TRY(cudaGetLastError());
    return 0;
}
extern "C" B200_API int b200_gemm_f16(const void* a, const void* w, int M, int N, int K, float* out, void* stream) {
    if (a == nullptr || w == nullptr || out == nullptr) return fail(B200_EINVAL, "[ERROR] null argument");
    GemmParams g{};
    g.A = a; g.W = reinterpret_cast<const __nv_bfloat16*>(w); g.M = M; g.N = N; g.K = K; g.out = out; g.epi = GEMM_STORE_F32;
    g.ab_f16 = 1;
    if (launch_gemm(g, (cudaStream_t)stream) != 0) return fail(B200_EINVAL, "[ERROR] %s", gemm_error());
    CU_TRY(cudaGetLastError());
    return 0;
}
extern "C" B200_API int b200_gemv_f16(const float* x, const void* w, int M, int N, int K, float* out, void* stream) {
    if (x == nullptr || w == nullptr || out == nullptr) return fail(B200_EINVAL, "[ERROR] null argument");
    if (M < 1 || M > kMaxGemvRows || K % 8 != 0 || K < 8 || gemv_smem_bytes(M, K) > 220 * 1024)
        return fail(B200_EINVAL, "[ERROR] gemv: need 1 <= M <= %d, K %% 8 == 0 and the token rows must fit in shared memory", kMaxGemvRows);
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    GemvParams g{};
    g.W = w; g.x = x; g.N = N; g.K = K; g.out = out;
    launch_gemv(g, M, EPI_STORE, false, sms, (cudaStream_t)stream);
    CU_TRY(cudaGetLastError());
    return 0;
}

extern "C" B200_API int b200_gemm_stream_f16(const void* x, const void* w, int M, int N, int K, int split, float* out, void* stream) {
    if (x == nullptr || w == nullptr || out == nullptr) return fail(B200_EINVAL, "[ERROR] null argument");
    if (split != 0 && split != 1 && split != 2 && split != 4 && split != 8) return fail(B200_EINVAL, "[ERROR] gemm_stream: split must be 0, 1, 2, 4 or 8");
    if (split > 1 && (K + 63) / 64 < split) return fail(B200_EINVAL, "[ERROR] gemm_stream: K too short for this split");
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    StreamGemmParams g{};
    g.W = w; g.X = x; g.M = M; g.N = N; g.K = K; g.split = split; g.out = out;
    if (launch_gemm_stream(g, EPI_STORE, sms, (cudaStream_t)stream) != 0) return fail(B200_EINVAL, "[ERROR] %s", gemm_stream_last_error());
    CU_TRY(cudaGetLastError());
    return 0;
}
