"""TEST INFRASTRUCTURE — CPU restatement of the reference's shard-execution arithmetic.

The reference (``/root/reference``) contains no arithmetic of its own on this path: it
delegates to Hugging Face ``transformers`` (third-party, pinned ``transformers==4.53.2``,
``/root/reference/requirements.txt:8``; 5.5.0 is what this image has), called from
``utils/shard_loader.py:36-38,50,66-77`` and ``utils/node_worker.py:152,184,241,262-264``.
This file restates that published algorithm op for op in plain torch-CPU tensor ops (no
``transformers`` import), in the same order and with the same dtype casts, so that in
fp32 on one machine it reproduces the live reference bit for bit.  SP: below means
``site-packages/transformers/models/llama/modeling_llama.py`` (5.5.0).

Pinned by: ``tests/golden/*.npz`` — outputs of the reference's own ``NodeWorker`` chain
run in the build container by ``oracle/make_golden.py`` (the reference ships no tests or
vectors of its own).  ``tests/test_oracle_golden.py`` checks this file
against them.

Behaviour that is deliberately reproduced: the reference passes no
attention mask, so prefill attention is NON-causal.
"""
from __future__ import annotations

import torch
import torch.nn.functional as F


# --------------------------------------------------------------------------- math
def rms_norm(x: torch.Tensor, weight: torch.Tensor, eps: float) -> torch.Tensor:
    """LlamaRMSNorm.forward, SP:62-67: fp32 mean-square, rsqrt, cast back, times weight."""
    dt = x.dtype
    h = x.to(torch.float32)
    var = h.pow(2).mean(-1, keepdim=True)
    h = h * torch.rsqrt(var + eps)
    return weight * h.to(dt)


def rope_inv_freq(head_dim: int, theta: float, scaling: dict | None = None) -> torch.Tensor:
    """compute_default_rope_parameters, SP:93-120; with ``scaling`` (rope_type "llama3": factor, low_freq_factor,
    high_freq_factor, original_max_position_embeddings) the Llama-3.x rescaling of
    site-packages/transformers/modeling_rope_utils.py:550-626 (_compute_llama3_parameters), op for op."""
    import math
    inv_freq = 1.0 / (theta ** (torch.arange(0, head_dim, 2, dtype=torch.int64).to(torch.float) / head_dim))
    if not scaling or scaling.get("rope_type", "default") == "default":
        return inv_freq
    if scaling["rope_type"] != "llama3":
        raise ValueError(f"[ERROR] unsupported rope_type {scaling['rope_type']}")
    factor, low, high = scaling["factor"], scaling["low_freq_factor"], scaling["high_freq_factor"]
    old_len = scaling["original_max_position_embeddings"]
    low_wavelen, high_wavelen = old_len / low, old_len / high
    wavelen = 2 * math.pi / inv_freq
    inv_llama = torch.where(wavelen > low_wavelen, inv_freq / factor, inv_freq)
    smooth = (old_len / wavelen - low) / (high - low)
    smoothed = (1 - smooth) * inv_llama / factor + smooth * inv_llama
    is_medium = ~(wavelen < high_wavelen) * ~(wavelen > low_wavelen)
    return torch.where(is_medium, smoothed, inv_llama)


def rope_cos_sin(position_ids: torch.Tensor, head_dim: int, theta: float, dtype: torch.dtype, scaling: dict | None = None):
    """LlamaRotaryEmbedding.forward, SP:124-135 (called utils/node_worker.py:241): fp32
    ``inv_freq (x) pos``, ``cat(freqs, freqs)``, cos/sin, cast to the activation dtype."""
    inv = rope_inv_freq(head_dim, theta, scaling)[None, :, None].float().expand(position_ids.shape[0], -1, 1)
    pos = position_ids[:, None, :].float()
    freqs = (inv @ pos).transpose(1, 2)
    emb = torch.cat((freqs, freqs), dim=-1)
    return emb.cos().to(dtype), emb.sin().to(dtype)


def rotate_half(x: torch.Tensor) -> torch.Tensor:
    """SP:138-142 (half-split convention)."""
    x1 = x[..., : x.shape[-1] // 2]
    x2 = x[..., x.shape[-1] // 2:]
    return torch.cat((-x2, x1), dim=-1)


def apply_rope(q, k, cos, sin):
    """apply_rotary_pos_emb, SP:146-168; cos/sin [B,S,d] broadcast over heads (dim 1)."""
    cos = cos.unsqueeze(1)
    sin = sin.unsqueeze(1)
    return (q * cos) + (rotate_half(q) * sin), (k * cos) + (rotate_half(k) * sin)


def build_position_ids(past_len: int, seq_len: int, batch_size: int) -> torch.Tensor:
    """utils/forwarding_utils.py:4-26."""
    pos = torch.arange(past_len, past_len + seq_len, dtype=torch.long)
    return pos.unsqueeze(0).expand(batch_size, -1).contiguous()


class KVCache:
    """DynamicCache / DynamicLayer.update, SP:transformers/cache_utils.py:102-121: per
    (shard-relative) layer, ``cat`` the new K,V on the sequence axis and return all of it."""

    def __init__(self):
        self.k: list[torch.Tensor] = []
        self.v: list[torch.Tensor] = []

    def get_seq_length(self) -> int:
        return 0 if not self.k else int(self.k[0].shape[-2])

    def update(self, k, v, layer_idx):
        if layer_idx == len(self.k):
            self.k.append(k)
            self.v.append(v)
        else:
            self.k[layer_idx] = torch.cat([self.k[layer_idx], k], dim=-2)
            self.v[layer_idx] = torch.cat([self.v[layer_idx], v], dim=-2)
        return self.k[layer_idx], self.v[layer_idx]


def attention(x, w, cos, sin, cache: KVCache, layer_idx: int, n_heads: int, n_kv: int, causal: bool = False):
    """LlamaAttention.forward SP:251-289 + eager_attention_forward SP:199-221.
    ``causal=False`` is the reference's behaviour (no mask is ever passed:
    utils/node_worker.py:253-258, utils/shard_loader.py:57,69)."""
    B, S, _ = x.shape
    d = w["self_attn.q_proj.weight"].shape[0] // n_heads
    q = F.linear(x, w["self_attn.q_proj.weight"]).view(B, S, n_heads, d).transpose(1, 2)
    k = F.linear(x, w["self_attn.k_proj.weight"]).view(B, S, n_kv, d).transpose(1, 2)
    v = F.linear(x, w["self_attn.v_proj.weight"]).view(B, S, n_kv, d).transpose(1, 2)
    q, k = apply_rope(q, k, cos, sin)
    k, v = cache.update(k, v, layer_idx)
    rep = n_heads // n_kv
    if rep > 1:  # repeat_kv, SP:187-196
        T = k.shape[2]
        k = k[:, :, None].expand(B, n_kv, rep, T, d).reshape(B, n_heads, T, d)
        v = v[:, :, None].expand(B, n_kv, rep, T, d).reshape(B, n_heads, T, d)
    scores = torch.matmul(q, k.transpose(2, 3)) * (d ** -0.5)
    if causal:
        T = k.shape[2]
        past = T - S
        i = torch.arange(S)[:, None] + past
        j = torch.arange(T)[None, :]
        scores = scores + torch.where(j <= i, 0.0, float("-inf")).to(scores.dtype)
    p = F.softmax(scores, dim=-1, dtype=torch.float32).to(q.dtype)
    o = torch.matmul(p, v).transpose(1, 2).contiguous().reshape(B, S, -1).contiguous()
    return F.linear(o, w["self_attn.o_proj.weight"])


def mlp(x, w):
    """LlamaMLP.forward, SP:182-184."""
    return F.linear(F.silu(F.linear(x, w["mlp.gate_proj.weight"])) * F.linear(x, w["mlp.up_proj.weight"]),
                    w["mlp.down_proj.weight"])


def decoder_layer(x, w, cos, sin, cache, layer_idx, n_heads, n_kv, eps, causal=False):
    """LlamaDecoderLayer.forward, SP:303-332."""
    res = x
    h = rms_norm(x, w["input_layernorm.weight"], eps)
    h = attention(h, w, cos, sin, cache, layer_idx, n_heads, n_kv, causal)
    x = res + h
    res = x
    h = rms_norm(x, w["post_attention_layernorm.weight"], eps)
    h = mlp(h, w)
    return res + h


# --------------------------------------------------------------------------- worker
class OracleShard:
    """LlamaShardPart, utils/shard_loader.py:8-78: layers with shard-relative ``layer_idx``
    (``:36-38``), optional final RMSNorm (``:49-55,76-77``)."""

    def __init__(self, layer_states: list[dict], n_heads: int, n_kv: int, eps: float,
                 final_norm_weight: torch.Tensor | None = None, dtype=torch.float32):
        self.layers = [{k: v.to(dtype) for k, v in st.items()} for st in layer_states]
        self.n_heads, self.n_kv, self.eps = n_heads, n_kv, eps
        self.final_norm_weight = None if final_norm_weight is None else final_norm_weight.to(dtype)

    def forward(self, hidden, cache: KVCache, cos, sin, causal=False):
        for rel_idx, w in enumerate(self.layers):
            hidden = decoder_layer(hidden, w, cos, sin, cache, rel_idx, self.n_heads, self.n_kv, self.eps, causal)
        if self.final_norm_weight is not None:
            hidden = rms_norm(hidden, self.final_norm_weight, self.eps)
        return hidden


class OracleWorker:
    """NodeWorker's compute methods (utils/node_worker.py:187-309) without sockets or
    tokenizer: same message dicts, same placement rules (embedding on entry nodes,
    RoPE on ``start==0``, final norm + lm_head on ``end==layer_num``)."""

    def __init__(self, cfg, seed: int, start: int, end: int, can_receive_user_request: bool,
                 dtype=torch.float32, causal: bool = False):
        from . import synth
        if start < 0 or start >= end or end > cfg.num_hidden_layers:
            raise ValueError("[ERROR] start or end is invalid")
        self.cfg, self.start, self.end, self.dtype, self.causal = cfg, start, end, dtype, causal
        self.can_receive_user_request = can_receive_user_request
        self.layer_num = cfg.num_hidden_layers
        self.embed = synth.embedding_state(cfg, seed)["weight"].to(dtype) if can_receive_user_request else None
        is_tail = end == self.layer_num
        self.lm_head = synth.lm_head_state(cfg, seed)["weight"].to(dtype) if is_tail else None
        self.shard = OracleShard(
            [synth.layer_state(cfg, seed, i) for i in range(start, end)],
            cfg.num_attention_heads, cfg.num_key_value_heads, cfg.rms_norm_eps,
            synth.final_norm_state(cfg, seed)["weight"] if is_tail else None, dtype)
        self.cache = KVCache()
        self.batch_size = 0
        self.generated_ids: list[torch.Tensor] = []
        self.last_logits = None  # [B,V] of the last position (for margin bookkeeping)

    @torch.inference_mode()
    def receive_user_request(self, input_ids: torch.Tensor) -> dict:
        """utils/node_worker.py:187-224 (``input_ids=`` branch, ``:205-209``)."""
        if not self.can_receive_user_request:
            raise RuntimeError("[ERROR] this node does not have embedding layer while receiving user request.")
        if input_ids.ndim != 2:
            raise ValueError("[ERROR] input_ids must be a 2D tensor with shape [batch_size, seq_len].")
        self.generated_ids = [input_ids]
        hidden = F.embedding(input_ids, self.embed)
        self.batch_size = hidden.shape[0]
        return {"hidden_states": hidden, "batch_size": hidden.shape[0], "seq_len": hidden.shape[1]}

    @torch.inference_mode()
    def pass_through_shard(self, state_info: dict):
        """utils/node_worker.py:226-272."""
        if isinstance(state_info, dict) and "batch_size" in state_info and "seq_len" in state_info:
            if self.start != 0:
                raise RuntimeError("[ERROR] after embedding layer, the states should first passing hidden layer 0!")
            self.batch_size = state_info["batch_size"]
            pos = build_position_ids(self.cache.get_seq_length(), state_info["seq_len"], state_info["batch_size"])
            cos, sin = rope_cos_sin(pos, self.cfg.head_dim, self.cfg.rope_theta, state_info["hidden_states"].dtype,
                                    getattr(self.cfg, "rope_scaling", None))
        elif isinstance(state_info, dict) and "cos" in state_info and "sin" in state_info:
            cos, sin = state_info["cos"], state_info["sin"]
        else:
            raise RuntimeError(f"[ERROR] Received unknown state_info type: {type(state_info)}")
        hidden = state_info["hidden_states"].to(self.dtype)
        out = self.shard.forward(hidden, self.cache, cos, sin, self.causal)
        if self.end == self.layer_num:
            logits = F.linear(out, self.lm_head)
            self.last_logits = logits[:, -1, :]
            return torch.argmax(logits[:, -1, :], dim=-1)
        return {"hidden_states": out, "cos": cos, "sin": sin}

    @torch.inference_mode()
    def receive_next_token(self, next_token_id: torch.Tensor, max_new_tokens: int = 1024):
        """utils/node_worker.py:274-309 without the tokenizer decode / EOS string test
        (random weights never emit EOS reliably)."""
        if not self.can_receive_user_request:
            raise RuntimeError('[ERROR] this node does not store the "generated token id" list, '
                               "but received a next_token_id.")
        self.generated_ids.append(next_token_id.unsqueeze(-1))
        if len(self.generated_ids) > max_new_tokens:
            return True, None
        hidden = F.embedding(next_token_id.unsqueeze(0), self.embed)
        return False, {"hidden_states": hidden, "batch_size": self.batch_size, "seq_len": 1}

    def clear_KV_cache(self):
        """utils/node_worker.py:319-355."""
        self.cache = KVCache()
        self.batch_size = 0
        self.generated_ids = []


def run_chain(workers: list[OracleWorker], input_ids: torch.Tensor, max_new_tokens: int, record: bool = True):
    """In-process chain, the pattern of utils/node_profiler.py:1174-1236 minus the sockets.
    Returns ``(token_ids [n_tokens,B], per_pass list of per-shard hidden states, margins)``."""
    head = workers[0]
    msg = head.receive_user_request(input_ids)
    tokens, states, margins = [], [], []
    while True:
        per_pass = []
        data = msg
        for w in workers:
            data = w.pass_through_shard(data)
            if isinstance(data, dict) and record:
                per_pass.append(data["hidden_states"].clone())
        states.append(per_pass)
        top2 = torch.topk(workers[-1].last_logits.float(), 2, dim=-1).values
        margins.append((top2[:, 0] - top2[:, 1]).clone())
        tokens.append(data.clone())
        # the reference's head appends [B] -> [B,1]; its own re-embed is batch-1 only
        # (utils/node_worker.py:302) — batch>1 is driven per row here.
        head.generated_ids.append(data.unsqueeze(-1))
        if len(head.generated_ids) > max_new_tokens:
            break
        hidden = F.embedding(data.unsqueeze(-1), head.embed)  # [B,1,H]
        msg = {"hidden_states": hidden, "batch_size": head.batch_size, "seq_len": 1}
    return torch.stack(tokens), states, torch.stack(margins)
