"""Thin object wrappers over the C ABI: ``Shard`` (weights + workspace of one layer range) and
``PagedKVCache`` (what ``NodeWorker.past_key_value`` holds instead of a ``DynamicCache``,
utils/node_worker.py:98,184).  Device memory handed in/out is plain torch tensors; torch is the
allocator and the stream provider, the kernels are the library's."""
from __future__ import annotations

import ctypes as C
import json
import os

import torch

from . import _lib

LAYER_KEYS = (  # LlamaDecoderLayer.state_dict() keys as saved in block_{i}.pth (utils/model_sharder.py:77-79)
    "self_attn.q_proj.weight", "self_attn.k_proj.weight", "self_attn.v_proj.weight", "self_attn.o_proj.weight",
    "mlp.gate_proj.weight", "mlp.up_proj.weight", "mlp.down_proj.weight",
    "input_layernorm.weight", "post_attention_layernorm.weight",
)


def rope_inv_freq(head_dim: int, theta: float, scaling: dict | None):
    """Inverse RoPE frequencies as fp32 torch ops, the way transformers computes them: default
    (modeling_llama.py:116-118) or the Llama-3.x "llama3" rescaling (modeling_rope_utils.py:606-624)."""
    import math
    inv = 1.0 / (theta ** (torch.arange(0, head_dim, 2, dtype=torch.int64).to(torch.float) / head_dim))
    if scaling:
        f, lo, hi, old = (scaling["factor"], scaling["low_freq_factor"], scaling["high_freq_factor"],
                          scaling["original_max_position_embeddings"])
        wavelen = 2 * math.pi / inv
        scaled = torch.where(wavelen > old / lo, inv / f, inv)
        smooth = (old / wavelen - lo) / (hi - lo)
        mid = ~(wavelen < old / hi) * ~(wavelen > old / lo)
        inv = torch.where(mid, (1 - smooth) * scaled / f + smooth * scaled, scaled)
    return inv.float().contiguous()


def read_model_config(shards_path: str) -> dict:
    """The fields of config.json the path needs (what LlamaConfig.from_pretrained(shards_path)
    supplies at utils/shard_loader.py:35 and utils/node_worker.py:88)."""
    with open(os.path.join(shards_path, "config.json")) as f:
        c = json.load(f)
    n_heads = c["num_attention_heads"]
    rope = c.get("rope_parameters") or {}
    scaling = c.get("rope_scaling") or ({} if rope.get("rope_type", "default") == "default" else rope)
    rope_type = (scaling or {}).get("rope_type", (scaling or {}).get("type", "default"))
    if rope_type not in ("default", "llama3"):
        raise ValueError(f"[ERROR] RoPE type {rope_type!r} is not built (default and llama3 are)")
    return {
        "rope_scaling": dict(scaling, rope_type=rope_type) if rope_type != "default" else None,
        "hidden_size": c["hidden_size"], "intermediate_size": c["intermediate_size"],
        "num_hidden_layers": c["num_hidden_layers"], "num_attention_heads": n_heads,
        "num_key_value_heads": c.get("num_key_value_heads") or n_heads,
        "head_dim": c.get("head_dim") or c["hidden_size"] // n_heads,
        "vocab_size": c["vocab_size"], "rms_norm_eps": c.get("rms_norm_eps", 1e-6),
        "rope_theta": float(rope.get("rope_theta", c.get("rope_theta", 10000.0))),
        "max_position_embeddings": c.get("max_position_embeddings", 4096),
    }


class Shard:
    def __init__(self, model_cfg: dict, n_layers: int, device: torch.device, *, has_final_norm=False, has_lm_head=False,
                 has_embedding=False, max_tokens=4096, max_batch=1, io_dtype: torch.dtype = torch.bfloat16):
        _lib.require_cuda()
        self.lib = _lib.load()
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("[ERROR] the B200 shard path runs on CUDA devices only (no CPU fallback)")
        self.dev_index = self.device.index if self.device.index is not None else torch.cuda.current_device()
        self.model_cfg = dict(model_cfg)
        self.n_layers = n_layers
        self.hidden_size = model_cfg["hidden_size"]
        self.vocab_size = model_cfg["vocab_size"]
        self.has_final_norm, self.has_lm_head, self.has_embedding = has_final_norm, has_lm_head, has_embedding
        self.max_tokens, self.max_batch = max_tokens, max_batch
        cfg = _lib.ShardConfig(
            model_cfg["hidden_size"], model_cfg["intermediate_size"], model_cfg["num_attention_heads"],
            model_cfg["num_key_value_heads"], model_cfg["head_dim"], model_cfg["vocab_size"],
            model_cfg["rms_norm_eps"], model_cfg["rope_theta"], n_layers,
            int(has_final_norm), int(has_lm_head), int(has_embedding), max_tokens, max_batch)
        h = C.c_void_p()
        _lib.check(self.lib.b200_shard_create(C.byref(cfg), self.dev_index, C.byref(h)))
        self.handle = h
        # the table torch would compute (bit-identical to the reference's LlamaRotaryEmbedding.inv_freq), scaled or not
        inv = rope_inv_freq(model_cfg["head_dim"], model_cfg["rope_theta"], model_cfg.get("rope_scaling"))
        _lib.check(self.lib.b200_shard_set_rope_inv_freq(self.handle, inv.numpy().ctypes.data_as(C.POINTER(C.c_float)), inv.numel()))
        if io_dtype not in _lib.TORCH_TO_DT:
            raise ValueError(f"[ERROR] dtype={io_dtype} is not built (bfloat16, float16 and float32 messages are)")
        self.io_dtype = io_dtype
        _lib.check(self.lib.b200_shard_set_io_dtype(self.handle, _lib.TORCH_TO_DT[io_dtype]))
        self.graphs = False
        self._step_bufs: dict = {}   # B -> persistent (hidden_in, hidden_out, tok, margin) for single-token calls

    # ---- weights
    def _stream(self):
        return _lib.stream_ptr(self.device)

    def set_layer(self, rel_idx: int, state: dict) -> None:
        ts = []
        dts = set()
        for k in LAYER_KEYS:
            if k not in state:
                raise RuntimeError(f"[ERROR] Missing key(s) in state_dict: \"{k}\"")
            t = state[k].to(self.device).contiguous()
            dts.add(t.dtype)
            ts.append(t)
        if len(dts) != 1 or next(iter(dts)) not in _lib.TORCH_TO_DT:
            raise ValueError(f"[ERROR] unsupported / mixed weight dtypes {dts}")
        _lib.check(self.lib.b200_shard_set_layer(self.handle, rel_idx, *[t.data_ptr() for t in ts],
                                                 _lib.TORCH_TO_DT[ts[0].dtype], self._stream()))

    def _set_single(self, fn, w: torch.Tensor) -> None:
        t = w.to(self.device).contiguous()
        if t.dtype not in _lib.TORCH_TO_DT:
            raise ValueError(f"[ERROR] unsupported weight dtype {t.dtype}")
        _lib.check(fn(self.handle, t.data_ptr(), _lib.TORCH_TO_DT[t.dtype], self._stream()))

    def set_final_norm(self, w): self._set_single(self.lib.b200_shard_set_final_norm, w)
    def set_lm_head(self, w): self._set_single(self.lib.b200_shard_set_lm_head, w)
    def set_embedding(self, w): self._set_single(self.lib.b200_shard_set_embedding, w)

    # ---- compute
    def new_kv(self, max_slots: int = 1, page_size: int = 64, n_pages: int | None = None, max_seq_len: int = 4096):
        return PagedKVCache(self, max_slots, page_size, n_pages, max_seq_len)

    def _slots(self, slots, B):
        if slots is None:
            slots = list(range(B))
        if len(slots) != B:
            raise ValueError("[ERROR] one KV slot per batch row is required")
        return (C.c_int32 * B)(*slots)

    def forward(self, hidden: torch.Tensor, kv: "PagedKVCache", slots=None, causal: bool = False,
                out: torch.Tensor | None = None) -> torch.Tensor:
        """LlamaShardPart.forward: [B,S,H] -> [B,S,H] in the shard's message dtype (bf16 unless chosen otherwise)."""
        hidden = self._check_hidden(hidden)
        B, S, _ = hidden.shape
        if S == 1 and self.graphs and out is None:
            # single-token calls replay a CUDA graph keyed by its pointers: run on persistent buffers and hand
            # back a fresh tensor (the reference's outputs are fresh tensors too)
            hin, hout, _, _ = self._bufs(B)
            hin.copy_(hidden)
            _lib.check(self.lib.b200_shard_forward(self.handle, kv.handle, self._slots(slots, B), hin.data_ptr(), B, 1,
                                                   int(causal), hout.data_ptr(), self._stream()))
            return hout.clone()
        if out is None:
            out = torch.empty_like(hidden)
        _lib.check(self.lib.b200_shard_forward(self.handle, kv.handle, self._slots(slots, B), hidden.data_ptr(), B, S,
                                               int(causal), out.data_ptr(), self._stream()))
        return out

    def _bufs(self, B: int):
        if B not in self._step_bufs:
            H = self.hidden_size
            self._step_bufs[B] = (torch.empty(B, 1, H, dtype=self.io_dtype, device=self.device),
                                  torch.empty(B, 1, H, dtype=self.io_dtype, device=self.device),
                                  torch.empty(B, dtype=torch.int64, device=self.device),
                                  torch.empty(B, dtype=torch.float32, device=self.device))
        return self._step_bufs[B]

    def forward_argmax(self, hidden: torch.Tensor, kv: "PagedKVCache", slots=None, causal: bool = False,
                       want_hidden: bool = False, want_margin: bool = False):
        hidden = self._check_hidden(hidden)
        B, S, _ = hidden.shape
        if S == 1 and self.graphs:
            hin, hout, tokb, marb = self._bufs(B)
            hin.copy_(hidden)
            _lib.check(self.lib.b200_shard_forward_argmax(
                self.handle, kv.handle, self._slots(slots, B), hin.data_ptr(), B, 1, int(causal),
                hout.data_ptr() if want_hidden else None, tokb.data_ptr(), marb.data_ptr() if want_margin else None,
                self._stream()))
            return tokb.clone(), (hout.clone() if want_hidden else None), (marb.clone() if want_margin else None)
        out = torch.empty_like(hidden) if want_hidden else None
        tok = torch.empty(B, dtype=torch.int64, device=self.device)
        margin = torch.empty(B, dtype=torch.float32, device=self.device) if want_margin else None
        _lib.check(self.lib.b200_shard_forward_argmax(self.handle, kv.handle, self._slots(slots, B), hidden.data_ptr(), B, S,
                                                      int(causal), _lib.ptr(out), tok.data_ptr(), _lib.ptr(margin),
                                                      self._stream()))
        return tok, out, margin

    def step_tokens(self, tok_in: torch.Tensor, kv: "PagedKVCache", slots=None, causal: bool = False,
                    tok_out: torch.Tensor | None = None, S: int = 1, hidden_out: torch.Tensor | None = None,
                    margin_out: torch.Tensor | None = None):
        """Token ids in -> token ids out (tail) / hidden out: embed + forward (+ lm_head, argmax) in one call.
        ``tok_in`` int64 [B*S] on the device; pass the same persistent ``tok_out`` buffer every step (it may
        alias ``tok_in``) so that the S == 1 CUDA graph is captured once and replayed."""
        B = tok_in.numel() // S
        if tok_out is None and self.has_lm_head:
            tok_out = torch.empty(B, dtype=torch.int64, device=self.device)
        _lib.check(self.lib.b200_shard_step_tokens(self.handle, kv.handle, self._slots(slots, B), tok_in.data_ptr(), B, S,
                                                   int(causal), _lib.ptr(hidden_out), _lib.ptr(tok_out),
                                                   _lib.ptr(margin_out), self._stream()))
        return tok_out if self.has_lm_head else hidden_out

    # ---- vocabulary-split tail (ring stages; include/b200shard.h)
    def partial_msg_bytes(self, B: int) -> int:
        return B * (self.hidden_size * 4 + 16)

    def forward_partial(self, hidden: torch.Tensor, kv: "PagedKVCache", split: int, msg_out: torch.Tensor, slots=None,
                        causal: bool = False) -> torch.Tensor:
        """Tail shard: layers + lm_head rows [0, split); ``msg_out`` (uint8, >= partial_msg_bytes(B)) receives the message for
        the stage that holds the other rows.  Pass the same persistent buffers every step (S == 1 replays a CUDA graph)."""
        hidden = self._check_hidden(hidden)
        B, S, _ = hidden.shape
        if msg_out.numel() * msg_out.element_size() < self.partial_msg_bytes(B):
            raise ValueError("[ERROR] message buffer too small")
        _lib.check(self.lib.b200_shard_forward_partial(self.handle, kv.handle, self._slots(slots, B), hidden.data_ptr(), B, S,
                                                       int(causal), int(split), msg_out.data_ptr(), self._stream()))
        return msg_out

    def set_vocab_tail(self, lm_head_rows: torch.Tensor, final_norm: torch.Tensor | None, split: int) -> None:
        """The stage after the tail: rows [split, V) of lm_head ([V - split, H]) and the final norm weight."""
        w = lm_head_rows.to(self.device).contiguous()
        n = None if final_norm is None else final_norm.to(device=self.device, dtype=w.dtype).contiguous()
        if w.dtype not in _lib.TORCH_TO_DT:
            raise ValueError(f"[ERROR] unsupported weight dtype {w.dtype}")
        _lib.check(self.lib.b200_shard_set_vocab_tail(self.handle, w.data_ptr(), _lib.ptr(n), int(split), _lib.TORCH_TO_DT[w.dtype],
                                                      self._stream()))

    def token_from_partial(self, msg: torch.Tensor, B: int, tok_out: torch.Tensor | None = None, want_margin: bool = False):
        if tok_out is None:
            tok_out = torch.empty(B, dtype=torch.int64, device=self.device)
        margin = torch.empty(B, dtype=torch.float32, device=self.device) if want_margin else None
        _lib.check(self.lib.b200_shard_token_from_partial(self.handle, msg.data_ptr(), B, tok_out.data_ptr(), _lib.ptr(margin),
                                                          self._stream()))
        return (tok_out, margin) if want_margin else tok_out

    def run_kernel(self, kv: "PagedKVCache", kind: int, layer: int, rows: int = 1) -> None:
        _lib.check(self.lib.b200_shard_run_kernel(self.handle, kv.handle, kind, layer, rows, self._stream()))

    def lm_head_argmax(self, hidden: torch.Tensor, want_margin: bool = False):
        hidden = self._check_hidden(hidden)
        B, S, _ = hidden.shape
        tok = torch.empty(B, dtype=torch.int64, device=self.device)
        margin = torch.empty(B, dtype=torch.float32, device=self.device) if want_margin else None
        _lib.check(self.lib.b200_shard_lm_head_argmax(self.handle, hidden.data_ptr(), B, S, tok.data_ptr(), _lib.ptr(margin),
                                                      self._stream()))
        return tok, margin

    def embed(self, ids: torch.Tensor) -> torch.Tensor:
        """embed_tokens(ids): int64 [...] -> [..., H] in the message dtype."""
        ids = ids.to(device=self.device, dtype=torch.long).contiguous()
        out = torch.empty(*ids.shape, self.hidden_size, dtype=self.io_dtype, device=self.device)
        _lib.check(self.lib.b200_shard_embed(self.handle, ids.data_ptr(), ids.numel(), out.data_ptr(), self._stream()))
        return out

    def _check_hidden(self, hidden: torch.Tensor) -> torch.Tensor:
        if hidden.ndim != 3 or hidden.shape[-1] != self.hidden_size:
            raise ValueError(f"[ERROR] hidden_states must be [B, S, {self.hidden_size}], got {tuple(hidden.shape)}")
        return hidden.to(device=self.device, dtype=self.io_dtype).contiguous()

    def counters(self) -> dict:
        arr = (C.c_uint64 * 4)()
        _lib.check(self.lib.b200_shard_counters(self.handle, arr))
        return {"launches": arr[0], "bytes": arr[1], "flops": arr[2], "calls": arr[3]}

    def reset_counters(self): _lib.check(self.lib.b200_shard_reset_counters(self.handle))
    def set_graphs(self, on: bool):
        _lib.check(self.lib.b200_shard_set_graphs(self.handle, int(on)))
        self.graphs = bool(on)

    def close(self):
        if getattr(self, "handle", None):
            self.lib.b200_shard_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class PagedKVCache:
    """Paged K/V storage of one shard.  Exposes ``get_seq_length()`` like the HF cache so that
    ``build_position_ids`` (utils/forwarding_utils.py:13-16) keeps working on it."""

    def __init__(self, shard: Shard, max_slots: int = 1, page_size: int = 64, n_pages: int | None = None,
                 max_seq_len: int = 4096):
        self.shard = shard
        self.lib = shard.lib
        self.max_slots, self.page_size = max_slots, page_size
        if n_pages is None:
            n_pages = max_slots * ((max_seq_len + page_size - 1) // page_size)
        self.n_pages = n_pages
        h = C.c_void_p()
        _lib.check(self.lib.b200_kv_create(shard.handle, max_slots, page_size, n_pages, C.byref(h)))
        self.handle = h

    def get_seq_length(self, slot: int = 0) -> int:
        return _lib.check(self.lib.b200_kv_seq_len(self.handle, slot))

    def reset(self, slot: int = -1) -> None:
        _lib.check(self.lib.b200_kv_reset(self.handle, slot, _lib.stream_ptr(self.shard.device)))

    def close(self):
        if getattr(self, "handle", None) and getattr(self.shard, "handle", None):
            self.lib.b200_kv_destroy(self.handle)
        self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
