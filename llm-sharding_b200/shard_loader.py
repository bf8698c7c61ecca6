"""``LlamaShardPart`` with the reference's constructor and ``forward`` signature
(utils/shard_loader.py:13-15,57), backed by ``libb200shard.so`` instead of a ``ModuleList`` of
Hugging Face ``LlamaDecoderLayer``."""
from __future__ import annotations

import os

import torch

from .engine import PagedKVCache, Shard, read_model_config


class LlamaShardPart:
    def __init__(self, shards_path: str, shard_weights: list[str], start: int, end: int,
                 device="cuda:0", dtype=torch.bfloat16,
                 add_final_norm=False, final_norm_weight=None, *,
                 lm_head_weight: str | None = None, embedding_weight: str | None = None,
                 max_tokens: int = 4096, max_batch: int = 1, state_provider=None):
        """Same positional parameters as the reference.  ``dtype`` is the dtype of the hidden states that enter and leave
        (the reference's module dtype: messages are cast to it, utils/node_worker.py:252); the arithmetic inside is the
        same for all three (fp32 residual stream, fp16 tensor-core operands, bf16 K/V).  Keyword-only extras: ``lm_head_weight`` /
        ``embedding_weight`` (file names) put the tail's ``lm_head`` and the entry node's embedding
        behind the same C handle; ``max_tokens`` / ``max_batch`` size the workspace; ``state_provider``
        (``file name -> state_dict``) replaces ``torch.load`` of the shard files (synthetic weights)."""
        self.shards_path = shards_path
        self.shard_weights = shard_weights
        self.start = start
        self.end = end
        self.device = torch.device(device)
        self.dtype = dtype
        if dtype not in (torch.bfloat16, torch.float16, torch.float32):
            raise ValueError(f"[ERROR] dtype={dtype} is not built (bfloat16, float16 and float32 are)")
        if len(self.shard_weights) != self.end - self.start:
            raise ValueError("[ERROR] String list: shard_weights length must be equal to (end - start)")
        if add_final_norm and final_norm_weight is None:
            raise ValueError("[ERROR] final_norm_weight is required when add_final_norm is True")
        self.config = read_model_config(shards_path)
        self.engine = Shard(self.config, end - start, self.device, has_final_norm=bool(add_final_norm),
                            has_lm_head=lm_head_weight is not None, has_embedding=embedding_weight is not None,
                            max_tokens=max_tokens, max_batch=max_batch, io_dtype=dtype)
        load = state_provider or (lambda fname: torch.load(os.path.join(shards_path, fname), map_location="cpu"))
        for rel_idx, fname in enumerate(self.shard_weights):  # shard-relative index, utils/shard_loader.py:36-45
            state = load(fname)
            self.engine.set_layer(rel_idx, state)
            del state
        if add_final_norm:
            self.final_norm_weight = final_norm_weight
            self.engine.set_final_norm(load(final_norm_weight)["weight"])
        if lm_head_weight is not None:
            self.engine.set_lm_head(load(lm_head_weight)["weight"])
        if embedding_weight is not None:
            self.engine.set_embedding(load(embedding_weight)["weight"])

    def eval(self):  # called at utils/node_worker.py:179; there is no training mode here
        return self

    def new_cache(self, max_slots: int = 1, max_seq_len: int | None = None, page_size: int = 64) -> PagedKVCache:
        return self.engine.new_kv(max_slots, page_size, None, max_seq_len or self.config["max_position_embeddings"])

    def forward(self, hidden_states, attention_mask=None, past_key_value=None, rotary_emb=None, *, slots=None):
        """hidden_states [B,S,H] -> [B,S,H] (in ``dtype``).

        ``past_key_value`` must be a ``PagedKVCache`` of this shard (the reference passes its
        ``DynamicCache``).  ``rotary_emb`` is accepted for signature parity and ignored: cos/sin
        are recomputed on the device from the cache length.  ``attention_mask=None`` is the
        reference's behaviour, i.e. NON-causal prefill; pass ``"causal"`` for
        real-Llama masking."""
        if not isinstance(past_key_value, PagedKVCache):
            raise ValueError("[ERROR] Unsupported past_key_value structure")
        if attention_mask is not None and not (isinstance(attention_mask, str) and attention_mask == "causal"):
            raise ValueError("[ERROR] attention_mask must be None (reference behaviour) or \"causal\"")
        return self.engine.forward(hidden_states, past_key_value, slots, causal=attention_mask is not None)

    __call__ = forward
