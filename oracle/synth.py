"""TEST INFRASTRUCTURE — deterministic synthetic Llama shards (not shipped, not timed).

Writes a shard folder in exactly the layout the reference's ``ModelSharder.save_shards``
produces (``/root/reference/utils/model_sharder.py:64-94``): ``config.json`` + tokenizer
files, ``embedding.pth``, ``block_{i}.pth`` (the ``LlamaDecoderLayer.state_dict()`` key set
listed in SURVEY.md §8 a17), ``final_norm.pth``, ``lm_head.pth``.  Both the reference's
``NodeWorker`` (``utils/node_worker.py:117-124,160-163``; ``utils/shard_loader.py:42-55``)
and this repo's mirror load it unchanged.

Values: numpy ``PCG64`` streams keyed by ``(seed, tensor id)`` — reproducible on any
machine — ``N(0, init_std)`` for matrices and ``1 + 0.1 N(0,1)`` for RMSNorm weights (HF
default init uses exactly 1; a non-trivial weight is used so that a mis-wired norm weight
cannot hide), every value rounded to bf16 so that the fp32 oracle and the bf16 GPU path
see *the same* weights.
"""
from __future__ import annotations

import hashlib
import json
import os
from dataclasses import dataclass, asdict

import numpy as np
import torch

LAYER_KEYS = (
    "self_attn.q_proj.weight",
    "self_attn.k_proj.weight",
    "self_attn.v_proj.weight",
    "self_attn.o_proj.weight",
    "mlp.gate_proj.weight",
    "mlp.up_proj.weight",
    "mlp.down_proj.weight",
    "input_layernorm.weight",
    "post_attention_layernorm.weight",
)


@dataclass(frozen=True)
class SynthConfig:
    name: str
    vocab_size: int
    hidden_size: int
    intermediate_size: int
    num_hidden_layers: int
    num_attention_heads: int
    num_key_value_heads: int
    max_position_embeddings: int = 4096
    rms_norm_eps: float = 1e-5
    rope_theta: float = 10000.0
    init_std: float = 0.02
    rope_scaling: dict | None = None   # e.g. the Llama-3.x "llama3" rescaling

    @property
    def head_dim(self) -> int:
        return self.hidden_size // self.num_attention_heads


PRESETS = {
    # head_dim is 128 everywhere: it is the only head size the CUDA path is built for.
    "tiny": SynthConfig("tiny", 1000, 256, 640, 4, 2, 2, max_position_embeddings=512),
    "small": SynthConfig("small", 2000, 512, 1408, 6, 4, 4, max_position_embeddings=1024),
    "gqa": SynthConfig("gqa", 1000, 512, 1024, 4, 4, 2, max_position_embeddings=512),
    # Llama-3.2-style: GQA + "llama3" RoPE rescaling (what start_node.py:14 of the reference actually points at);
    # original_max_position_embeddings is small so that test positions reach the rescaled / smoothed bands
    "llama3ish": SynthConfig("llama3ish", 1000, 512, 1024, 4, 4, 2, max_position_embeddings=2048, rope_theta=500000.0,
                             rope_scaling={"rope_type": "llama3", "factor": 8.0, "low_freq_factor": 1.0,
                                           "high_freq_factor": 4.0, "original_max_position_embeddings": 64}),
    "llama2_7b": SynthConfig("llama2_7b", 32000, 4096, 11008, 32, 32, 32),
    # eight small layers: one per GPU of an 8-way ring (the parity check bench_chain.py runs before its timed region)
    "ring8": SynthConfig("ring8", 1000, 256, 640, 8, 2, 2, max_position_embeddings=512),
    # 28 small layers: the layer count the reference's go_through_every_shards hard-codes (0-7 / 7-14 / 14-21 / 21-28,
    # utils/node_profiler.py:1207-1210, i.e. Llama-3.2-3B's)
    "l28": SynthConfig("l28", 1000, 256, 640, 28, 2, 2, max_position_embeddings=512),
    # Llama-2-7B LAYER dimensions (what bench.py and BASELINE configs[3] run) with three layers and a small vocabulary, so
    # that the CPU oracle finishes 512..4096-token prompts in seconds: reaches the CTA-pair GEMM epilogues and the
    # steady state of the prefill-attention pipeline (>= 3 key tiles) that the small presets never touch
    "l7b3": SynthConfig("l7b3", 8192, 4096, 11008, 3, 32, 32),
    # same with grouped-query attention (Llama-3-8B-like head layout: 32 query heads on 8 K/V heads)
    "l7b3_gqa": SynthConfig("l7b3_gqa", 8192, 4096, 11008, 3, 32, 8),
}
_BIG_CACHE: dict = {}   # (cfg.name, seed, layer) -> state dict, for the few-layer 7B-dimension presets only


def _rng(seed: int, *tag: int) -> np.random.Generator:
    return np.random.Generator(np.random.PCG64(np.random.SeedSequence([seed, *tag])))


def _bf16_round(a: np.ndarray) -> torch.Tensor:
    """fp32 ndarray -> torch bf16 (round-to-nearest-even)."""
    return torch.from_numpy(a).to(torch.bfloat16)


def matrix(seed: int, tag: tuple, rows: int, cols: int, std: float) -> torch.Tensor:
    a = _rng(seed, *tag).standard_normal((rows, cols), dtype=np.float32)
    a *= np.float32(std)
    return _bf16_round(a)


def norm_weight(seed: int, tag: tuple, n: int) -> torch.Tensor:
    a = _rng(seed, *tag).standard_normal((n,), dtype=np.float32)
    return _bf16_round(np.float32(1.0) + np.float32(0.1) * a)


def layer_state(cfg: SynthConfig, seed: int, layer: int) -> dict:
    """bf16 tensors for absolute layer index ``layer`` keyed like ``block_i.pth``."""
    cacheable = cfg.hidden_size >= 4096 and cfg.num_hidden_layers <= 4   # generated twice per test (oracle + GPU chain)
    if cacheable and (cfg.name, seed, layer) in _BIG_CACHE:
        return dict(_BIG_CACHE[(cfg.name, seed, layer)])
    state = _layer_state(cfg, seed, layer)
    if cacheable:
        if len(_BIG_CACHE) >= 6:
            _BIG_CACHE.clear()
        _BIG_CACHE[(cfg.name, seed, layer)] = state
    return dict(state)


def _layer_state(cfg: SynthConfig, seed: int, layer: int) -> dict:
    H, I, d = cfg.hidden_size, cfg.intermediate_size, cfg.head_dim
    nq, nkv = cfg.num_attention_heads * d, cfg.num_key_value_heads * d
    s = cfg.init_std
    t = lambda k: (1, layer, k)
    return {
        "self_attn.q_proj.weight": matrix(seed, t(0), nq, H, s),
        "self_attn.k_proj.weight": matrix(seed, t(1), nkv, H, s),
        "self_attn.v_proj.weight": matrix(seed, t(2), nkv, H, s),
        "self_attn.o_proj.weight": matrix(seed, t(3), H, nq, s),
        "mlp.gate_proj.weight": matrix(seed, t(4), I, H, s),
        "mlp.up_proj.weight": matrix(seed, t(5), I, H, s),
        "mlp.down_proj.weight": matrix(seed, t(6), H, I, s),
        "input_layernorm.weight": norm_weight(seed, t(7), H),
        "post_attention_layernorm.weight": norm_weight(seed, t(8), H),
    }


def embedding_state(cfg: SynthConfig, seed: int) -> dict:
    return {"weight": matrix(seed, (2, 0, 0), cfg.vocab_size, cfg.hidden_size, cfg.init_std)}


def final_norm_state(cfg: SynthConfig, seed: int) -> dict:
    return {"weight": norm_weight(seed, (3, 0, 0), cfg.hidden_size)}


def lm_head_state(cfg: SynthConfig, seed: int) -> dict:
    return {"weight": matrix(seed, (4, 0, 0), cfg.vocab_size, cfg.hidden_size, cfg.init_std)}


def prompt_ids(cfg: SynthConfig, seed: int, batch: int, seq_len: int) -> torch.Tensor:
    """int64 [B,S]; ids >= 3 so the special ids 0..2 never appear in a prompt."""
    ids = _rng(seed, 9, batch, seq_len).integers(3, cfg.vocab_size, size=(batch, seq_len), dtype=np.int64)
    return torch.from_numpy(ids)


def tensor_digest(t: torch.Tensor) -> str:
    return hashlib.sha256(t.contiguous().view(torch.uint8).numpy().tobytes()).hexdigest()[:16]


def hf_config_dict(cfg: SynthConfig) -> dict:
    return {
        "architectures": ["LlamaForCausalLM"],
        "model_type": "llama",
        "vocab_size": cfg.vocab_size,
        "hidden_size": cfg.hidden_size,
        "intermediate_size": cfg.intermediate_size,
        "num_hidden_layers": cfg.num_hidden_layers,
        "num_attention_heads": cfg.num_attention_heads,
        "num_key_value_heads": cfg.num_key_value_heads,
        "max_position_embeddings": cfg.max_position_embeddings,
        "rms_norm_eps": cfg.rms_norm_eps,
        "rope_theta": cfg.rope_theta,
        **({"rope_scaling": dict(cfg.rope_scaling)} if cfg.rope_scaling else {}),
        "hidden_act": "silu",
        "attention_bias": False,
        "mlp_bias": False,
        "tie_word_embeddings": False,
        "bos_token_id": 1,
        "eos_token_id": 2,
        "pad_token_id": 0,
    }


def write_tokenizer(path: str, vocab_size: int) -> None:
    """Synthetic WordLevel tokenizer, ids 0..V-1, loadable by ``AutoTokenizer.from_pretrained``
    (what ``utils/node_worker.py:117`` calls)."""
    from tokenizers import Tokenizer, models, pre_tokenizers
    from transformers import PreTrainedTokenizerFast

    vocab = {"<unk>": 0, "<s>": 1, "</s>": 2}
    for i in range(3, vocab_size):
        vocab[f"t{i}"] = i
    tok = Tokenizer(models.WordLevel(vocab=vocab, unk_token="<unk>"))
    tok.pre_tokenizer = pre_tokenizers.WhitespaceSplit()
    fast = PreTrainedTokenizerFast(tokenizer_object=tok, unk_token="<unk>", bos_token="<s>",
                                   eos_token="</s>", pad_token="<unk>")
    fast.save_pretrained(path)


def state_provider(cfg: SynthConfig, seed: int):
    """``file name -> state_dict`` straight from the generators (no .pth round trip): what the NodeWorker mirror accepts
    through ``state_provider=`` for the 7B-dimension presets, whose shard files would be gigabytes."""
    def provide(name: str) -> dict:
        if name.startswith("block_"):
            return layer_state(cfg, seed, int(name[len("block_"):-len(".pth")]))
        if name == "embedding.pth":
            return embedding_state(cfg, seed)
        if name == "final_norm.pth":
            return final_norm_state(cfg, seed)
        if name == "lm_head.pth":
            return lm_head_state(cfg, seed)
        raise FileNotFoundError(name)
    return provide


def write_model_dir(path: str, cfg: SynthConfig) -> str:
    """config.json + tokenizer + an empty embedding.pth placeholder (for use with ``state_provider``)."""
    os.makedirs(path, exist_ok=True)
    with open(os.path.join(path, "config.json"), "w") as f:
        json.dump(hf_config_dict(cfg), f, indent=1)
    write_tokenizer(path, cfg.vocab_size)
    open(os.path.join(path, "embedding.pth"), "wb").close()
    return path


def write_shards(path: str, cfg: SynthConfig, seed: int, dtype: torch.dtype = torch.float32,
                 layers: range | None = None, with_tokenizer: bool = True) -> dict:
    """Write the shard folder; returns ``{file: digest}`` of the bf16 source values.

    ``dtype`` is the storage dtype of the ``.pth`` files (the reference loads them into
    modules of its own ``dtype``); values are bf16-representable either way.
    """
    os.makedirs(path, exist_ok=True)
    with open(os.path.join(path, "config.json"), "w") as f:
        json.dump(hf_config_dict(cfg), f, indent=1)
    with open(os.path.join(path, "synth.json"), "w") as f:
        json.dump({"seed": seed, "dtype": str(dtype), **asdict(cfg)}, f, indent=1)
    if with_tokenizer:
        write_tokenizer(path, cfg.vocab_size)
    digests = {}

    def save(name: str, state: dict) -> None:
        h = hashlib.sha256()
        for k in sorted(state):
            h.update(tensor_digest(state[k]).encode())
        digests[name] = h.hexdigest()[:16]
        torch.save({k: v.to(dtype) for k, v in state.items()}, os.path.join(path, name))

    save("embedding.pth", embedding_state(cfg, seed))
    for i in (layers if layers is not None else range(cfg.num_hidden_layers)):
        save(f"block_{i}.pth", layer_state(cfg, seed, i))
    save("final_norm.pth", final_norm_state(cfg, seed))
    save("lm_head.pth", lm_head_state(cfg, seed))
    return digests
