"""Random-init Llama weights for benchmarking and demos (no checkpoint or network needed): a shard
folder with only ``config.json`` + tokenizer files, and a ``state_provider`` that fabricates each
``block_i.pth`` / ``embedding.pth`` / ``final_norm.pth`` / ``lm_head.pth`` state dict directly on
the GPU (HF default init: N(0, 0.02) matrices, unit norm weights).  ``LlamaShardPart`` /
``NodeWorker`` take it through their ``state_provider=`` keyword instead of ``torch.load``."""
from __future__ import annotations

import json
import os

import torch

LLAMA2_7B = dict(vocab_size=32000, hidden_size=4096, intermediate_size=11008, num_hidden_layers=32,
                 num_attention_heads=32, num_key_value_heads=32, max_position_embeddings=4096,
                 rms_norm_eps=1e-5, rope_theta=10000.0)


def write_model_dir(path: str, cfg: dict = LLAMA2_7B) -> str:
    os.makedirs(path, exist_ok=True)
    full = dict(cfg, architectures=["LlamaForCausalLM"], model_type="llama", hidden_act="silu",
                attention_bias=False, mlp_bias=False, tie_word_embeddings=False,
                bos_token_id=1, eos_token_id=2, pad_token_id=0)
    with open(os.path.join(path, "config.json"), "w") as f:
        json.dump(full, f, indent=1)
    from tokenizers import Tokenizer, models, pre_tokenizers
    from transformers import PreTrainedTokenizerFast

    vocab = {"<unk>": 0, "<s>": 1, "</s>": 2}
    vocab.update({f"t{i}": i for i in range(3, cfg["vocab_size"])})
    tok = Tokenizer(models.WordLevel(vocab=vocab, unk_token="<unk>"))
    tok.pre_tokenizer = pre_tokenizers.WhitespaceSplit()
    PreTrainedTokenizerFast(tokenizer_object=tok, unk_token="<unk>", bos_token="<s>", eos_token="</s>",
                            pad_token="<unk>").save_pretrained(path)
    open(os.path.join(path, "embedding.pth"), "wb").close()  # placeholder: the provider supplies the tensor
    return path


class RandomStateProvider:
    """``provider(file_name) -> state_dict`` with bf16 tensors on ``device``; deterministic per file."""

    def __init__(self, cfg: dict = LLAMA2_7B, device="cuda:0", seed: int = 0, std: float = 0.02):
        self.cfg, self.device, self.seed, self.std = cfg, torch.device(device), seed, std

    def _mat(self, g, rows, cols):
        return (torch.randn(rows, cols, device=self.device, generator=g, dtype=torch.float32) * self.std).bfloat16()

    def __call__(self, name: str) -> dict:
        c = self.cfg
        H, I, V = c["hidden_size"], c["intermediate_size"], c["vocab_size"]
        d = H // c["num_attention_heads"]
        nq, nkv = c["num_attention_heads"] * d, c["num_key_value_heads"] * d
        g = torch.Generator(device=self.device)
        g.manual_seed(self.seed * 1000003 + sum(ord(ch) * (i + 1) for i, ch in enumerate(name)))
        ones = lambda: torch.ones(H, device=self.device, dtype=torch.bfloat16)
        if name.startswith("block_"):
            return {
                "self_attn.q_proj.weight": self._mat(g, nq, H), "self_attn.k_proj.weight": self._mat(g, nkv, H),
                "self_attn.v_proj.weight": self._mat(g, nkv, H), "self_attn.o_proj.weight": self._mat(g, H, nq),
                "mlp.gate_proj.weight": self._mat(g, I, H), "mlp.up_proj.weight": self._mat(g, I, H),
                "mlp.down_proj.weight": self._mat(g, H, I),
                "input_layernorm.weight": ones(), "post_attention_layernorm.weight": ones(),
            }
        if name == "final_norm.pth":
            return {"weight": ones()}
        if name in ("lm_head.pth", "embedding.pth"):
            return {"weight": self._mat(g, V, H)}
        raise FileNotFoundError(name)
