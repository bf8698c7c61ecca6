"""-m gpu: the vocabulary-split tail (b200_shard_forward_partial / b200_shard_set_vocab_tail / b200_shard_token_from_partial).
The tail shard scores lm_head rows [0, split) and the stage after it the rest; token ids and top-1 margins must be the very
bits the unsplit tail produces (same GEMV, same fused final norm, arg-max = first index of the maximum)."""
import os

import pytest
import torch

from gpu_utils import build_chain, close_chain, shard_dir
from oracle import synth

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def tmp_root(tmp_path_factory):
    return str(tmp_path_factory.mktemp("vsplit"))


@pytest.mark.parametrize("preset,B,S,split", [
    ("tiny", 1, 1, 500), ("tiny", 1, 7, 500), ("tiny", 3, 1, 500), ("tiny", 2, 5, 1), ("tiny", 2, 1, 999), ("tiny", 1, 1, 333),
    ("gqa", 3, 4, 512),
])
def test_split_tail_reproduces_the_unsplit_tail(preset, B, S, split, tmp_root):
    cfg = synth.PRESETS[preset]
    seed = 140 + B + S
    workers = build_chain(tmp_root, preset, seed, [(0, 2), (2, 4)], max_batch=2 * B, max_tokens=2 * B * max(S, 8))
    try:
        head, tail = workers
        eng_t, kv_t, eng_h = tail.shard.engine, tail.past_key_value, head.shard.engine
        d = shard_dir(tmp_root, preset, seed)
        lm = torch.load(os.path.join(d, "lm_head.pth"), map_location="cpu")["weight"]
        fn = torch.load(os.path.join(d, "final_norm.pth"), map_location="cpu")["weight"]
        eng_h.set_vocab_tail(lm[split:], fn, split)
        g = torch.Generator().manual_seed(seed)
        msg = torch.empty(eng_t.partial_msg_bytes(B), dtype=torch.uint8, device="cuda:0")
        tokb = torch.empty(B, dtype=torch.int64, device="cuda:0")
        for step, s_len in enumerate([S, 1, 1]):   # a prompt-like call, then two single-token calls (CUDA-graph capture and replay)
            hidden = (torch.randn(B, s_len, cfg.hidden_size, generator=g) * 0.7).bfloat16().cuda()
            tok_ref, _, mar_ref = eng_t.forward_argmax(hidden, kv_t, slots=list(range(B)), want_margin=True)
            eng_t.forward_partial(hidden, kv_t, split, msg, slots=list(range(B, 2 * B)))
            tok, mar = eng_h.token_from_partial(msg, B, tok_out=tokb, want_margin=True)
            torch.cuda.synchronize()
            assert torch.equal(tok.cpu(), tok_ref.cpu()), (step, tok, tok_ref)
            assert torch.equal(mar.cpu(), mar_ref.cpu()), (step, mar, mar_ref)
            x = msg[:B * cfg.hidden_size * 4].view(torch.float32).view(B, cfg.hidden_size)
            assert torch.isfinite(x).all()
    finally:
        close_chain(workers)


@pytest.mark.parametrize("B,S", [(4, 1), (8, 1), (16, 3), (32, 1)])
def test_split_tail_batched_rows(B, S, tmp_root):
    """More than three rows: both halves score through fp16 operands on the tensor cores, the unsplit tail through the folded
    norm — same arithmetic up to fp16 rounding of the operand, so tokens agree wherever the top-1 margin is not a tie at that
    precision, and the margins agree to 1e-2 of the logit scale."""
    cfg = synth.PRESETS["tiny"]
    seed, split = 160 + B, 488
    workers = build_chain(tmp_root, "tiny", seed, [(0, 2), (2, 4)], max_batch=2 * B, max_tokens=2 * B * 8)
    try:
        head, tail = workers
        eng_t, kv_t, eng_h = tail.shard.engine, tail.past_key_value, head.shard.engine
        d = shard_dir(tmp_root, "tiny", seed)
        lm = torch.load(os.path.join(d, "lm_head.pth"), map_location="cpu")["weight"]
        fn = torch.load(os.path.join(d, "final_norm.pth"), map_location="cpu")["weight"]
        eng_h.set_vocab_tail(lm[split:], fn, split)
        g = torch.Generator().manual_seed(seed)
        msg = torch.empty(eng_t.partial_msg_bytes(B), dtype=torch.uint8, device="cuda:0")
        for s_len in (S, 1):
            hidden = (torch.randn(B, s_len, cfg.hidden_size, generator=g) * 0.7).bfloat16().cuda()
            tok_ref, _, mar_ref = eng_t.forward_argmax(hidden, kv_t, slots=list(range(B)), want_margin=True)
            eng_t.forward_partial(hidden, kv_t, split, msg, slots=list(range(B, 2 * B)))
            tok, mar = eng_h.token_from_partial(msg, B, want_margin=True)
            torch.cuda.synchronize()
            tie = mar_ref.cpu() < 5e-3
            assert torch.equal(tok.cpu()[~tie], tok_ref.cpu()[~tie]), (tok, tok_ref, mar_ref)
            assert (mar.cpu() - mar_ref.cpu()).abs().max() < 2e-2, (mar, mar_ref)
    finally:
        close_chain(workers)


def test_split_tail_argument_checks(tmp_root):
    from llm_sharding_b200 import _lib
    workers = build_chain(tmp_root, "tiny", 150, [(0, 2), (2, 4)], max_batch=8, max_tokens=64)
    try:
        head, tail = workers
        eng_t, kv_t, eng_h = tail.shard.engine, tail.past_key_value, head.shard.engine
        V, H = synth.PRESETS["tiny"].vocab_size, synth.PRESETS["tiny"].hidden_size
        msg = torch.empty(eng_t.partial_msg_bytes(4), dtype=torch.uint8, device="cuda:0")
        hidden = torch.zeros(4, 1, H, dtype=torch.bfloat16, device="cuda:0")
        with pytest.raises(ValueError):   # more rows than the shard was created for
            eng_t.forward_partial(torch.zeros(9, 1, H, dtype=torch.bfloat16, device="cuda:0"), kv_t, V // 2,
                                  torch.empty(eng_t.partial_msg_bytes(9), dtype=torch.uint8, device="cuda:0"), slots=list(range(8)) + [0])
        with pytest.raises(ValueError):   # a split outside the vocabulary
            eng_t.forward_partial(hidden[:1].contiguous(), kv_t, V, msg, slots=[5])
        with pytest.raises(RuntimeError):   # no vocabulary tail loaded on this shard
            eng_h.token_from_partial(msg, 1)
        with pytest.raises(RuntimeError):   # the head shard holds no lm_head: it cannot produce the partial message
            eng_h.forward_partial(hidden[:1].contiguous(), head.past_key_value, V // 2, msg)
        assert _lib is not None
    finally:
        close_chain(workers)
