"""-m gpu: oracle parity at the shapes bench.py and BASELINE configs[3] actually run (round-1 VERDICT, "What's weak" #1).

The small presets keep prompts below 256 tokens, so the prefill attention kernel's K/V rings never wrap, its score
buffers in tensor memory are never reused, and the CTA-pair GEMM never runs with the fused QKV+RoPE or residual
epilogues.  Here: Llama-2-7B LAYER dimensions (H = 4096, I = 11008, 32 heads; three layers, split 2 + 1), prompts of
512 .. 2 x 2048 tokens, compared with the CPU oracle (per-shard hidden state <= 2e-2 relative L2, greedy tokens equal
wherever the fp32 margin is above bf16 noise), plus the prefill attention kernel alone against fp64 torch.

Which kernels each case launches (tile counts from csrc/gemm_tcgen05.cu: the CTA-pair kernel needs M > 128 and >= 64
tiles of 256 x 256):
  1 x 512    QKV pair<QKV_ROPE> (96 tiles), gate/up pair<SWIGLU> (172), o_proj / down_proj single-CTA<RESID> (128 tiles
             of 128 x 128), attention 4 key tiles
  2 x 2048   every GEMM on the pair kernel (RESID: 256 tiles), attention 16 key tiles, causal and not
  1 x 2000   ragged M (2000 = 7 x 256 + 208), context not page aligned, then decode steps at T = 2001, 2002
  3 x 300    M = 900 (ragged last pair tile: 132 rows), RESID pair with exactly 64 tiles, 3 key tiles (partial last one)
  GQA 1 x 640 causal: 32 query heads on 8 K/V heads, QKV pair<QKV_ROPE> with N = 6144 (72 tiles)
"""
import pytest
import torch

from gpu_utils import build_chain, close_chain
from oracle import synth
from oracle.llama_oracle import OracleWorker, run_chain

pytestmark = pytest.mark.gpu

HIDDEN_TOL = 2e-2   # BASELINE.json north_star: bf16 path against the fp32 reference
TIE_TOL = 0.02      # top-1 margins below this are ties at bf16 precision (same rule as test_gpu_parity.py)


def rel_l2(a, b):
    return float((a.double() - b.double()).norm() / b.double().norm().clamp_min(1e-30))


@pytest.fixture(scope="module")
def tmp_root(tmp_path_factory):
    return str(tmp_path_factory.mktemp("shards7b"))


@pytest.mark.parametrize("preset,seed,B,S,n_new,causal", [
    ("l7b3", 301, 1, 512, 2, False),       # bench.py's prompt
    ("l7b3", 302, 2, 2048, 2, False),      # BASELINE configs[3] micro-batch, reference (non-causal) attention
    ("l7b3", 302, 2, 2048, 2, True),       # same, causal
    ("l7b3", 303, 1, 2000, 3, False),      # page-unaligned context, then decode on top of it
    ("l7b3", 304, 3, 300, 2, False),       # ragged tiles everywhere
    ("l7b3_gqa", 305, 1, 640, 2, True),    # grouped-query attention at pair-GEMM sizes
])
def test_prefill_at_bench_shapes_matches_oracle(preset, seed, B, S, n_new, causal, tmp_root):
    from llm_sharding_b200.chain import generate
    cfg = synth.PRESETS[preset]
    splits = [(0, 2), (2, 3)]
    ids = synth.prompt_ids(cfg, seed, B, S)
    o_tokens, o_states, o_margins = run_chain([OracleWorker(cfg, seed, s, e, i == 0, causal=causal) for i, (s, e) in enumerate(splits)],
                                              ids, n_new)
    workers = build_chain(tmp_root, preset, seed, splits, max_batch=B, max_tokens=B * S, causal=causal, max_seq_len=S + 64)
    try:
        tokens, states = generate(workers, ids, n_new, record_states=True, teacher_tokens=o_tokens)
        worst = 0.0
        for p, per_pass in enumerate(states):
            for i, h in enumerate(per_pass):
                assert torch.isfinite(h).all()
                worst = max(worst, rel_l2(h, o_states[p][i]))
        flips = [(p, b, float(o_margins[p, b])) for p in range(o_tokens.shape[0]) for b in range(B)
                 if int(tokens[p, b]) != int(o_tokens[p, b])]
        print(f"{preset} B={B} S={S} causal={causal}: worst hidden rel-L2 {worst:.2e}; token flips {flips}")
        assert worst <= HIDDEN_TOL, worst
        assert not [f for f in flips if f[2] >= TIE_TOL], flips
    finally:
        close_chain(workers)


def _paged(B, n_kv, T, page_size, g, extra_pages=3):
    per = (T + page_size - 1) // page_size
    n_pages = B * per + extra_pages
    # NaN everywhere the sequences never wrote: a kernel that lets a masked key leak into a sum is caught
    k_pool = torch.full((n_pages, n_kv, page_size, 128), float("nan"), device="cuda", dtype=torch.bfloat16)
    v_pool = torch.full_like(k_pool, float("nan"))
    perm = torch.randperm(n_pages, generator=torch.Generator().manual_seed(11)).tolist()
    bt = torch.full((B, per), -1, dtype=torch.int32)
    ks, vs = [], []
    for b in range(B):
        k = torch.randn(n_kv, T, 128, device="cuda", generator=g).bfloat16()
        v = torch.randn(n_kv, T, 128, device="cuda", generator=g).bfloat16()
        ks.append(k); vs.append(v)
        for pi in range(per):
            page = perm.pop()
            bt[b, pi] = page
            lo, hi = pi * page_size, min(T, (pi + 1) * page_size)
            k_pool[page, :, : hi - lo] = k[:, lo:hi]
            v_pool[page, :, : hi - lo] = v[:, lo:hi]
            if hi - lo < page_size:   # the tail of a sequence's last page is read (and masked): finite, but huge
                k_pool[page, :, hi - lo:] = 1e4
                v_pool[page, :, hi - lo:] = 1e4
    return k_pool, v_pool, bt.cuda(), ks, vs, per, n_pages


@pytest.mark.parametrize("B,n_heads,n_kv,base,S,causal", [
    (1, 4, 4, 0, 129, False), (1, 4, 4, 0, 129, True),
    (1, 4, 2, 0, 384, False), (1, 4, 2, 0, 385, True),
    (2, 8, 8, 0, 1024, False), (2, 8, 8, 0, 1024, True),
    (1, 32, 32, 0, 2048, False), (1, 32, 8, 0, 2048, True),
    (2, 4, 4, 300, 100, False), (2, 4, 4, 300, 100, True),     # a chunk of new tokens on top of a cached context
    (1, 2, 2, 1000, 1, False),                                  # one query row over a long context
])
def test_prefill_attention_kernel_against_fp64(B, n_heads, n_kv, base, S, causal):
    """The tcgen05 / TMA prefill attention alone, through b200_attn_prefill.  Error budget: the probabilities are rounded
    to bf16 (2^-9 relative each) before the second MMA and the output to bf16; q arrives as a hi + lo bf16 pair
    (~2^-17); K/V are exact bf16 inputs.  Averaged over a row that is ~2e-3 relative L2; bound 4e-3."""
    from llm_sharding_b200 import _lib
    _lib.require_cuda()
    lib = _lib.load()
    g = torch.Generator(device="cuda").manual_seed(B * 1000 + S + base)
    T = base + S
    k_pool, v_pool, bt, ks, vs, per, n_pages = _paged(B, n_kv, T, 64, g)
    q = torch.randn(B, S, n_heads, 128, device="cuda", generator=g) * 2.0
    q_hi = q.bfloat16()
    q_lo = (q - q_hi.float()).bfloat16()
    base_len = torch.full((B,), base, dtype=torch.int32, device="cuda")
    out = torch.full((B, S, n_heads, 128), float("nan"), device="cuda", dtype=torch.bfloat16)
    _lib.check(lib.b200_attn_prefill(q_hi.data_ptr(), q_lo.data_ptr(), k_pool.data_ptr(), v_pool.data_ptr(), bt.data_ptr(),
                                     base_len.data_ptr(), B, S, n_heads, n_kv, 64, per, n_pages, int(causal), 0,
                                     out.data_ptr(), torch.cuda.current_stream().cuda_stream))
    torch.cuda.synchronize()
    assert torch.isfinite(out.float()).all()
    rep = n_heads // n_kv
    qq = (q_hi.double() + q_lo.double())
    worst = 0.0
    for b in range(B):
        k = ks[b].double().repeat_interleave(rep, dim=0)   # [n_heads, T, 128]
        v = vs[b].double().repeat_interleave(rep, dim=0)
        sc = torch.einsum("shd,htd->hst", qq[b], k) * 128 ** -0.5
        if causal:
            i = torch.arange(S, device="cuda")[:, None] + base
            j = torch.arange(T, device="cuda")[None, :]
            sc = sc.masked_fill(~(j <= i), float("-inf"))
        ref = torch.einsum("hst,htd->shd", torch.softmax(sc, dim=-1), v)
        worst = max(worst, rel_l2(out[b].float(), ref))
    assert worst < 4e-3, worst
