"""-m gpu: size-independent properties at Llama-2-7B layer dimensions (H=4096, 32 heads, I=11008, 4 layers of
on-device random-init weights), where a CPU oracle run would take minutes: the prefill path (tcgen05 GEMMs + tiled
attention) and the decode path (weight-streaming kernels + split-K attention) must agree with each other, decode must
be bit-reproducible, identical rows must give identical results, and the KV bookkeeping must hold at its limits."""
import pytest
import torch

pytestmark = pytest.mark.gpu

TOL = 2e-2  # north-star hidden-state tolerance; the two paths differ only by 16-bit operand roundings


def rel_l2(a, b):
    return float((a.double() - b.double()).norm() / b.double().norm().clamp_min(1e-30))


@pytest.fixture(scope="module")
def shard7b():
    from llm_sharding_b200 import synthetic
    from llm_sharding_b200.engine import Shard
    cfg = dict(synthetic.LLAMA2_7B, num_hidden_layers=4)
    mc = {"hidden_size": 4096, "intermediate_size": 11008, "num_attention_heads": 32, "num_key_value_heads": 32,
          "head_dim": 128, "vocab_size": 32000, "rms_norm_eps": 1e-5, "rope_theta": 10000.0, "num_hidden_layers": 4}
    prov = synthetic.RandomStateProvider(cfg, "cuda:0", seed=3)
    sh = Shard(mc, 4, torch.device("cuda:0"), has_final_norm=True, has_lm_head=True, has_embedding=True,
               max_tokens=1024, max_batch=8)
    for l in range(4):
        sh.set_layer(l, prov(f"block_{l}.pth"))
    sh.set_final_norm(prov("final_norm.pth")["weight"])
    sh.set_lm_head(prov("lm_head.pth")["weight"])
    sh.set_embedding(prov("embedding.pth")["weight"])
    yield sh
    sh.close()


def test_causal_prefill_equals_token_by_token_decode(shard7b):
    """With causal masking, pushing S tokens at once must give the same last hidden state as pushing them one at a
    time through the decode kernels (up to operand rounding): couples every prefill kernel to its decode twin."""
    sh = shard7b
    S = 200  # crosses the 64- and 128-token page boundaries, not a multiple of any tile
    g = torch.Generator(device="cuda").manual_seed(1)
    ids = torch.randint(3, 32000, (1, S), device="cuda", generator=g)
    x = sh.embed(ids)
    kv_a, kv_b = sh.new_kv(1, max_seq_len=256), sh.new_kv(1, max_seq_len=256)
    out_prefill = sh.forward(x, kv_a, [0], causal=True)
    outs = [sh.forward(x[:, t:t + 1].contiguous(), kv_b, [0], causal=True) for t in range(S)]
    out_steps = torch.cat(outs, dim=1)
    torch.cuda.synchronize()
    assert kv_a.get_seq_length(0) == kv_b.get_seq_length(0) == S
    assert rel_l2(out_prefill[:, -1], out_steps[:, -1]) < TOL
    assert rel_l2(out_prefill, out_steps) < TOL
    # and the two caches are interchangeable: one more token on either gives the same result
    nxt = sh.embed(torch.tensor([[17]], device="cuda"))
    a, b = sh.forward(nxt, kv_a, [0], causal=True), sh.forward(nxt, kv_b, [0], causal=True)
    assert rel_l2(a, b) < TOL
    kv_a.close(); kv_b.close()


def test_decode_is_bit_reproducible_and_rows_are_independent(shard7b):
    sh = shard7b
    g = torch.Generator(device="cuda").manual_seed(2)
    ids = torch.randint(3, 32000, (1, 70), device="cuda", generator=g)
    x = sh.embed(ids)
    results = []
    for rep in range(2):
        kv = sh.new_kv(4, max_seq_len=128)
        xb = x.expand(4, -1, -1).contiguous()          # four identical rows
        sh.forward(xb, kv, [0, 1, 2, 3])
        step = sh.embed(torch.full((4, 1), 99, device="cuda"))
        tok, hid, margin = sh.forward_argmax(step, kv, [0, 1, 2, 3], want_hidden=True, want_margin=True)
        results.append((tok.clone(), hid.clone()))
        assert torch.equal(hid[0], hid[1]) and torch.equal(hid[0], hid[3])   # same input rows -> same bits
        assert len(set(tok.tolist())) == 1
        kv.close()
    assert torch.equal(results[0][0], results[1][0]) and torch.equal(results[0][1], results[1][1])
    # a single row through the 1-row (CUDA-core) kernels agrees with the 4-row (tensor-core) kernels
    kv = sh.new_kv(1, max_seq_len=128)
    sh.forward(x, kv, [0])
    tok1, hid1, _ = sh.forward_argmax(sh.embed(torch.full((1, 1), 99, device="cuda")), kv, [0], want_hidden=True)
    assert rel_l2(hid1[0], results[0][1][0]) < TOL
    kv.close()


def test_kv_limits_and_ragged_batches_are_rejected(shard7b):
    sh = shard7b
    kv = sh.new_kv(2, page_size=64, n_pages=3)
    x = torch.zeros(1, 128, 4096, device="cuda", dtype=torch.bfloat16)
    sh.forward(x, kv, [0])                                   # exactly two pages
    assert kv.get_seq_length(0) == 128
    sh.forward(x[:, :64].contiguous(), kv, [0])              # third and last page, filled to the brim
    assert kv.get_seq_length(0) == 192
    with pytest.raises(torch.cuda.OutOfMemoryError):
        sh.forward(x[:, :1].contiguous(), kv, [0])           # nothing left
    with pytest.raises(ValueError):                           # rows with different cached lengths (192 vs 0)
        sh.forward(torch.zeros(2, 1, 4096, device="cuda", dtype=torch.bfloat16), kv, [0, 1])
    with pytest.raises(ValueError):                           # the same slot twice in one batch
        sh.forward(torch.zeros(2, 1, 4096, device="cuda", dtype=torch.bfloat16), kv, [1, 1])
    with pytest.raises(ValueError):                           # empty sequence
        sh.forward(torch.zeros(1, 0, 4096, device="cuda", dtype=torch.bfloat16), kv, [1])
    with pytest.raises(ValueError):                           # more rows than the workspace was sized for
        sh.forward(torch.zeros(9, 1, 4096, device="cuda", dtype=torch.bfloat16), kv, list(range(9)))
    kv.reset(0)
    assert kv.get_seq_length(0) == 0
    sh.forward(x, kv, [1])                                    # the pages are back in the pool
    kv.close()
