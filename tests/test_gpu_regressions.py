"""-m gpu: regression tests for faults found in review (round-1 ADVICE.md): stale CUDA graphs after a KV cache was
destroyed, unbounded graph cache, the hop using its own ring geometry for the peer's mailbox, the device-side KV
error word that nobody read, pickled messages on a network socket."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def rel_l2(a, b):
    return float((a.double() - b.double()).norm() / b.double().norm().clamp_min(1e-30))


@pytest.fixture(scope="module")
def small_shard():
    from llm_sharding_b200.engine import Shard
    from oracle import synth
    cfg = synth.PRESETS["tiny"]
    mc = {"hidden_size": cfg.hidden_size, "intermediate_size": cfg.intermediate_size,
          "num_attention_heads": cfg.num_attention_heads, "num_key_value_heads": cfg.num_key_value_heads,
          "head_dim": 128, "vocab_size": cfg.vocab_size, "rms_norm_eps": cfg.rms_norm_eps, "rope_theta": cfg.rope_theta,
          "num_hidden_layers": 2}
    sh = Shard(mc, 2, torch.device("cuda:0"), max_tokens=128, max_batch=2)
    for l in range(2):
        sh.set_layer(l, synth.layer_state(cfg, 28, l))
    yield sh
    sh.close()


def test_graphs_do_not_outlive_the_kv_cache_they_were_captured_on(small_shard):
    """kv_destroy followed by kv_create often returns the same host address; a graph keyed on that address would replay
    with the freed pools baked in.  The key is a generation number now and kv_destroy drops the graphs."""
    sh = small_shard
    sh.set_graphs(True)
    g = torch.Generator(device="cuda").manual_seed(3)
    prompt = (torch.randn(1, 40, sh.hidden_size, device="cuda", generator=g) * 0.5).bfloat16()
    step = (torch.randn(1, 1, sh.hidden_size, device="cuda", generator=g) * 0.5).bfloat16()
    hin, hout = step.clone(), torch.empty_like(step)   # persistent pointers: the graph key repeats exactly
    expected = None
    for rep in range(6):
        kv = sh.new_kv(1, max_seq_len=128)
        sh.forward(prompt, kv, [0])
        out = sh.forward(hin, kv, [0], out=hout).clone()
        torch.cuda.synchronize()
        if expected is None:
            expected = out
        assert torch.equal(out, expected), rep
        kv.close()
        # occupy the memory the pools just gave back, so a stale graph would read garbage rather than lucky leftovers
        junk = torch.full((1 << 20,), float("nan"), device="cuda")
        del junk
    sh.set_graphs(False)


def test_graph_cache_is_bounded(small_shard):
    sh = small_shard
    sh.set_graphs(True)
    kv = sh.new_kv(1, max_seq_len=256)
    g = torch.Generator(device="cuda").manual_seed(4)
    x = (torch.randn(1, 1, sh.hidden_size, device="cuda", generator=g) * 0.5).bfloat16()
    outs = [torch.empty_like(x) for _ in range(100)]   # a fresh output pointer per step = a fresh graph per step
    for o in outs:
        sh.forward(x, kv, [0], out=o)
    torch.cuda.synchronize()
    assert kv.get_seq_length(0) == 100
    assert all(torch.isfinite(o.float()).all() for o in outs)
    # replay through an evicted key: must be re-captured, not crash
    sh.forward(x, kv, [0], out=outs[0])
    torch.cuda.synchronize()
    kv.close()
    sh.set_graphs(False)


def test_kv_error_word_is_readable(small_shard):
    from llm_sharding_b200 import _lib
    sh = small_shard
    kv = sh.new_kv(1, max_seq_len=64)
    assert sh.lib.b200_kv_error(kv.handle) == 0
    kv.close()


def test_hop_uses_the_peers_ring_geometry():
    """Sender ring (2 slots of 4 KiB) != receiver ring (5 slots of 64 KiB): payload and flag must land where the
    RECEIVER looks for them (the handle carries its geometry)."""
    from llm_sharding_b200.chain import P2PHop
    rx = P2PHop(torch.device("cuda:0"), 5, 64 * 1024)
    tx = P2PHop(torch.device("cuda:0"), 2, 4096)
    tx.connect(rx.export(), 0, same_process=True)
    g = torch.Generator(device="cuda").manual_seed(9)
    for seq in range(7):
        payload = torch.randint(0, 255, (40_000,), device="cuda", dtype=torch.uint8, generator=g)  # > tx's own slot size
        out = torch.zeros_like(payload)
        tx.send(payload, seq)
        rx.recv(out, seq, max_wait_us=1_000_000)
        torch.cuda.synchronize()
        assert torch.equal(out, payload), seq
    assert rx.error() == 0
    with pytest.raises(ValueError):
        tx.send(torch.zeros(64 * 1024 + 256, dtype=torch.uint8, device="cuda"), 7)
    bogus = bytearray(tx.export())
    bogus[72:80] = b"\0" * 8   # geometry word wiped
    with pytest.raises(ValueError):
        tx.connect(bytes(bogus), 0, same_process=True)
    tx.close(); rx.close()
