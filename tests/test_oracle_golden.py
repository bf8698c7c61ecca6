"""The oracle restatement against the fixtures produced by the real reference chain
(oracle/make_golden.py).  CPU only."""
import pytest
import torch

from golden_utils import golden_names, golden_states, load_golden, oracle_replay, rel_l2
from oracle import synth


@pytest.mark.parametrize("name", golden_names())
def test_oracle_matches_reference_fixture(name):
    z, meta = load_golden(name)
    cfg = synth.PRESETS[meta["preset"]]
    # the weights the fixture was made from are the ones synth regenerates now
    import hashlib
    st = synth.layer_state(cfg, meta["seed"], 0)
    h = hashlib.sha256()
    for k in sorted(st):
        h.update(synth.tensor_digest(st[k]).encode())
    assert h.hexdigest()[:16] == meta["digests"]["block_0.pth"], "synthetic weight generator drifted"

    tokens, states, margins = oracle_replay(meta)
    assert tokens.tolist() == z["tokens"].tolist()
    gold = golden_states(z, len(states), len(meta["splits"]))
    worst = 0.0
    for p, per_pass in enumerate(states):
        assert len(per_pass) == len(gold[p])
        for i, h_ in enumerate(per_pass):
            assert h_.shape == gold[p][i].shape
            worst = max(worst, rel_l2(h_, gold[p][i]))
    # same op order, same machine class: fp32 results agree to rounding noise
    assert worst < 1e-5, worst
    assert torch.allclose(margins, torch.from_numpy(z["margins"]), atol=1e-4)


def test_prompt_has_no_special_ids():
    ids = synth.prompt_ids(synth.PRESETS["tiny"], 1, 4, 64)
    assert int(ids.min()) >= 3 and ids.dtype == torch.int64


def test_oracle_causal_flag_changes_prefill_only():
    """the reference is non-causal; the causal switch must differ on S>1 and
    agree on S=1 decode (no future positions exist)."""
    from oracle.llama_oracle import OracleWorker
    cfg = synth.PRESETS["tiny"]
    ids = synth.prompt_ids(cfg, 3, 1, 8)
    a = OracleWorker(cfg, 3, 0, 1, True, causal=False)  # one layer: deeper rows mix earlier rows
    b = OracleWorker(cfg, 3, 0, 1, True, causal=True)
    ha = a.pass_through_shard(a.receive_user_request(ids))["hidden_states"]
    hb = b.pass_through_shard(b.receive_user_request(ids))["hidden_states"]
    assert rel_l2(ha, hb) > 1e-3
    assert torch.equal(ha[:, -1], hb[:, -1])  # last row sees every key in both modes


def test_oracle_errors_match_reference_classes():
    from oracle.llama_oracle import OracleWorker
    cfg = synth.PRESETS["tiny"]
    with pytest.raises(ValueError):
        OracleWorker(cfg, 1, 3, 2, False)
    w = OracleWorker(cfg, 1, 1, 2, False)
    with pytest.raises(RuntimeError):
        w.receive_user_request(synth.prompt_ids(cfg, 1, 1, 4))
    with pytest.raises(RuntimeError):
        w.pass_through_shard({"hidden_states": torch.zeros(1, 1, 256), "batch_size": 1, "seq_len": 1})


def test_llama3_rope_rescaling_matches_transformers():
    """The oracle's and the product host code's inv_freq against the third-party module that carries the arithmetic
    (transformers LlamaRotaryEmbedding with rope_type llama3) — the dependency the reference calls at
    utils/node_worker.py:152."""
    from transformers import LlamaConfig
    from transformers.models.llama.modeling_llama import LlamaRotaryEmbedding

    from llm_sharding_b200.engine import rope_inv_freq as product_inv_freq
    from oracle.llama_oracle import rope_inv_freq
    for old_len in (64, 8192):
        sc = {"rope_type": "llama3", "factor": 8.0, "low_freq_factor": 1.0, "high_freq_factor": 4.0,
              "original_max_position_embeddings": old_len}
        cfg = LlamaConfig(hidden_size=512, num_attention_heads=4, rope_theta=500000.0, rope_scaling=dict(sc),
                          max_position_embeddings=4096)
        ref = LlamaRotaryEmbedding(cfg).inv_freq
        assert torch.equal(rope_inv_freq(128, 500000.0, sc), ref)
        assert torch.equal(product_inv_freq(128, 500000.0, sc), ref)
        assert not torch.equal(ref, rope_inv_freq(128, 500000.0, None))
    assert torch.equal(product_inv_freq(128, 10000.0, None), rope_inv_freq(128, 10000.0))
