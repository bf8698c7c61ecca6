"""-m gpu: where the fp16-operand design could bite (round-1 VERDICT, evidence item 10).  The library holds bf16 model weights
as fp16 and feeds fp16 activations to the tensor cores with saturating casts.  Pinned here, against the CPU oracle on one
shard of custom weights: (1) outlier weights up to |w| = 1e3 stay exact and parity holds; (2) weights far below fp16's
normal range (bf16 values of 1e-6 .. 1e-9, fp16-subnormal or flushed) cost at most their absolute rounding error;
(3) activations beyond 65504 saturate — the output is finite and reproducible, never inf / NaN — on the prefill, the
batched-decode and the single-row decode kernels alike."""
import pytest
import torch

from oracle import synth
from oracle.llama_oracle import KVCache, OracleShard, build_position_ids, rope_cos_sin

pytestmark = pytest.mark.gpu

CFG = synth.PRESETS["small"]   # H = 512, 4 heads, I = 1408


def rel_l2(a, b):
    return float((a.double() - b.double()).norm() / b.double().norm().clamp_min(1e-30))


def _model_cfg(n_layers):
    return {"hidden_size": CFG.hidden_size, "intermediate_size": CFG.intermediate_size, "num_attention_heads": CFG.num_attention_heads,
            "num_key_value_heads": CFG.num_key_value_heads, "head_dim": 128, "vocab_size": CFG.vocab_size,
            "rms_norm_eps": CFG.rms_norm_eps, "rope_theta": CFG.rope_theta, "num_hidden_layers": n_layers}


def _run_both(states, x, steps=0):
    """x [B,S,H] bf16-representable fp32 -> (gpu outputs, oracle outputs) for the prompt and `steps` single-token calls."""
    from llm_sharding_b200.engine import Shard
    B, S, H = x.shape
    sh = Shard(_model_cfg(len(states)), len(states), torch.device("cuda:0"), max_tokens=B * S, max_batch=B)
    for l, st in enumerate(states):
        sh.set_layer(l, st)
    kv = sh.new_kv(B, max_seq_len=S + steps + 8)
    oracle = OracleShard(states, CFG.num_attention_heads, CFG.num_key_value_heads, CFG.rms_norm_eps)
    cache = KVCache()
    outs_g, outs_o = [], []
    g = torch.Generator().manual_seed(5)
    chunks = [x] + [(torch.randn(B, 1, H, generator=g) * 0.5).bfloat16().float() for _ in range(steps)]
    for c in chunks:
        pos = build_position_ids(cache.get_seq_length(), c.shape[1], B)
        cos, sin = rope_cos_sin(pos, 128, CFG.rope_theta, torch.float32)
        with torch.inference_mode():
            outs_o.append(oracle.forward(c, cache, cos, sin))
        outs_g.append(sh.forward(c.cuda().bfloat16(), kv, list(range(B))).float().cpu())
    torch.cuda.synchronize()
    kv.close(); sh.close()
    return outs_g, outs_o


@pytest.mark.parametrize("B,S", [(1, 70), (6, 1), (1, 1)])   # tile GEMMs + tcgen05 attention / streaming GEMM / GEMV
def test_outlier_weights_up_to_1e3_keep_parity(B, S):
    g = torch.Generator().manual_seed(B * 10 + S)
    states = [synth.layer_state(CFG, 41, l) for l in range(2)]
    # a few large entries in the projections whose outputs are used linearly.  (Not q / k: logits of order 1e3 turn the
    # softmax into an arg-max over keys that the bf16 K cache — the model dtype, not the fp16 weight store — cannot rank
    # to 2e-2; not gate / up: their product would leave fp16's range, which the saturation test below covers.)
    for st in states:
        for k in ("self_attn.v_proj.weight", "self_attn.o_proj.weight", "mlp.down_proj.weight"):
            w = st[k].float()
            idx = torch.randint(0, w.numel(), (16,), generator=g)
            w.view(-1)[idx] = torch.tensor([1000.0, -1000.0, 512.0, -768.0] * 4)
            st[k] = w.bfloat16()
    x = (torch.randn(B, S, CFG.hidden_size, generator=g) * 0.5).bfloat16().float()
    outs_g, outs_o = _run_both(states, x, steps=2)
    for a, b in zip(outs_g, outs_o):
        assert torch.isfinite(a).all()
        assert rel_l2(a, b) <= 2e-2, rel_l2(a, b)


def test_weights_below_fp16_normal_range_cost_only_their_rounding_error():
    """lm-head-like GEMV with bf16 weights of 1e-6 .. 1e-9 (fp16 subnormals, spacing 6e-8): |error| <= K * 3e-8 * max|x|."""
    from llm_sharding_b200 import _lib
    lib = _lib.load()
    g = torch.Generator(device="cuda").manual_seed(3)
    N, K = 256, 512
    mag = 10.0 ** (-6 - 3 * torch.rand(N, K, device="cuda", generator=g))
    w_bf16 = (mag * torch.sign(torch.randn(N, K, device="cuda", generator=g))).bfloat16()
    w_held = w_bf16.float().clamp(-65504, 65504).half()          # what b200_shard_set_layer stores (round to nearest)
    x = torch.randn(1, K, device="cuda", generator=g)
    out = torch.empty(1, N, device="cuda")
    _lib.check(lib.b200_gemv_f16(x.data_ptr(), w_held.data_ptr(), 1, N, K, out.data_ptr(), torch.cuda.current_stream().cuda_stream))
    torch.cuda.synchronize()
    ref = x.double() @ w_bf16.double().t()
    assert float((out.double() - ref).abs().max()) <= K * 3.0e-8 * float(x.abs().max())


@pytest.mark.parametrize("B,S", [(1, 70), (6, 1), (1, 1)])
def test_activations_beyond_fp16_range_saturate_and_stay_finite(B, S):
    g = torch.Generator().manual_seed(7)
    states = [synth.layer_state(CFG, 42, l) for l in range(2)]
    for st in states:   # gate / up weights x 400: silu(g) * u reaches ~1e7, far beyond 65504
        st["mlp.gate_proj.weight"] = (st["mlp.gate_proj.weight"].float() * 400).bfloat16()
        st["mlp.up_proj.weight"] = (st["mlp.up_proj.weight"].float() * 400).bfloat16()
    x = (torch.randn(B, S, CFG.hidden_size, generator=g) * 0.5).bfloat16().float()
    a1, oracle = _run_both(states, x, steps=1)
    a2, _ = _run_both(states, x, steps=1)
    assert float(oracle[0].abs().max()) > 65504 / 8     # the fp32 reference really does leave fp16's range inside the MLP
    for u, v in zip(a1, a2):
        assert torch.isfinite(u).all()                   # saturated, not inf / NaN
        assert torch.equal(u, v)                         # and reproducible
