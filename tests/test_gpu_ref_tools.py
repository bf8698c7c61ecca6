"""-m gpu: the reference's OWN acceptance harness, `utils/node_profiler.py`, loaded unchanged from a checkout of the
reference and run on this package's NodeWorker (round-1 VERDICT 'missing' #5): the latency sweeps
with its timing protocol and fits, the assisted two-device mode, and its four-node in-process ring over ZeroMQ.

The reference's source is needed at test time, so these tests run where a checkout exists (/root/reference in the build
container, or the git-ignored copy `baseline/_ref` that travels to the GPU box) and are skipped elsewhere.  matplotlib
(absent from this image) is replaced by a do-nothing stand-in: only the PNG plots are lost."""
import contextlib
import io
import os
import re
import sys
import threading
import types

import pytest
import torch

from gpu_utils import shard_dir
from oracle import synth

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_ROOTS = ["/root/reference", os.path.join(ROOT, "baseline", "_ref")]
REF_ROOT = next((r for r in REF_ROOTS if os.path.isfile(os.path.join(r, "utils", "node_profiler.py"))), None)
needs_reference = pytest.mark.skipif(REF_ROOT is None, reason="no checkout of the reference on this machine")


class _Nothing(types.ModuleType):
    def __getattr__(self, name):
        return lambda *a, **k: None


@pytest.fixture()
def ref_profiler_cls(monkeypatch, tmp_path):
    if "matplotlib" not in sys.modules:
        try:
            import matplotlib  # noqa: F401
        except ImportError:
            mpl = _Nothing("matplotlib")
            mpl.pyplot = _Nothing("matplotlib.pyplot")
            monkeypatch.setitem(sys.modules, "matplotlib", mpl)
            monkeypatch.setitem(sys.modules, "matplotlib.pyplot", mpl.pyplot)
    monkeypatch.chdir(tmp_path)   # the profiler writes results/profiling/* relative to the working directory
    from llm_sharding_b200.ref_compat import load_reference_profiler
    return load_reference_profiler(REF_ROOT)


def _quick(prof):
    prof.PROFILE_INTERVAL_SLEEP_TIME = 0
    prof.PROFILE_REPEAT_NUM = 2
    prof.PROFILE_PREFILL_INPUT_TOKEN_LENGTHS = [8, 16, 32, 64]
    prof.PROFILE_DECODE_OUTPUT_TOKEN_LENGTHS = [4, 8]
    return prof


@needs_reference
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
def test_reference_profiler_runs_unchanged_on_the_b200_worker(ref_profiler_cls, tmp_path, dtype):
    d = shard_dir(str(tmp_path), "tiny", 28)
    prof = _quick(ref_profiler_cls(d, device="cuda:0", dtype=dtype))
    out = io.StringIO()
    with contextlib.redirect_stdout(out):
        assert prof.profile_max_layer_num() == 4
        prof.profile_compute_capability(max_layer_num=-1)        # prefill sweep, decode sweep, fits, similarity report
        prof.profile_cold_start_latency(max_layer_num=-1)
    text = out.getvalue()
    assert "[ERROR]" not in text
    assert os.path.isdir(tmp_path / "results")                    # its reports were written where it writes them


@needs_reference
def test_reference_profiler_assisted_mode_two_workers(ref_profiler_cls, tmp_path):
    """Target (layers 0..1) and assistor (layers 2..3 + embedding + lm_head) exactly as profiling.py /
    assist_profiling.py start them, here as two threads on one GPU talking over loopback ZeroMQ."""
    d = shard_dir(str(tmp_path), "tiny", 28)
    target = _quick(ref_profiler_cls(d, device="cuda:0", dtype=torch.float16))
    assistor = _quick(ref_profiler_cls(d, device="cuda:0", dtype=torch.float16))
    errors = []

    def run(fn, *a, **k):
        try:
            with contextlib.redirect_stdout(io.StringIO()):
                fn(*a, **k)
        except BaseException as e:   # noqa: BLE001 - reported by the main thread
            errors.append(repr(e))

    t = threading.Thread(target=run, args=(target.profile_compute_capability,),
                         kwargs=dict(max_layer_num=3, assisted=True, src_addr="tcp://*:40810", dst_addr="tcp://127.0.0.1:40811"))
    a = threading.Thread(target=run, args=(assistor.assist_profile_compute_capability,),
                         kwargs=dict(target_max_layer_num=3, src_addr="tcp://*:40811", dst_addr="tcp://127.0.0.1:40810"))
    t.start(); a.start()
    t.join(180); a.join(180)
    assert not t.is_alive() and not a.is_alive(), "assisted profiling did not finish"
    assert not errors, errors


@needs_reference
def test_reference_four_node_ring_produces_the_oracle_tokens(ref_profiler_cls, tmp_path):
    """go_through_every_shards (utils/node_profiler.py:1174-1236): four NodeWorkers, 7 layers each, ZeroMQ between them.
    It prints the tokens; they must be the CPU oracle's for the same prompt."""
    from oracle.llama_oracle import OracleWorker, run_chain
    from transformers import AutoTokenizer
    cfg = synth.PRESETS["l28"]
    d = shard_dir(str(tmp_path), "l28", 3)
    ids = AutoTokenizer.from_pretrained(d)("Why the sky blue", return_tensors="pt")["input_ids"]
    n = 5
    o_tok, _, o_margin = run_chain([OracleWorker(cfg, 3, s, e, i == 0) for i, (s, e) in enumerate([(0, 7), (7, 14), (14, 21), (21, 28)])],
                                   ids, n, record=False)
    prof = ref_profiler_cls(d, device="cuda:0", dtype=torch.bfloat16)
    out = io.StringIO()
    with contextlib.redirect_stdout(out):
        prof.go_through_every_shards(out_token_num=n)
    printed = re.findall(r"'t(\d+)'|'(<unk>|<s>|</s>)'", out.getvalue())
    special = {"<unk>": 0, "<s>": 1, "</s>": 2}
    got = [int(a) if a else special[b] for a, b in printed][:n]
    assert len(got) == n, out.getvalue()[-400:]
    for p in range(n):
        if float(o_margin[p, 0]) >= 0.02:
            assert got[p] == int(o_tok[p, 0]), (p, got, o_tok[:, 0].tolist(), o_margin[:, 0].tolist())
        if got[p] != int(o_tok[p, 0]):
            break   # a near-tie flipped: nothing after it is comparable
