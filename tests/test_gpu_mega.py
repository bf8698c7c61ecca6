"""-m gpu: the opt-in persistent decode kernel (csrc/decode_mega.cu, B200SHARD_MEGA=1) against the same oracle and golden
fixtures as the default path.  The switch is read once per process, so the checks run in a child interpreter: the
single-row parity cases of test_gpu_parity.py / test_gpu_prefill_parity.py (tiny and Llama-2-7B layer dimensions, GQA,
page-crossing contexts) with the kernel on, and a direct count of launches that proves the kernel was the one that ran."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(args, extra_env):
    env = dict(os.environ, **extra_env)
    return subprocess.run([sys.executable] + args, cwd=ROOT, env=env, capture_output=True, text=True, timeout=900)


def test_parity_cases_with_the_persistent_kernel():
    r = _run(["-m", "pytest", "tests/test_gpu_parity.py", "tests/test_gpu_prefill_parity.py", "-m", "gpu", "-x", "-q", "-p", "no:cacheprovider",
              "-k", "fixture or tiny-101 or tiny-105 or gqa-108 or 301 or 305"], {"B200SHARD_MEGA": "1"})
    tail = (r.stdout + r.stderr)[-1500:]
    assert r.returncode == 0, tail
    assert " passed" in r.stdout and "failed" not in r.stdout, tail


COUNT_SNIPPET = r"""
import sys, tempfile, torch
sys.path.insert(0, "tests")
from gpu_utils import build_chain, close_chain
from oracle import synth
root = tempfile.mkdtemp()
workers = build_chain(root, "tiny", 101, [(0, 4)], max_batch=1, max_tokens=8)
from llm_sharding_b200.chain import generate
ids = synth.prompt_ids(synth.PRESETS["tiny"], 101, 1, 1)
before = workers[0].shard.engine.counters()["launches"]
generate(workers, ids, 3)
print("LAUNCHES", workers[0].shard.engine.counters()["launches"] - before)
close_chain(workers)
"""


def test_the_persistent_kernel_is_the_one_that_runs():
    """Four layers, three single-token passes: 5 launches per layer without the kernel, 1 per pass for all layers with it."""
    on = _run(["-c", COUNT_SNIPPET], {"B200SHARD_MEGA": "1"})
    off = _run(["-c", COUNT_SNIPPET], {"B200SHARD_MEGA": "0"})
    assert on.returncode == 0 and off.returncode == 0, (on.stderr[-800:], off.stderr[-800:])
    n_on = int([l for l in on.stdout.splitlines() if l.startswith("LAUNCHES")][0].split()[1])
    n_off = int([l for l in off.stdout.splitlines() if l.startswith("LAUNCHES")][0].split()[1])
    assert n_off - n_on == 3 * (4 * 5 - 1), (n_on, n_off)
