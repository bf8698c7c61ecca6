"""Run the reference's own scripts and tools on top of this package, unchanged.

The reference imports its runtime as ``utils.node_worker`` / ``utils.shard_loader`` / ``utils.forwarding_utils`` /
``utils.config_sender`` (e.g. utils/node_profiler.py:8-10, start_node.py, send_config.py).  ``reference_modules()``
binds those names to this package's mirrors for the duration of a ``with`` block, so that a reference file — most
usefully its acceptance harness ``utils/node_profiler.py`` (latency sweeps, fits, assisted two-device profiling, the
four-node in-process ring) — can be loaded from a checkout of the reference and driven as it is:

    from llm_sharding_b200.ref_compat import load_reference_profiler
    NodeProfiler = load_reference_profiler("/path/to/llm-sharding")      # the reference's class, its code untouched
    NodeProfiler(shards, device="cuda:0", dtype=torch.float16).profile_compute_capability(max_layer_num=-1)

Nothing of the reference is copied or needed by the product path: this module only arranges ``sys.modules``.
"""
from __future__ import annotations

import contextlib
import importlib.util
import os
import sys
import types

_NAMES = ("node_worker", "shard_loader", "forwarding_utils", "config_sender")


@contextlib.contextmanager
def reference_modules():
    """``utils`` and ``utils.<mirror>`` resolve to this package inside the block; previous bindings are restored."""
    from . import config_sender, forwarding_utils, node_worker, shard_loader
    mine = {"node_worker": node_worker, "shard_loader": shard_loader, "forwarding_utils": forwarding_utils,
            "config_sender": config_sender}
    saved = {k: sys.modules.get(k) for k in ["utils"] + [f"utils.{n}" for n in _NAMES]}
    pkg = types.ModuleType("utils")
    pkg.__path__ = []   # a package without a directory: only the names bound below exist
    sys.modules["utils"] = pkg
    for n in _NAMES:
        sys.modules[f"utils.{n}"] = mine[n]
        setattr(pkg, n, mine[n])
    try:
        yield pkg
    finally:
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v


def load_reference_file(path: str, module_name: str):
    """Execute one reference source file with the bindings above; returns the module object."""
    if not os.path.isfile(path):
        raise FileNotFoundError(path)
    with reference_modules():
        spec = importlib.util.spec_from_file_location(module_name, path)
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
    return mod


def load_reference_profiler(reference_root: str):
    """The reference's ``NodeProfiler`` class (utils/node_profiler.py), bound to this package's ``NodeWorker``."""
    return load_reference_file(os.path.join(reference_root, "utils", "node_profiler.py"), "b200_ref_node_profiler").NodeProfiler
