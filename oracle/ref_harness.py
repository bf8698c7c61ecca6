"""TEST INFRASTRUCTURE — drives the REAL reference (``/root/reference``) in-process.

In the build container the reference is imported from ``/root/reference``: that is how
``tests/golden/*.npz`` are produced (``oracle/make_golden.py``) and how the restatement in
``llama_oracle.py`` is validated.  The GPU box has no ``/root/reference``; there the same six
files are found under ``baseline/_ref/utils/`` — a git-ignored, verbatim copy made by
``__graft_entry__.build()`` (an output directory like a compiled ``.so``: never committed,
never edited) — and are used ONLY by ``bench.py``'s CPU-baseline legs
(``oracle/ref_chain_bench.py``), never by the ``-m gpu`` tests or ``smoke()``.
The reference's files are imported unmodified; two
compat points are patched from outside because this image has transformers 5.5.0 while
the reference pins 4.53.2:

1. ``LlamaDecoderLayer.forward`` is wrapped to accept ``past_key_value=`` (the 4.53.2
   keyword used at ``utils/shard_loader.py:67-73``) and to return a tuple (``outputs[0]``
   at ``:74``).
2. ``NodeWorker.clear_KV_cache`` is called under ``torch.inference_mode()`` (5.5.0's
   ``DynamicCache.reset`` zeroes inference tensors in place).

The chain pattern is the reference's own in-process one
(``utils/node_profiler.py:1174-1236``): N ``NodeWorker``s on loopback ZMQ ports forming a
ring, ``pass_through_shard`` -> ``transfer_data`` -> next ``receive_data`` ...
"""
from __future__ import annotations

import contextlib
import io
import os
import sys

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
REFERENCE_ROOT = "/root/reference" if os.path.isfile("/root/reference/utils/node_worker.py") else os.path.join(os.path.dirname(_HERE), "baseline", "_ref")
_shim_applied = False


def reference_available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_ROOT, "utils", "node_worker.py"))


def apply_shim() -> None:
    global _shim_applied
    if _shim_applied:
        return
    from transformers.models.llama import modeling_llama

    orig = modeling_llama.LlamaDecoderLayer.forward

    def forward_4_53(self, *args, past_key_value=None, **kwargs):
        if past_key_value is not None:
            kwargs["past_key_values"] = past_key_value
        out = orig(self, *args, **kwargs)
        return out if isinstance(out, tuple) else (out,)

    modeling_llama.LlamaDecoderLayer.forward = forward_4_53
    _shim_applied = True


def import_reference():
    """Returns the reference's ``utils.node_worker`` module (unmodified source)."""
    if not reference_available():
        raise RuntimeError("the reference tree is not present on this machine")
    apply_shim()
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    import utils.node_worker as nw  # noqa: the reference's package is literally called `utils`
    return nw


class ReferenceChain:
    """N reference ``NodeWorker``s in one process, ring-connected over loopback ZMQ."""

    def __init__(self, shards_path: str, splits: list[tuple[int, int]], dtype=torch.float32,
                 base_port: int = 41800, workdir: str | None = None, quiet: bool = True):
        nw = import_reference()
        self.nw = nw
        self.quiet = quiet
        self._cwd = os.getcwd()
        if workdir is not None:  # Communicator writes results/*.pt relative to cwd (node_worker.py:45,57)
            os.makedirs(workdir, exist_ok=True)
            os.chdir(workdir)
        n = len(splits)
        self.workers = []
        with self._mute():
            for i, (s, e) in enumerate(splits):
                w = nw.NodeWorker(
                    src_addr=f"tcp://127.0.0.1:{base_port + i}",
                    dst_addr=f"tcp://127.0.0.1:{base_port + (i + 1) % n}",
                    can_receive_user_request=(i == 0),
                    shards_path=shards_path, device="cpu", dtype=dtype)
                w.load_shards(s, e)
                self.workers.append(w)
        self.tail_logits = None
        self.workers[-1].lm_head.register_forward_hook(self._grab_logits)

    def _grab_logits(self, mod, inp, out):
        self.tail_logits = out[:, -1, :].detach().clone()

    def _mute(self):
        return contextlib.redirect_stdout(io.StringIO()) if self.quiet else contextlib.nullcontext()

    def close(self):
        for w in self.workers:
            w.communicator.recv_socket.close(0)
            w.communicator.send_socket.close(0)
        os.chdir(self._cwd)

    def clear(self):
        with torch.inference_mode(), self._mute():
            for w in self.workers:
                w.clear_KV_cache()

    def one_pass(self, msg, over_sockets: bool = True):
        """msg through every worker; returns (token ids [B], [hidden states leaving each
        non-tail shard], margin [B])."""
        states = []
        data = msg
        with self._mute():
            for i, w in enumerate(self.workers):
                out = w.pass_through_shard(data)
                if isinstance(out, dict):
                    states.append(out["hidden_states"].detach().clone().float())
                if over_sockets:
                    w.communicator.transfer_data(out)
                    data = self.workers[(i + 1) % len(self.workers)].communicator.receive_data()
                else:
                    data = out
        top2 = torch.topk(self.tail_logits.float(), 2, dim=-1).values
        return data, states, (top2[:, 0] - top2[:, 1])

    def generate(self, input_ids: torch.Tensor, max_new_tokens: int):
        """The reference's batch-1 loop (receive_user_request -> passes -> receive_next_token)."""
        assert input_ids.shape[0] == 1, "the reference's generation loop is batch-1 only (SURVEY.md F5)"
        head = self.workers[0]
        with self._mute():
            msg = head.receive_user_request(input_ids=input_ids)
        tokens, states, margins = [], [], []
        while True:
            tok, st, mg = self.one_pass(msg)
            tokens.append(tok.clone())
            states.append(st)
            margins.append(mg)
            with self._mute():
                reached_end, msg = head.receive_next_token(tok, max_new_tokens)
            if reached_end:
                break
        return torch.stack(tokens), states, torch.stack(margins)

    def batched_passes(self, input_ids: torch.Tensor, n_decode: int):
        """pass_through_shard is batch-generic (node_worker.py:262-264) even though the loop
        is not: drive a [B,S] prefill and ``n_decode`` [B,1] steps through it by hand."""
        head = self.workers[0]
        B, S = input_ids.shape
        with torch.inference_mode():
            hidden = head.embed_tokens(input_ids)
        msg = {"hidden_states": hidden, "batch_size": B, "seq_len": S}
        tokens, states, margins = [], [], []
        for _ in range(n_decode + 1):
            tok, st, mg = self.one_pass(msg, over_sockets=False)
            tokens.append(tok.clone())
            states.append(st)
            margins.append(mg)
            with torch.inference_mode():
                hidden = head.embed_tokens(tok.unsqueeze(-1))
            msg = {"hidden_states": hidden, "batch_size": B, "seq_len": 1}
        return torch.stack(tokens), states, torch.stack(margins)
