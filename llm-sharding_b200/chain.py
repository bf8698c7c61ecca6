"""Chain runtime: the reference's in-process N-worker ring (utils/node_profiler.py:1174-1236) and
the transports that replace its ZeroMQ hop on one NVSwitch box.

* ``LoopbackCommunicator`` — ``Communicator``-shaped (transfer_data / receive_data / change_*),
  in-process queues; device tensors are handed over by reference (same GPU) — used when several
  shards share one GPU and by the single-GPU benchmark.
* ``P2PHop`` — the NVLink hop: ``b200_hop_*`` mailbox + in-kernel 128-bit peer stores + flag.
* ``generate`` — drives a list of ``NodeWorker``s exactly like ``go_through_every_shards``.
"""
from __future__ import annotations

import collections
import ctypes as C

import torch
import zmq

from . import _lib

_QUEUES: dict[str, collections.deque] = {}


def _norm(addr: str) -> str:
    return addr.replace("tcp://*:", "tcp://127.0.0.1:")


class LoopbackCommunicator:
    def __init__(self, src_addr: str, dst_addr: str):
        self.src_addr, self.dst_addr = src_addr, dst_addr
        self.actual_src_addr = _norm(src_addr)
        _QUEUES.setdefault(self.actual_src_addr, collections.deque())

    def change_src_addr(self, new_src_addr: str) -> str:
        self.src_addr, self.actual_src_addr = new_src_addr, _norm(new_src_addr)
        _QUEUES.setdefault(self.actual_src_addr, collections.deque())
        return self.src_addr

    def change_dst_addr(self, new_dst_addr: str) -> str:
        self.dst_addr = new_dst_addr
        return self.dst_addr

    def transfer_data(self, data, data_path: str = "", keep_data: bool = False):
        _QUEUES.setdefault(_norm(self.dst_addr), collections.deque()).append(data)
        return None

    def receive_data(self, no_block: bool = False, data_path: str = "", keep_data: bool = False):
        q = _QUEUES[self.actual_src_addr]
        if not q:
            raise zmq.Again()
        return q.popleft()

    def close(self) -> None:
        _QUEUES.pop(self.actual_src_addr, None)


class P2PHop:
    """One receive mailbox on this GPU and a mapping of the next stage's mailbox."""

    def __init__(self, device: torch.device, n_ring_slots: int, slot_bytes: int):
        _lib.require_cuda()
        self.lib = _lib.load()
        self.device = torch.device(device)
        self.dev_index = self.device.index if self.device.index is not None else torch.cuda.current_device()
        h = C.c_void_p()
        _lib.check(self.lib.b200_hop_create(self.dev_index, n_ring_slots, slot_bytes, C.byref(h)))
        self.handle = h
        self.n_ring_slots, self.slot_bytes = n_ring_slots, slot_bytes

    def export(self) -> bytes:
        buf = C.create_string_buffer(_lib.HOP_HANDLE_BYTES)
        _lib.check(self.lib.b200_hop_export(self.handle, buf))
        return buf.raw

    def connect(self, peer_handle: bytes, peer_device: int, same_process: bool) -> None:
        _lib.check(self.lib.b200_hop_connect(self.handle, peer_handle, peer_device, int(same_process)))

    def send(self, payload: torch.Tensor, seq: int) -> None:
        t = payload.contiguous()
        _lib.check(self.lib.b200_hop_send(self.handle, t.data_ptr(), t.numel() * t.element_size(), seq,
                                          _lib.stream_ptr(self.device)))

    def recv(self, out: torch.Tensor, seq: int, max_wait_us: int = 2_000_000) -> torch.Tensor:
        _lib.check(self.lib.b200_hop_recv(self.handle, out.data_ptr(), out.numel() * out.element_size(), seq,
                                          max_wait_us, _lib.stream_ptr(self.device)))
        return out

    def error(self) -> int:
        return self.lib.b200_hop_error(self.handle)

    def close(self):
        if getattr(self, "handle", None):
            self.lib.b200_hop_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def generate(workers, input_ids: torch.Tensor, max_new_tokens: int, record_states: bool = False,
             teacher_tokens: torch.Tensor | None = None):
    """The reference's in-process loop: request -> every shard in order -> next token -> re-embed.
    Returns ``(tokens [n, B] int64 (cpu), states per pass (list of lists, fp32 cpu))``.
    ``teacher_tokens`` [n, B] forces the token fed back at each step (used by parity tests to keep
    comparing after a near-tie flips an arg-max)."""
    head = workers[0]
    msg = head.receive_user_request(input_ids=input_ids)
    tokens, states = [], []
    n = 0
    while True:
        per_pass = []
        data = msg
        for w in workers:
            data = w.pass_through_shard(data)
            if isinstance(data, dict) and record_states:
                per_pass.append(data["hidden_states"].float().cpu())
        tokens.append(data.cpu())
        states.append(per_pass)
        fed = data if teacher_tokens is None else teacher_tokens[n].to(data.device)
        n += 1
        head.generated_ids.append(fed.unsqueeze(-1))
        if len(head.generated_ids) > max_new_tokens:
            break
        msg = {"hidden_states": head.embed_tokens(fed.reshape(-1, 1)), "batch_size": head.batch_size, "seq_len": 1}
    return torch.stack(tokens), states
