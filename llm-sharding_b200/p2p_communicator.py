"""``P2PCommunicator``: the reference's ``Communicator`` interface (utils/node_worker.py:13-67 —
``transfer_data`` / ``receive_data(no_block)`` / ``change_src_addr`` / ``change_dst_addr``) with the NVLink mailbox
underneath, for ``NodeController`` processes that sit on one NVSwitch box (one per GPU, started like ``run_this.sh``).

What travels where
* Data path: every one of the reference's message kinds — ``input_token_info`` {hidden_states, batch_size, seq_len},
  ``next_state_info`` {hidden_states, cos, sin (+ pos_offset, slot)}, the next-token ``Tensor`` and the
  ``clear_KV_cache`` command — goes through ``b200_hop_send_msg``: a 256-byte header (kind, dtype, shape, batch_size,
  seq_len, pos_offset, KV slots) and the tensor are stored by ONE kernel straight into the next GPU's mailbox slot,
  followed by a release of the slot's sequence flag.  No ``.cpu()``, no ``torch.save``, no file, no pickle.
* Control path: the ZeroMQ PUSH/PULL pair of the reference, on the same ``src_addr`` / ``dst_addr``, carries only small
  JSON objects: the CUDA-IPC handle of a node's mailbox travels forward around the ring until it reaches the node whose
  ``dst_addr`` it belongs to (PUSH/PULL is one-way, so the successor cannot simply answer its predecessor).

``receive_data(no_block=True)`` raises ``zmq.Again`` when nothing has arrived, exactly like the reference's socket, so
``NodeController.step`` / ``run_worker_loop`` work unchanged.  Messages sent before the successor's handle is known are
queued (references held) and flushed on the next call, so a ring that is stepped by one thread cannot deadlock.
"""
from __future__ import annotations

import base64
import ctypes as C
import json
import os
import struct
import time

import torch
import zmq

from . import _lib
from .chain import P2PHop, tensor_view

KIND_TOKEN, KIND_INPUT, KIND_STATE, KIND_JSON = 0, 1, 2, 3
_MAGIC = b"B2HP"
_HDR = struct.Struct("<4sIII4qiiqqi")      # magic, kind, dtype, ndim, shape[4], batch_size, seq_len, pos_offset, payload bytes, n_slot
_DTYPES = [torch.bfloat16, torch.float32, torch.float16, torch.int64, torch.uint8]
_MAX_SLOTS_IN_HEADER = (_lib.HOP_HEADER_BYTES - _HDR.size) // 2


def _norm(addr: str) -> str:
    return addr.replace("tcp://*:", "tcp://127.0.0.1:").replace("tcp://localhost:", "tcp://127.0.0.1:").replace("tcp://0.0.0.0:", "tcp://127.0.0.1:")


class P2PCommunicator:
    def __init__(self, src_addr: str, dst_addr: str, device="cuda:0", slot_bytes: int = 4096 * 4096 * 2 + 4096,
                 n_ring_slots: int = 4, connect_timeout_s: float = 60.0):
        _lib.require_cuda()
        self.lib = _lib.load()
        self.device = torch.device(device)
        self.src_addr, self.dst_addr = src_addr, dst_addr
        self.connect_timeout_s = connect_timeout_s
        self.hop = P2PHop(self.device, n_ring_slots, slot_bytes + _lib.HOP_HEADER_BYTES)
        self.seq_in = 0
        self.seq_out = 0
        self.connected = False
        self._queue = []          # messages waiting for the successor's handle: (header bytes, payload tensor or None)
        self._inbox = []          # user requests addressed to this node (entry -> head, utils/node_worker.py:476-491)
        self._hdr_buf = C.create_string_buffer(_lib.HOP_HEADER_BYTES)
        # control sockets, same shape as the reference's (utils/node_worker.py:20-29)
        self._rctx = zmq.Context()
        self.recv_socket = self._rctx.socket(zmq.PULL)
        self.recv_socket.bind(self.src_addr)
        self.actual_src_addr = self.recv_socket.getsockopt_string(zmq.LAST_ENDPOINT)
        self._sctx = zmq.Context()
        self.send_socket = self._sctx.socket(zmq.PUSH)
        self.send_socket.connect(self.dst_addr)
        self._announce()

    # ------------------------------------------------------------------ control plane (JSON over ZeroMQ)
    def _my_names(self):
        return {_norm(self.src_addr), _norm(self.actual_src_addr)}

    @staticmethod
    def _ctl_send(sock, msg: dict, blob: bytes | None = None) -> None:
        frames = [json.dumps(msg).encode()] + ([blob] if blob is not None else [])
        sock.send_multipart(frames)

    def _announce(self, ttl: int = 64) -> None:
        self._ctl_send(self.send_socket, {"p2p": "handle", "owner": sorted(self._my_names()), "pid": os.getpid(),
                                          "device": self.hop.dev_index, "ttl": ttl,
                                          "handle": base64.b64encode(self.hop.export()).decode()})

    def _pump_control(self) -> None:
        while True:
            try:
                frames = self.recv_socket.recv_multipart(flags=zmq.NOBLOCK)
            except zmq.Again:
                return
            try:
                msg = json.loads(frames[0].decode())
            except (UnicodeDecodeError, ValueError):
                continue   # not ours (never unpickled, never executed)
            if not isinstance(msg, dict) or "p2p" not in msg:
                continue
            if msg["p2p"] == "request":   # an entry node's embedded prompt for the head of the chain (raw bytes, no pickle)
                if msg["want"] in self._my_names() and len(frames) == 2:
                    t = torch.frombuffer(bytearray(frames[1]), dtype=_DTYPES[msg["dtype"]]).view(msg["shape"]).to(self.device)
                    self._inbox.append({"hidden_states": t, "batch_size": msg["batch_size"], "seq_len": msg["seq_len"]})
                continue
            if msg["p2p"] == "handle":
                if _norm(self.dst_addr) in msg["owner"]:
                    if not self.connected:
                        self.hop.connect(base64.b64decode(msg["handle"]), int(msg["device"]), same_process=(msg["pid"] == os.getpid()))
                        self.connected = True
                        self._flush()
                elif msg.get("ttl", 0) > 1 and not (set(msg["owner"]) & self._my_names()):
                    self._ctl_send(self.send_socket, dict(msg, ttl=msg["ttl"] - 1))     # not for me: pass it on around the ring
            elif msg["p2p"] == "hello":   # somebody re-pointed its dst_addr and asks the owner of that address to announce itself
                if msg["want"] in self._my_names():
                    self._announce()
                elif msg.get("ttl", 0) > 1:
                    self._ctl_send(self.send_socket, dict(msg, ttl=msg["ttl"] - 1))

    def change_src_addr(self, new_src_addr: str) -> str:
        self.recv_socket.unbind(self.actual_src_addr)
        self.recv_socket.bind(new_src_addr)
        self.src_addr = new_src_addr
        self.actual_src_addr = self.recv_socket.getsockopt_string(zmq.LAST_ENDPOINT)
        self._announce()
        return self.src_addr

    def change_dst_addr(self, new_dst_addr: str) -> str:
        self.send_socket.disconnect(self.dst_addr)
        self.send_socket.connect(new_dst_addr)
        self.dst_addr = new_dst_addr
        self.connected = False
        self.seq_out = 0   # a fresh successor counts from zero
        self._ctl_send(self.send_socket, {"p2p": "hello", "want": _norm(new_dst_addr), "ttl": 64})
        self._announce()
        return self.dst_addr

    # ------------------------------------------------------------------ data plane (mailbox)
    @staticmethod
    def _header(kind, t: torch.Tensor | None, batch_size=0, seq_len=0, pos_offset=-1, slots=None, payload_bytes=0) -> bytes:
        shape = list(t.shape) if t is not None else []
        if len(shape) > 4:
            raise ValueError("[ERROR] tensors of more than 4 dimensions do not fit the hop header")
        dt = _DTYPES.index(t.dtype) if t is not None else 4
        slots = list(slots or [])
        if len(slots) > _MAX_SLOTS_IN_HEADER:
            raise ValueError("[ERROR] too many KV slots for one hop header")
        head = _HDR.pack(_MAGIC, kind, dt, len(shape), *(shape + [0] * (4 - len(shape))), int(batch_size), int(seq_len),
                         int(pos_offset), int(payload_bytes), len(slots))
        return head + struct.pack(f"<{len(slots)}h", *slots)

    def _encode(self, data):
        if isinstance(data, torch.Tensor):
            t = data.detach().to(self.device).contiguous()
            return self._header(KIND_TOKEN, t, payload_bytes=t.numel() * t.element_size()), t
        if isinstance(data, dict) and "hidden_states" in data:
            t = data["hidden_states"].detach().to(self.device).contiguous()
            kind = KIND_INPUT if ("batch_size" in data and "seq_len" in data) else KIND_STATE
            return self._header(kind, t, data.get("batch_size", 0), data.get("seq_len", 0), data.get("pos_offset", -1),
                                data.get("slot"), t.numel() * t.element_size()), t
        raw = json.dumps(data).encode()   # the clear_KV_cache command and other small control dicts
        t = torch.frombuffer(bytearray(raw + b"\0" * (-len(raw) % 16)), dtype=torch.uint8).to(self.device)
        return self._header(KIND_JSON, None, payload_bytes=len(raw)), t

    def _send_now(self, header: bytes, t: torch.Tensor | None) -> None:
        nbytes = 0 if t is None else t.numel() * t.element_size()
        if t is not None and nbytes % 16:   # the copy kernel moves 16-byte words: give it a padded source
            pad = torch.zeros((nbytes + 15) // 16 * 16, dtype=torch.uint8, device=self.device)
            pad[:nbytes] = t.view(-1).view(torch.uint8)
            t = pad
        _lib.check(self.lib.b200_hop_send_msg(self.hop.handle, header, len(header), _lib.ptr(t), nbytes, self.seq_out,
                                              _lib.stream_ptr(self.device)))
        self.seq_out += 1
        self._keep = t   # alive until the next send: the kernel reads it asynchronously

    def _flush(self) -> None:
        while self.connected and self._queue:
            header, t = self._queue.pop(0)
            self._send_now(header, t)

    def transfer_data(self, data, data_path: str = "", keep_data: bool = False):
        """Communicator.transfer_data (utils/node_worker.py:44-54) without the file and the socket."""
        header, t = self._encode(data)
        self._pump_control()
        if self.connected and not self._queue:
            self._send_now(header, t)
        else:
            self._queue.append((header, t))
            self._flush()
        return None

    def _decode(self, header: bytes, slot_ptr: int):
        magic, kind, dt, ndim, s0, s1, s2, s3, batch_size, seq_len, pos_offset, nbytes, n_slot = _HDR.unpack_from(header)
        if magic != _MAGIC:
            raise RuntimeError("[ERROR] malformed hop message (bad magic)")
        payload = tensor_view(slot_ptr + _lib.HOP_HEADER_BYTES, max(int(nbytes), 1), self.device)[:nbytes]
        if kind == KIND_JSON:
            return json.loads(bytes(payload.cpu().numpy()).decode())
        shape = [s0, s1, s2, s3][:ndim]
        t = payload.view(_DTYPES[dt]).view(shape).clone()   # the ring slot is reused: hand out an owned tensor
        if kind == KIND_TOKEN:
            return t
        slots = list(struct.unpack_from(f"<{n_slot}h", header, _HDR.size))
        if kind == KIND_INPUT:
            out = {"hidden_states": t, "batch_size": batch_size, "seq_len": seq_len}
        else:
            out = {"hidden_states": t, "cos": None, "sin": None}
            if pos_offset >= 0:
                out["pos_offset"] = pos_offset
        if slots:
            out["slot"] = slots
        return out

    def receive_data(self, no_block: bool = False, data_path: str = "", keep_data: bool = False):
        """Communicator.receive_data (utils/node_worker.py:56-67): ``zmq.Again`` when ``no_block`` and nothing is there."""
        deadline = time.time() + 3600.0
        while True:
            self._pump_control()
            self._flush()
            if self._inbox:
                return self._inbox.pop(0)
            got = _lib.check(self.lib.b200_hop_poll(self.hop.handle, self.seq_in, self._hdr_buf))
            if got == 1:
                slot_ptr = self.lib.b200_hop_local_slot(self.hop.handle, self.seq_in)
                self.seq_in += 1
                return self._decode(self._hdr_buf.raw, slot_ptr)
            if no_block:
                raise zmq.Again()
            if time.time() > deadline:
                raise RuntimeError("[ERROR] no message arrived within an hour")
            time.sleep(0.0002)

    def forward_request(self, data: dict, first_node_addr: str, request_socket=None) -> None:
        """NodeController._forward_request (utils/node_worker.py:476-485): hand the embedded prompt to the head of the
        chain.  The head being this very node (the usual placement) it goes straight to the inbox; otherwise it travels
        to the head's control socket as a JSON frame + the raw tensor bytes."""
        if _norm(first_node_addr) in self._my_names():
            self._inbox.append(data)
            return
        t = data["hidden_states"].detach().contiguous()
        blob = t.cpu().view(torch.uint8).numpy().tobytes()
        msg = {"p2p": "request", "want": _norm(first_node_addr), "dtype": _DTYPES.index(t.dtype), "shape": list(t.shape),
               "batch_size": int(data["batch_size"]), "seq_len": int(data["seq_len"])}
        self._ctl_send(request_socket if request_socket is not None else self.send_socket, msg, blob)

    def wait_connected(self) -> None:
        """Optional: block until the successor's mailbox is mapped (set-up time; the data path never needs it)."""
        t0 = time.time()
        while not self.connected:
            self._pump_control()
            if time.time() - t0 > self.connect_timeout_s:
                raise RuntimeError(f"[ERROR] no mailbox handle from {self.dst_addr} within {self.connect_timeout_s:.0f}s")
            time.sleep(0.001)

    def close(self) -> None:
        self.recv_socket.close(0)
        self.send_socket.close(0)
        self.hop.close()
