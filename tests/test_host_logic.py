"""CPU-only: the host-side mirror of the reference interface (message kinds, config wire format,
Communicator, placement/validation rules) — no kernels involved."""
import json
import os
import time

import pytest
import torch
import zmq

from llm_sharding_b200.chain import LoopbackCommunicator
from llm_sharding_b200.config_sender import CONFIG_KEYS, ConfigSender
from llm_sharding_b200.engine import read_model_config
from llm_sharding_b200.forwarding_utils import build_position_ids
from llm_sharding_b200.node_worker import Communicator, NodeWorker
from oracle import synth


def test_message_kind_predicates_match_reference_keys():
    # utils/node_worker.py:311-317,374-376
    assert NodeWorker.is_input_token_info({"hidden_states": 0, "batch_size": 1, "seq_len": 3})
    assert not NodeWorker.is_input_token_info({"hidden_states": 0, "cos": 0, "sin": 0})
    assert NodeWorker.is_next_state_info({"hidden_states": 0, "cos": None, "sin": None, "pos_offset": 4})
    assert not NodeWorker.is_next_state_info(torch.zeros(1))
    assert NodeWorker.is_clear_KV_cache_command({"command": "clear_KV_cache", "origin_node": {}})
    assert not NodeWorker.is_clear_KV_cache_command({"command": "other"})


def test_config_sender_wire_format_and_validation():
    s = ConfigSender(node_port=45711)
    with pytest.raises(ValueError):
        s.build_config(0, 4, True, "tcp://*:45811", "tcp://127.0.0.1:45812")  # entry node needs first_node_addr
    s.build_config(shards_start=4, shards_end=16, can_receive_user_request=False,
                   src_addr="tcp://*:45811", dst_addr="tcp://127.0.0.1:45812")
    assert tuple(s.config) == CONFIG_KEYS
    ctx = zmq.Context()
    pull = ctx.socket(zmq.PULL)
    pull.bind("tcp://127.0.0.1:45711")
    s.send_config("127.0.0.1")
    assert pull.poll(5000)
    got = pull.recv_json()
    assert got == {"src_addr": "tcp://*:45811", "dst_addr": "tcp://127.0.0.1:45812", "can_receive_user_request": False,
                   "first_node_addr": "", "shards_start": 4, "shards_end": 16}
    pull.close(0)
    s.close()


def test_communicator_roundtrip_dict_and_tensor():
    a = Communicator("tcp://127.0.0.1:45821", "tcp://127.0.0.1:45822")
    b = Communicator("tcp://127.0.0.1:45822", "tcp://127.0.0.1:45821")
    try:
        with pytest.raises(zmq.Again):
            b.receive_data(no_block=True)
        msg = {"hidden_states": torch.randn(1, 3, 8).bfloat16(), "cos": None, "sin": None, "pos_offset": 5}
        a.transfer_data(msg)
        got = b.receive_data()
        assert torch.equal(got["hidden_states"], msg["hidden_states"]) and got["pos_offset"] == 5
        b.transfer_data(torch.tensor([7], dtype=torch.int64))
        t = a.receive_data()
        assert isinstance(t, torch.Tensor) and t.tolist() == [7]
        assert a.change_dst_addr("tcp://127.0.0.1:45823") == "tcp://127.0.0.1:45823"
    finally:
        a.close()
        b.close()


def test_loopback_communicator_is_communicator_shaped():
    a = LoopbackCommunicator("tcp://*:45831", "tcp://127.0.0.1:45832")
    b = LoopbackCommunicator("tcp://*:45832", "tcp://127.0.0.1:45831")
    with pytest.raises(zmq.Again):
        b.receive_data(no_block=True)
    a.transfer_data({"x": 1})
    assert b.receive_data() == {"x": 1}
    a.close(); b.close()


def test_build_position_ids_semantics():
    class Cache:
        def get_seq_length(self):
            return 5
    p = build_position_ids(Cache(), 3, "cpu", batch_size=2)
    assert p.dtype == torch.long and p.tolist() == [[5, 6, 7], [5, 6, 7]]
    assert build_position_ids(None, 2, "cpu").tolist() == [[0, 1]]
    k = torch.zeros(1, 2, 9, 4)
    assert build_position_ids((k, k), 1, "cpu").tolist() == [[9]]
    with pytest.raises(ValueError):
        build_position_ids(object(), 1, "cpu")


def test_read_model_config_and_cpu_device_is_refused(tmp_path):
    cfg = synth.PRESETS["tiny"]
    d = str(tmp_path / "tiny_bfloat16")
    synth.write_shards(d, cfg, 1, torch.bfloat16, layers=range(0))
    mc = read_model_config(d)
    assert mc["head_dim"] == 128 and mc["num_key_value_heads"] == 2 and mc["rope_theta"] == 10000.0
    assert set(os.listdir(d)) >= {"config.json", "embedding.pth", "final_norm.pth", "lm_head.pth", "tokenizer.json"}
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        NodeWorker("tcp://*:45841", "tcp://127.0.0.1:45842", False, d, device="cpu",
                   communicator=LoopbackCommunicator("tcp://*:45841", "tcp://127.0.0.1:45842"))


def test_shard_layout_matches_model_sharder_keys(tmp_path):
    cfg = synth.PRESETS["tiny"]
    d = str(tmp_path / "s")
    synth.write_shards(d, cfg, 2, torch.float32, layers=range(1))
    st = torch.load(os.path.join(d, "block_0.pth"))
    from llm_sharding_b200.engine import LAYER_KEYS
    assert tuple(st) == LAYER_KEYS == synth.LAYER_KEYS
    assert st["mlp.down_proj.weight"].shape == (cfg.hidden_size, cfg.intermediate_size)


def test_profiler_latency_fit_recovers_models():
    from llm_sharding_b200.node_profiler import NodeProfiler
    s = [8, 16, 32, 64, 128, 256, 512]
    fit = NodeProfiler._fit_latency_models(s, [2e-6 * x * x + 3e-4 * x + 0.01 for x in s])
    a, b, c = fit["quadratic"]["coefficients"]
    assert abs(a - 2e-6) < 1e-10 and abs(b - 3e-4) < 1e-8 and abs(c - 0.01) < 1e-6
    assert fit["quadratic"]["rmse"] < 1e-9 and fit["quadratic"]["r_squared"] > 0.999999
    assert fit["linear"]["r_squared"] < fit["quadratic"]["r_squared"]
    with pytest.raises(ValueError):
        NodeProfiler._fit_latency_models([1, 2], [1.0, 2.0])
    with pytest.raises(ValueError):
        NodeProfiler._fit_latency_models([1, 2, 3], [1.0, 2.0])


def test_communicator_does_not_unpickle_code():
    """The ZeroMQ PULL socket is bound to a network address: a pickle that would execute code must be refused, while the
    three message kinds of the reference still pass."""
    import io
    import pickle
    import zmq
    from llm_sharding_b200.node_worker import Communicator
    a = Communicator("tcp://127.0.0.1:45990", "tcp://127.0.0.1:45990")
    try:
        msgs = [{"hidden_states": torch.ones(1, 2, 4, dtype=torch.bfloat16), "batch_size": 1, "seq_len": 2},
                {"hidden_states": torch.ones(1, 1, 4), "cos": None, "sin": None, "pos_offset": 3, "slot": [0]},
                torch.tensor([5], dtype=torch.int64),
                {"command": "clear_KV_cache", "origin_node": {"src_addr": "x", "dst_addr": "y", "shards_start": 0, "shards_end": 2}}]
        for m in msgs:
            a.transfer_data(m)
            got = a.receive_data()
            if isinstance(m, torch.Tensor):
                assert torch.equal(got, m)
            else:
                assert set(got) == set(m)

        class Evil:
            def __reduce__(self):
                return (eval, ("1+1",))
        buf = io.BytesIO()
        torch.save({"x": Evil()}, buf)
        a.send_socket.send(buf.getvalue())
        with pytest.raises(pickle.UnpicklingError):
            a.receive_data()
    finally:
        a.close()
