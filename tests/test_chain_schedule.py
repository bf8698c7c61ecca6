"""CPU-only: the host-side protocol of the multi-GPU ring (layer split, sequence numbers, mailbox slots), single
process and across two real processes on the gloo backend (world_size 2)."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from llm_sharding_b200.chain import RingSchedule, layer_range


def test_layer_range_covers_all_layers_contiguously():
    for n_layers, world in [(32, 1), (32, 2), (32, 4), (32, 8), (32, 3), (28, 5), (4, 4)]:
        spans = [layer_range(r, world, n_layers) for r in range(world)]
        assert spans[0][0] == 0 and spans[-1][1] == n_layers
        for (a0, a1), (b0, b1) in zip(spans, spans[1:]):
            assert a1 == b0 and a1 > a0
        sizes = [e - s for s, e in spans]
        assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        layer_range(0, 5, 4)


def test_ring_schedule_slots_are_stable_per_stream():
    world, n_streams = 4, 4
    scheds = [RingSchedule(r, world, n_streams) for r in range(world)]
    sent = {r: [] for r in range(world)}   # messages produced by stage r, in order
    for i in range(n_streams):
        for r in range(world):
            it = scheds[r].prefill(i)
            assert (it["wait"] is None) == (r == 0)
            sent[r].append((it["send"], i))
    for _ in range(3):
        for i in range(n_streams):
            for r in range(world):
                it = scheds[r].decode(i)
                assert it["wait"][1] == i and it["send"][1] == i  # stream i always uses mailbox slot i
                sent[r].append((it["send"], i))
    # what stage r consumes as its k-th message is what stage r-1 produced as its k-th message
    for r in range(world):
        seqs = [s for (s, _), _ in sent[r]]
        assert seqs == list(range(len(seqs)))
    with pytest.raises(RuntimeError):
        scheds[0].prefill(0)
    with pytest.raises(RuntimeError):
        RingSchedule(0, 2, 2).decode(1)


def test_clear_keeps_both_ends_of_the_token_hop_counting_alike():
    """bench_chain.py prefills, clears (no token was consumed) and prefills again: afterwards the head must wait for the
    very message the tail sends next, not for one whose flag a dropped request already published."""
    world, n_streams = 3, 2
    scheds = [RingSchedule(r, world, n_streams) for r in range(world)]
    for _ in range(2):
        for i in range(n_streams):
            for s in scheds:
                s.prefill(i)
        for s in scheds:
            s.clear()
    tail_sends = []
    for i in range(n_streams):
        for s in scheds:
            it = s.prefill(i)
        tail_sends.append(it["send"])
    for i in range(n_streams):
        assert scheds[0].decode(i)["wait"] == tail_sends[i]


def test_finish_keeps_the_ring_counts_aligned_under_a_mixed_schedule():
    """Streams are retired and given new prompts while the others keep decoding (BASELINE configs[4]): every message a
    stage produces is consumed by its successor, in order, for the stream it was meant for, and the messages in flight
    on one hop never share a mailbox slot."""
    world, n_streams = 3, 3
    scheds = [RingSchedule(r, world, n_streams) for r in range(world)]
    queues = [[] for _ in range(world)]   # queues[r]: messages produced by stage r, not yet consumed by stage r+1

    def run(kind, stream):
        for r in range(world):
            it = getattr(scheds[r], kind)(stream)
            if it["wait"] is not None:
                seq_slot, st = queues[(r - 1) % world].pop(0)
                assert seq_slot == it["wait"] and st == stream, (kind, stream, r)
            if it["send"] is not None:
                q = queues[r]
                q.append((it["send"], stream))
                assert len({slot for (_, slot), _ in q}) == len(q) <= n_streams
    for i in range(n_streams):
        run("prefill", i)
    # streams are served round-robin; at its turn a stream decodes, or finishes (and may take a new prompt)
    plan = [("decode", 0), ("decode", 1), ("decode", 2),
            ("decode", 0), ("finish", 1), ("prefill", 1), ("decode", 2),
            ("finish", 0), ("decode", 1), ("finish", 2), ("prefill", 2),
            ("decode", 1), ("decode", 2),
            ("prefill", 0), ("decode", 1), ("decode", 2), ("decode", 0)]
    for kind, stream in plan:
        run(kind, stream)
    # what is left in flight: exactly one token per live stream, waiting in the head's mailbox
    assert [len(q) for q in queues] == [0, 0, 3]
    with pytest.raises(RuntimeError):   # the oldest outstanding token belongs to stream 1, not 2
        scheds[0].decode(2)
    with pytest.raises(RuntimeError):
        scheds[0].finish(0)


@pytest.mark.parametrize("seed", range(8))
def test_random_round_robin_schedules_keep_the_ring_consistent(seed):
    """Property test: any schedule in which the live streams take turns in a fixed order — each turn being a decode, a
    finish, or a finish followed at once by a new prefill, with idle streams re-entering at any time — keeps every
    hop's FIFO, stream and mailbox-slot bookkeeping consistent on every stage, for several ring sizes."""
    import random
    rng = random.Random(seed)
    world, n_streams = rng.choice([1, 2, 3, 5, 8]), rng.choice([1, 2, 3, 4, 7])
    scheds = [RingSchedule(r, world, n_streams) for r in range(world)]
    queues = [[] for _ in range(world)]

    def run(kind, stream):
        for r in range(world):
            it = getattr(scheds[r], kind)(stream)
            if it["wait"] is not None:
                assert queues[(r - 1) % world], (kind, stream, r, "nothing to consume")
                seq_slot, st = queues[(r - 1) % world].pop(0)
                assert seq_slot == it["wait"] and st == stream, (kind, stream, r)
            if it["send"] is not None:
                q = queues[r]
                q.append((it["send"], stream))
                assert len({slot for (_, slot), _ in q}) == len(q) <= n_streams
    order = []                      # live streams in the order their tokens come back
    idle = list(range(n_streams))
    for _ in range(200):
        if idle and (not order or rng.random() < 0.3):
            s_new = idle.pop(rng.randrange(len(idle)))
            run("prefill", s_new)
            order.append(s_new)
            continue
        s_cur = order.pop(0)        # the stream whose token is the oldest one outstanding
        action = rng.random()
        if action < 0.7:
            run("decode", s_cur)
            order.append(s_cur)
        elif action < 0.85:
            run("finish", s_cur)
            idle.append(s_cur)
        else:
            run("finish", s_cur)
            run("prefill", s_cur)
            order.append(s_cur)
        if len(order) > 1:          # any other live stream is refused: its token is not at the head of the FIFO
            with pytest.raises(RuntimeError):
                scheds[0].decode(order[-1]) if order[-1] != order[0] else scheds[0].finish(idle[0] if idle else n_streams + 5)
    # every token still in flight sits in the head's mailbox (the tail -> head hop), one per live stream
    assert [len(q) for q in queues[:-1]] == [0] * (world - 1)
    assert len(queues[-1]) == len(order)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _ring_worker(rank, world, port, n_streams, rounds, out_q):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        sched = RingSchedule(rank, world, n_streams)
        nxt, prv = (rank + 1) % world, (rank - 1) % world
        log, pending = [], []
        # every "message" carries (seq, slot, stream, hops so far); the consumer checks it is the one it expects.
        # Sends are non-blocking, like a mailbox store: the head does not consume a prefill's token until it decodes.
        def handle(item):
            if item["wait"] is not None:
                m = torch.zeros(4, dtype=torch.int64)
                dist.recv(m, src=prv)
                assert (int(m[0]), int(m[1])) == item["wait"], (rank, m.tolist(), item)
                assert int(m[2]) == item["stream"]
                hops = int(m[3]) + 1
            else:
                hops = 0
            msg = torch.tensor([item["send"][0], item["send"][1], item["stream"], hops])
            pending.append((dist.isend(msg, dst=nxt), msg))
            log.append(hops)
        for i in range(n_streams):
            handle(sched.prefill(i))
        for _ in range(rounds):
            for i in range(n_streams):
                handle(sched.decode(i))
        # drain: the head still has the last round's tokens in its mailbox
        if rank == 0:
            for _ in range(n_streams):
                m = torch.zeros(4, dtype=torch.int64)
                dist.recv(m, src=prv)
        for req, _ in pending:
            req.wait()
        gathered = [None] * world
        dist.all_gather_object(gathered, (rank, layer_range(rank, world, 32), log[-1]))
        if rank == 0:
            out_q.put(gathered)
    finally:
        dist.destroy_process_group()


def test_ring_protocol_two_processes_gloo():
    world, n_streams, rounds = 2, 2, 5
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_ring_worker, args=(r, world, port, n_streams, rounds, q)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(90)
    hung = [p for p in procs if p.is_alive()]
    for p in hung:
        p.terminate()
    assert not hung, "ring protocol deadlocked"
    assert all(p.exitcode == 0 for p in procs)
    gathered = q.get(timeout=10)
    assert [g[1] for g in gathered] == [(0, 16), (16, 32)]
    # the last message a stage handled has travelled (rounds * world + rank) hops around the ring
    assert [g[2] for g in gathered] == [rounds * world + r for r in range(world)]


# ---------------------------------------------------------------------------------------------- stop rules in the ring
class _FakeStage:
    """Records the work items a RingServer enqueues; tokens are made up by `token_fn(stream, k)` on the head."""

    def __init__(self, rank, world, n_streams, rows=1, token_fn=None):
        from llm_sharding_b200.chain import RingSchedule
        self.sched = RingSchedule(rank, world, n_streams)
        self.rank, self.world, self.n_streams, self.rows = rank, world, n_streams, rows
        self.device = torch.device("cpu")
        self.dist = dist
        self.tokens, self.log = [], []
        self.token_fn = token_fn or (lambda stream, nth, k: 100 * stream + k)
        self._count = [0] * n_streams
        self._nth = [-1] * n_streams     # how many requests this stream has started, minus one

    def _tok(self, stream):
        k = self._count[stream]
        self._count[stream] += 1
        self.tokens.append(torch.tensor([self.token_fn(stream, self._nth[stream], k)] * self.rows))

    def prefill(self, stream, prompt_len, ids=None):
        self.sched.prefill(stream)
        self._count[stream] = 0
        self._nth[stream] += 1
        self.log.append(("prefill", stream, prompt_len))

    def decode(self, stream, record=False, via_host=False):
        self.sched.decode(stream)
        if record:
            self._tok(stream)
        self.log.append(("decode", stream))

    def finish(self, stream, record=False):
        self.sched.finish(stream)
        if record:
            self._tok(stream)
        self.log.append(("finish", stream))


class _NoEvent:
    def record(self, *a): pass
    def query(self): return True
    def synchronize(self): pass


def test_ring_server_applies_the_reference_stop_rules_per_stream(monkeypatch):
    """max_new_tokens gives exactly that many tokens per request (utils/node_worker.py:291: the last token is not fed
    back); an EOS ends the request early and cuts its output there; a freed stream takes the next queued prompt at once."""
    from llm_sharding_b200 import chain
    monkeypatch.setattr(torch.cuda, "Event", lambda *a, **k: _NoEvent())
    monkeypatch.setattr(torch.cuda, "current_stream", lambda *a, **k: None)
    EOS = 2
    # the second request served by stream 1 (r2) emits EOS as its third token although it may produce six

    def token_fn(stream, nth, k):
        return EOS if (stream == 1 and nth == 1 and k == 2) else 10 * (stream + 1) + k + 1

    stage = _FakeStage(0, 1, 2, token_fn=token_fn)
    srv = chain.RingServer(stage, None, eos_token_id=EOS)
    r0 = srv.submit(torch.zeros(1, 5, dtype=torch.long), 4)
    r1 = srv.submit(torch.zeros(1, 9, dtype=torch.long), 2)
    r2 = srv.submit(torch.zeros(1, 3, dtype=torch.long), 6)   # lands on whichever stream frees first (stream 1)
    out = srv.serve()
    assert [t for t in out[r0][:, 0].tolist()] == [11, 12, 13, 14]
    assert [t for t in out[r1][:, 0].tolist()] == [21, 22]
    assert out[r2][:, 0].tolist() == [21, 22, EOS]           # cut at the EOS although max_new_tokens was 6
    kinds = [e[0] for e in stage.log]
    assert kinds.count("prefill") == 3 and kinds.count("finish") == 3
    # request r0: prefill + 3 decodes + finish = 4 tokens; nothing is decoded for a finished stream
    assert [e for e in stage.log if e[1] == 0][:5] == [("prefill", 0, 5), ("decode", 0), ("decode", 0), ("decode", 0), ("finish", 0)]


def test_ring_planner_eos_cuts_the_output_and_retires_the_stream():
    from llm_sharding_b200.chain import ACT_DECODE, ACT_IDLE, RingPlanner
    p = RingPlanner(1, rows=2, eos_token_id=2)
    rid = p.submit(4, 10)
    assert p.plan() == [0, 4]
    p.observe(0, rid, [7, 2])            # one row hit EOS: the micro-batch goes on until every row has
    assert p.plan() == [0, ACT_DECODE]
    p.observe(0, rid, [2, 9])
    assert p.plan() == [1, ACT_IDLE]     # retired at the first round that knows
    p.observe(0, rid, [5, 5])            # a token that was already on its way: dropped
    assert p.done[rid] == [[7, 2], [2, 9]] and p.idle()
    with pytest.raises(ValueError):
        p.submit(0, 3)


def _server_proc(rank, world, port, q):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from llm_sharding_b200 import chain
        torch.cuda.Event = lambda *a, **k: _NoEvent()
        torch.cuda.current_stream = lambda *a, **k: None
        stage = _FakeStage(rank, world, 2)
        srv = chain.RingServer(stage, dist.group.WORLD)
        if rank == 0:
            srv.submit(torch.zeros(1, 6, dtype=torch.long), 3)
            srv.submit(torch.zeros(1, 2, dtype=torch.long), 5)
            srv.submit(torch.zeros(1, 8, dtype=torch.long), 1)
        out = srv.serve()
        q.put((rank, stage.log, {k: v.tolist() for k, v in out.items()}))
    finally:
        dist.destroy_process_group()


def test_ring_server_plan_reaches_every_stage_gloo():
    """world_size 2 on gloo: the head's per-round plan is broadcast and both stages enqueue the same work items."""
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    procs = [ctx.Process(target=_server_proc, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = dict()
    for _ in range(2):
        rank, log, out = q.get(timeout=120)
        got[rank] = (log, out)
    for p in procs:
        p.join(60)
    assert all(p.exitcode == 0 for p in procs)
    assert got[0][0] == got[1][0] and len(got[0][0]) > 0
    assert {k: len(v) for k, v in got[0][1].items()} == {0: 3, 1: 5, 2: 1} and got[1][1] == {}
