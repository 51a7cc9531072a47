"""Scheduler / selector / scaler tables and model-store semantics (reference
``synchronous_scheduler_test.cc:28-83``, ``asynchronous_scheduler_test.cc:28-72``,
``scheduled_cardinality_test.cc:28-72``, ``model_store_test.cc:390-545``)."""
import socket
import threading

import numpy as np
import pytest

from metisfl_b200 import _core
from helpers import model_bytes

LEARNERS = [f"learner{i}" for i in range(1, 6)]


def test_synchronous_scheduler_barrier_and_no_double_schedule():
    s = _core.SynchronousScheduler()
    for l in LEARNERS[:-1]:
        assert s.schedule_next(l, LEARNERS) == []
    assert s.schedule_next(LEARNERS[0], LEARNERS) == []  # the same learner reporting twice does not count twice
    assert sorted(s.schedule_next(LEARNERS[-1], LEARNERS)) == LEARNERS
    assert s.schedule_next(LEARNERS[0], LEARNERS) == []  # barrier state was reset


def test_synchronous_scheduler_barrier_follows_active_set():
    s = _core.SynchronousScheduler()
    assert s.schedule_next("learner1", LEARNERS[:3]) == []
    assert s.schedule_next("learner2", LEARNERS[:3]) == []
    # learner3 left: the barrier compares against the current active set
    assert sorted(s.schedule_next("learner2", LEARNERS[:2])) == LEARNERS[:2]


def test_asynchronous_scheduler_returns_caller():
    s = _core.AsynchronousScheduler()
    for l in LEARNERS:
        assert s.schedule_next(l, LEARNERS) == [l]


@pytest.mark.parametrize("n_scheduled,expected", [(0, 5), (1, 5), (2, 2), (5, 5)])
def test_scheduled_cardinality(n_scheduled, expected):
    assert len(_core.scheduled_cardinality_select(LEARNERS[:n_scheduled], LEARNERS)) == expected


def test_scalers():
    parts = [("a", 100.0, 10.0), ("b", 300.0, 30.0)]
    assert _core.compute_scaling_factors(3, 2, parts) == {"a": 0.25, "b": 0.75}     # NUM_TRAINING_EXAMPLES
    assert _core.compute_scaling_factors(1, 2, parts) == {"a": 0.25, "b": 0.75}     # NUM_COMPLETED_BATCHES
    assert _core.compute_scaling_factors(2, 2, parts) == {"a": 0.5, "b": 0.5}       # NUM_PARTICIPANTS
    assert _core.compute_scaling_factors(3, 1, parts[:1]) == {"a": 1.0}             # federation of one
    assert _core.compute_scaling_factors(3, 5, parts[:1]) == {"a": 100.0}           # single participant: raw
    assert _core.compute_scaling_factors(1, 5, parts[1:]) == {"b": 30.0}


def _models(n):
    return [model_bytes([np.full(16, i, dtype=np.float64)]) for i in range(n)]


def _check_store(store_factory):
    ms = _models(5)
    # no eviction: everything is kept, oldest -> newest
    s = store_factory(False, 0)
    for m in ms:
        s.insert_model("l1", m)
    assert s.get_learner_lineage_length("l1") == 5
    assert s.select_models([("l1", 0)])["l1"] == ms
    assert s.select_models([("l1", -1)])["l1"] == ms
    assert s.select_models([("l1", 2)])["l1"] == ms[-2:]
    assert s.select_models([("l1", 6)])["l1"] == []          # over-ask -> empty
    assert s.select_models([("nobody", 1)]) == {}
    s.reset_state()
    s.erase_models(["l1"])
    assert s.get_learner_lineage_length("l1") == 0
    # lineage-length eviction k = 2 (k = 0 behaves as 1)
    s = store_factory(True, 2)
    for m in ms:
        s.insert_model("l1", m)
        s.insert_model("l2", m)
    assert s.get_learner_lineage_length("l1") == 2
    assert s.select_models([("l1", 0), ("l2", 1)]) == {"l1": ms[-2:], "l2": ms[-1:]}
    s0 = store_factory(True, 0)
    for m in ms:
        s0.insert_model("l1", m)
    assert s0.select_models([("l1", 0)])["l1"] == ms[-1:]
    s.expunge()
    assert s.get_learner_lineage_length("l1") == 0 and s.get_learner_lineage_length("l2") == 0


def test_hash_map_model_store():
    _check_store(lambda evict, k: _core.HashMapModelStore(evict, k))


class _MiniRedis(threading.Thread):
    """Just enough RESP2 for the store: RPUSH, LRANGE, DEL, FLUSHDB, MULTI/EXEC."""

    def __init__(self):
        super().__init__(daemon=True)
        self.sock = socket.socket()
        self.sock.bind(("127.0.0.1", 0))
        self.sock.listen(4)
        self.port = self.sock.getsockname()[1]
        self.data = {}
        self.commands = []

    def run(self):
        while True:
            try:
                conn, _ = self.sock.accept()
            except OSError:
                return
            threading.Thread(target=self._serve, args=(conn,), daemon=True).start()

    def _serve(self, conn):
        f = conn.makefile("rb")
        queued = None
        while True:
            line = f.readline()
            if not line:
                return
            n = int(line[1:])
            args = []
            for _ in range(n):
                ln = int(f.readline()[1:])
                args.append(f.read(ln))
                f.read(2)
            cmd = args[0].upper()
            self.commands.append(cmd)
            if cmd == b"MULTI":
                queued = []
                conn.sendall(b"+OK\r\n")
            elif cmd == b"EXEC":
                out = b"*%d\r\n" % len(queued) + b"".join(self._exec(a) for a in queued)
                queued = None
                conn.sendall(out)
            elif queued is not None:
                queued.append(args)
                conn.sendall(b"+QUEUED\r\n")
            else:
                conn.sendall(self._exec(args))

    def _exec(self, args):
        cmd = args[0].upper()
        if cmd == b"RPUSH":
            self.data.setdefault(args[1], []).extend(args[2:])
            return b":%d\r\n" % len(self.data[args[1]])
        if cmd == b"LRANGE":
            items = self.data.get(args[1], [])
            return b"*%d\r\n" % len(items) + b"".join(b"$%d\r\n%s\r\n" % (len(i), i) for i in items)
        if cmd == b"DEL":
            return b":%d\r\n" % (1 if self.data.pop(args[1], None) is not None else 0)
        if cmd == b"FLUSHDB":
            self.data.clear()
            return b"+OK\r\n"
        return b"-ERR unknown command\r\n"


def test_redis_model_store_against_resp_server():
    srv = _MiniRedis()
    srv.start()
    _check_store(lambda evict, k: _core.RedisModelStore("127.0.0.1", srv.port, evict, k))
    assert {b"RPUSH", b"MULTI", b"LRANGE", b"EXEC", b"DEL", b"FLUSHDB"} <= set(srv.commands)
    srv.sock.close()


def test_redis_model_store_connection_failure_is_an_error():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    with pytest.raises(RuntimeError):
        _core.RedisModelStore("127.0.0.1", port, False, 0)
