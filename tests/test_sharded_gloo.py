"""Multi-rank host logic on CPU: world_size-2 gloo.  The row kernel is replaced by the oracle's edge
test so that only the partition + all-gather plumbing of poseestimator_b200.sharded is exercised here; the GPU test
`test_row_sharded_graph_equals_full` covers the kernel side."""
import os
import socket

import numpy as np
import pytest

from poseestimator_b200 import sharded, synth


def test_partitions_cover_everything_once():
    for n, world in [(10, 1), (10, 3), (64, 8), (20000, 8), (7, 8), (1, 2)]:
        seen = np.zeros(n, int)
        for r in range(world):
            for i in sharded.shard_round_robin(n, r, world):
                seen[i] += 1
        assert (seen == 1).all()
        rows = np.zeros(n, int)
        c = sharded.row_chunk(n, world)
        for r in range(world):
            b, e = sharded.row_shard(n, r, world)
            assert 0 <= b <= e <= n and e - b <= c
            rows[b:e] += 1
        assert (rows == 1).all()
    with pytest.raises(ValueError):
        sharded.shard_round_robin(4, 2, 2)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n, seed, q):
    import torch
    import torch.distributed as dist

    import oracle
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        pb = synth.make_problem(n, 0.8, seed)
        W = (n + 31) // 32
        pitch = 4 * ((n + 127) // 128)
        full_ref = oracle.tim_graph(pb.src, pb.dst, pb.noise_bound)

        def rows(b, e, out):  # CPU stand-in for b2r_graph_rows
            blk = np.zeros((e - b, pitch), np.uint32)
            blk[:, :W] = full_ref[b:e]
            out.copy_(torch.from_numpy(blk.view(np.int32)))

        full = sharded.allgather_adjacency(n, pitch, rows, torch.device("cpu"))
        got = full.numpy().view(np.uint32)[:n, :W]
        ok = bool(np.array_equal(got, full_ref)) and full.shape[0] == world * sharded.row_chunk(n, world)
        # independent detections: each rank takes its round-robin share, results gathered on the host side only
        mine = list(sharded.shard_round_robin(13, rank, world))
        gathered = [None] * world
        dist.all_gather_object(gathered, mine)
        ok = ok and sorted(sum(gathered, [])) == list(range(13))

        # peer exchange (b2r_solve_sharded): handles gathered in rank order, attach is all-or-nothing over the ranks
        class FakeSolver:
            def __init__(self, fail):
                self.fail, self.attached = fail, None
            def shard_export(self, n):
                return bytes([rank]) * 152
            def shard_attach(self, r, w, handles):
                if self.fail:
                    raise RuntimeError("no peer access")
                self.attached = (r, w, [h[0] for h in handles])
        good = FakeSolver(False)
        sharded.attach_peers(good, n)
        ok = ok and good.attached == (rank, world, list(range(world))) and good._shard_n == n
        sharded.attach_peers(good, n)                      # idempotent per (solver, n)
        half = FakeSolver(rank == 1)                       # one rank cannot attach: EVERY rank must see ExchangeUnavailable
        try:
            sharded.attach_peers(half, n)
            ok = False
        except sharded.ExchangeUnavailable as e:
            ok = ok and "rank 1" in str(e) and getattr(half, "_shard_n", None) is None
        q.put((rank, ok))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n", [257, 1000])
def test_row_sharded_allgather_world2(oracle_mod, n):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, n, 5, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert res == [(0, True), (1, True)]
