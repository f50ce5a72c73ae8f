"""Multi-GPU host logic of the registration path (BASELINE.json configs 4 and 5).

Two cases, one process per GPU (torchrun / torch.distributed):

* independent registrations (detections, templates, frames) -- `shard_round_robin`: static assignment, NO collective
  on the data path; the reference's own loop over templates is independent per item
  (/root/reference/src/pose_estimator/EstimHelpers/PoseEstimator.py:90-114).
* one very large problem (config 5, N = 20 000) -- `solve_row_sharded`: every rank computes full rows
  [r*chunk, (r+1)*chunk) of the consistency graph from the replicated 3xN operands (b2r_graph_rows), ONE all-gather of
  the bit rows (NCCL over NVLink on the GPU box; gloo in the CPU tests), then every rank continues the solve from the
  complete adjacency (b2r_solve_from_adjacency) -- replicated, so no broadcast of the result is needed.

The collective plumbing is `torch.distributed`; the row kernel and the solve are libb2r.so.  `compute_rows` is
injectable so that the partition / gather logic is testable on CPU with world_size 2 (tests/test_sharded_gloo.py).
"""
from __future__ import annotations

from typing import Callable, Tuple


def shard_round_robin(n_items: int, rank: int, world: int) -> range:
    """Static round-robin assignment of independent work items to ranks -- no collective needed."""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError(f"bad rank/world {rank}/{world}")
    return range(rank, n_items, world)


def row_chunk(n: int, world: int) -> int:
    """Rows per rank: equal chunks (all_gather_into_tensor needs equal sizes); the tail of the last chunk is padding."""
    return (n + world - 1) // world


def row_shard(n: int, rank: int, world: int) -> Tuple[int, int]:
    """[begin, end) rows of the consistency graph owned by `rank` (end clipped to n; may be empty for tiny n)."""
    c = row_chunk(n, world)
    b = min(rank * c, n)
    e = min(b + c, n)
    return b, e


def allgather_adjacency(n: int, pitch_words: int, compute_rows: Callable, device, group=None):
    """Row-sharded consistency graph -> complete bit matrix on every rank.

    compute_rows(row_begin, row_end, out) must fill `out` (an int32 tensor view [row_end-row_begin, pitch_words] on
    `device`) with the bit rows.  Returns an int32 tensor [world*chunk, pitch_words] whose first n rows are the graph.
    """
    import torch
    import torch.distributed as dist

    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    c = row_chunk(n, world)
    full = torch.zeros((world * c, pitch_words), dtype=torch.int32, device=device)
    b, e = row_shard(n, rank, world)
    if e > b:
        compute_rows(b, e, full[b:e])
    if world > 1:
        mine = full[rank * c:(rank + 1) * c]
        dist.all_gather_into_tensor(full, mine, group=group)  # in place: every rank's chunk lands at its row offset
    return full


EXCHANGE_MODES = ("nccl", "peer")


class ExchangeUnavailable(RuntimeError):
    """The requested exchange of adjacency rows cannot run on this box (e.g. no peer access between the GPUs)."""


def exchange_handles(mine: bytes, group=None):
    """Every rank's exported handle, in rank order (all_gather_object: plumbing, 152 bytes per rank, once)."""
    import torch.distributed as dist

    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return [mine]
    out = [None] * dist.get_world_size(group)
    dist.all_gather_object(out, mine, group=group)
    return out


def attach_peers(solver, n: int, group=None) -> None:
    """Export this rank's matrix, gather everybody's handles, map the peers (idempotent per (solver, n))."""
    import torch.distributed as dist

    if getattr(solver, "_shard_n", None) == n:
        return
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    err = None
    try:
        handles = exchange_handles(solver.shard_export(n), group)
        solver.shard_attach(rank, world, handles)
    except Exception as e:  # noqa: BLE001
        err = str(e)
    # the solve is collective: either every rank is attached or none uses the path
    errs = exchange_handles(err, group) if world > 1 else [err]
    bad = [f"rank {r}: {e}" for r, e in enumerate(errs) if e]
    if bad:
        raise ExchangeUnavailable("; ".join(bad))
    solver._shard_n = n


def solve_row_sharded(solver, src_dev, dst_dev, n: int, group=None, exchange: str = "nccl"):
    """Config 5 on the GPU: `solver` is a batch.BatchSolver, src_dev/dst_dev packed 3xN float64 CUDA tensors
    (replicated on every rank).  Returns (RegistrationSolution, adjacency tensor)."""
    import torch

    from . import batch

    if exchange not in EXCHANGE_MODES:
        raise ValueError(f"exchange must be one of {EXCHANGE_MODES}")
    if exchange == "peer":
        # tile-sharded tensor-core filter, every word stored into every rank's matrix over NVLink while it is computed
        # (b2r_solve_sharded, csrc/shard.cu); torch.distributed only carries the 152-byte IPC handles, once per (solver, n)
        attach_peers(solver, n, group)
        sol = solver.solve_sharded(src_dev.data_ptr(), dst_dev.data_ptr(), n)
        return sol, None
    pitch = batch.pitch_words(n)

    def rows(b, e, out):
        torch.cuda.current_stream().synchronize()
        solver.graph_rows(src_dev.data_ptr(), dst_dev.data_ptr(), n, b, e, out.data_ptr(), pitch)
        solver.sync()

    full = allgather_adjacency(n, pitch, rows, src_dev.device, group)
    torch.cuda.current_stream().synchronize()
    sol = solver.solve_from_adjacency(src_dev.data_ptr(), dst_dev.data_ptr(), n, full.data_ptr(), pitch)
    return sol, full
