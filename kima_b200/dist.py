"""One process per GPU: lane sharding and DNest4's level-creation barrier.

Lanes (DNest4 "threads", each with its own RNG and particles) are independent
between barriers (DNest4/code/SamplerImpl.h:134-165), so the likelihood path
shards with no data-path collective.  The only exchange is the barrier merge
(SamplerImpl.h:306-329): gather every rank's `above` list of
(log-likelihood, tiebreaker) pairs and sum the per-level counter deltas, after
which every rank can run do_bookkeeping (SamplerImpl.h:377-440) identically.

Two transports with the same semantics:
  * torch.distributed (gloo on CPU for tests, nccl on GPUs), used by bench.py;
  * libkima_cuda's own kc_allgather_above / kc_allreduce_levels (NCCL bound
    directly), which is what a C++ host shim would call.
"""
from __future__ import annotations

import ctypes
import os
from typing import List, Tuple

import numpy as np


def shard_lanes(total: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous block [start, stop) of lanes owned by `rank`."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError("bad rank/world")
    base, rem = divmod(total, world)
    start = rank * base + min(rank, rem)
    return start, start + base + (1 if rank < rem else 0)


def env_rank_world() -> Tuple[int, int, int]:
    return (int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")),
            int(os.environ.get("LOCAL_RANK", "0")))


class TorchBarrierExchange:
    """Barrier merge over an initialised torch.distributed process group."""

    def __init__(self, device=None):
        import torch
        import torch.distributed as dist
        if not dist.is_initialized():
            raise RuntimeError("torch.distributed is not initialised")
        self.torch, self.dist = torch, dist
        self.world = dist.get_world_size()
        self.rank = dist.get_rank()
        self.device = device if device is not None else torch.device("cpu")

    def allgather_above(self, pairs: np.ndarray) -> Tuple[np.ndarray, List[int]]:
        """pairs[n,2] of this rank -> (all pairs, rank-major; counts per rank)."""
        torch, dist = self.torch, self.dist
        pairs = np.ascontiguousarray(pairs, np.float64).reshape(-1, 2)
        n = torch.tensor([pairs.shape[0]], dtype=torch.int64, device=self.device)
        counts = [torch.zeros_like(n) for _ in range(self.world)]
        dist.all_gather(counts, n)
        counts = [int(c.item()) for c in counts]
        mx = max(counts)
        if mx == 0:
            return np.zeros((0, 2)), counts
        send = torch.zeros((mx, 2), dtype=torch.float64, device=self.device)
        if pairs.shape[0]:
            send[:pairs.shape[0]] = torch.from_numpy(pairs).to(self.device)
        recv = [torch.zeros_like(send) for _ in range(self.world)]
        dist.all_gather(recv, send)
        out = [recv[r][:counts[r]].cpu().numpy() for r in range(self.world)]
        return np.concatenate(out, axis=0), counts

    def allreduce_levels(self, counters: np.ndarray) -> np.ndarray:
        """Sum of per-level (accepts, tries, visits, exceeds) deltas over ranks."""
        torch, dist = self.torch, self.dist
        c = np.ascontiguousarray(counters, np.uint64)
        t = torch.from_numpy(c.astype(np.int64)).to(self.device)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return t.cpu().numpy().astype(np.uint64).reshape(c.shape)

    def max_float(self, v: float) -> float:
        torch, dist = self.torch, self.dist
        t = torch.tensor([v], dtype=torch.float64, device=self.device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sum_float(self, v: float) -> float:
        torch, dist = self.torch, self.dist
        t = torch.tensor([v], dtype=torch.float64, device=self.device)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    def barrier(self) -> None:
        self.dist.barrier()


class NcclBarrierExchange:
    """The same merge through libkima_cuda's own NCCL binding (kc_comm_*).

    The 128-byte NCCL unique id is created on rank 0 and handed to the other
    ranks by `broadcast_bytes`, any callable that broadcasts a bytes object
    from rank 0 (e.g. over the torch.distributed store)."""

    def __init__(self, engine, rank: int, world: int, broadcast_bytes):
        from . import _lib
        self._lib = _lib
        self.engine, self.rank, self.world = engine, rank, world
        lib = engine.lib
        buf = ctypes.create_string_buffer(128)
        if rank == 0:
            _lib.check(lib.kc_comm_unique_id(buf))
        raw = broadcast_bytes(bytes(buf.raw) if rank == 0 else None)
        buf = ctypes.create_string_buffer(raw, 128)
        _lib.check(lib.kc_comm_init(engine._ctx, rank, world, buf))

    def allgather_above(self, pairs: np.ndarray, max_total: int = 1 << 20):
        pairs = np.ascontiguousarray(pairs, np.float64).reshape(-1, 2)
        out = np.zeros((max_total, 2))
        counts = (ctypes.c_int * self.world)()
        dp = ctypes.POINTER(ctypes.c_double)
        self._lib.check(self.engine.lib.kc_allgather_above(
            self.engine._ctx, pairs.ctypes.data_as(dp), pairs.shape[0], out.ctypes.data_as(dp),
            max_total, counts))
        counts = list(counts)
        return out[:sum(counts)].copy(), counts

    def allreduce_levels(self, counters: np.ndarray) -> np.ndarray:
        c = np.ascontiguousarray(counters, np.uint64).copy()
        self._lib.check(self.engine.lib.kc_allreduce_levels(
            self.engine._ctx, c.ctypes.data_as(ctypes.POINTER(ctypes.c_uint64)), c.size))
        return c


def merge_above_and_create_level(all_above: np.ndarray, new_level_interval: int,
                                 compression: float = float(np.e)):
    """The level-creation rule of do_bookkeeping (SamplerImpl.h:382-399): once
    `all_above` holds at least new_level_interval entries, sort them by
    (value, tiebreaker), cut the new level at index (1 - 1/compression) * size
    (`compression` is the run's -c option, e by default, SamplerImpl.h:389) and keep
    the entries above it.  Returns (new_level or None, remaining above)."""
    a = np.asarray(all_above, np.float64).reshape(-1, 2)
    if a.shape[0] < new_level_interval:
        return None, a
    order = np.lexsort((a[:, 1], a[:, 0]))
    a = a[order]
    idx = int((1.0 - 1.0 / compression) * a.shape[0])
    return a[idx].copy(), a[idx + 1:].copy()
