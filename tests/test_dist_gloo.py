"""CPU, world size 2 over gloo: the N>1 host path (lane sharding + DNest4's level-creation
barrier merge, DNest4/code/SamplerImpl.h:306-329, 382-399) without any GPU."""
import os
import subprocess
import sys
import textwrap

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = textwrap.dedent('''
    import os, sys
    sys.path.insert(0, %r)
    import numpy as np
    import torch.distributed as dist
    from kima_b200.dist import (TorchBarrierExchange, env_rank_world, merge_above_and_create_level,
                                shard_lanes)
    dist.init_process_group("gloo")
    rank, world, _ = env_rank_world()
    ex = TorchBarrierExchange()
    assert ex.world == 2 and ex.rank == rank
    # lanes shard contiguously and exhaustively
    lo, hi = shard_lanes(11, rank, world)
    spans = [shard_lanes(11, r, world) for r in range(world)]
    assert spans[0] == (0, 6) and spans[1] == (6, 11) and (lo, hi) == spans[rank]
    # every rank contributes a different number of (logL, tiebreaker) pairs -- rank 1 none at first
    rng = np.random.default_rng(100 + rank)
    mine = rng.normal(size=(7 if rank == 0 else 0, 2))
    allp, counts = ex.allgather_above(mine)
    assert counts == [7, 0] and allp.shape == (7, 2)
    ref0 = np.random.default_rng(100).normal(size=(7, 2))
    assert np.array_equal(allp, ref0)                      # rank-major, bit-exact
    mine = rng.normal(size=(5 + 4 * rank, 2))
    allp, counts = ex.allgather_above(mine)
    assert counts == [5, 9] and allp.shape == (14, 2)
    # the level-creation rule gives the same level on every rank
    lvl, rest = merge_above_and_create_level(allp, 10)
    assert lvl is not None and rest.shape[0] == 14 - int((1 - 1 / np.e) * 14) - 1
    chk = np.array([lvl[0], lvl[1], rest.sum()])
    import torch
    t = torch.from_numpy(chk.copy()); dist.broadcast(t, 0)
    assert np.array_equal(t.numpy(), chk)
    none, same = merge_above_and_create_level(allp, 15)
    assert none is None and same.shape == (14, 2)
    # per-level counters (accepts, tries, visits, exceeds) sum over ranks, as uint64
    c = np.arange(12, dtype=np.uint64).reshape(3, 4) * (rank + 1) + (2 ** 40 if rank else 0)
    tot = ex.allreduce_levels(c)
    assert tot.dtype == np.uint64 and np.array_equal(tot, np.arange(12, dtype=np.uint64).reshape(3, 4) * 3 + 2 ** 40)
    assert ex.max_float(float(rank)) == 1.0 and ex.sum_float(1.5) == 3.0
    ex.barrier()
    dist.destroy_process_group()
    sys.stdout.write("rank{}-ok;".format(rank))
''') % ROOT


def test_barrier_merge_world2_gloo(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(WORKER)
    env = dict(os.environ, OMP_NUM_THREADS="1")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                        "--master-addr", "127.0.0.1", "--master-port", "29517", str(script)],
                       capture_output=True, text=True, timeout=300, env=env)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "rank0-ok" in r.stdout and "rank1-ok" in r.stdout
