"""Worker for tests/test_dist_gloo.py: one rank of a world_size-2 `gloo` group on CPU.

Exercises the host logic of the N>1 path: the shard partition, the MAX
all-reduce that completes the sharded occupancy update, and bench.py's max/sum aggregation.
The per-rank "kernel" is the CPU oracle (test infrastructure) because the product kernels need a GPU.
"""
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import oracle  # noqa: E402
from bench import aggregate_over_ranks, make_rays  # noqa: E402
from nerfacc_b200 import _occ_ops  # noqa: E402


def main():
    out_path = sys.argv[1]
    dist.init_process_group("gloo")
    world, rank = _occ_ops.dist_info()
    assert (world, rank) == (int(os.environ["WORLD_SIZE"]), int(os.environ["RANK"]))

    # the same cells, offsets and field on every rank (same seed), as OccGridEstimator._update does
    rng = np.random.default_rng(77)
    n_cells, n = 4096, 3000
    occs0 = rng.random(n_cells).astype(np.float32)
    occs0[rng.random(n_cells) < 0.1] = -1.0  # invisible cells: never sampled
    visible = np.flatnonzero(occs0 >= 0)
    cell_ids = rng.choice(visible, n, replace=True).astype(np.int64)  # duplicates on purpose
    field = rng.random(n).astype(np.float32) * 2
    decay = 0.95

    # single-process answer
    want = oracle.occ_ema_update(occs0.copy(), cell_ids, field, decay)

    # sharded: every rank decays the full sample but max-es in only its slice of evaluations
    lo, hi = _occ_ops.shard_range(n, world, rank)
    masked = np.full(n, -np.inf, np.float32)
    masked[lo:hi] = field[lo:hi]
    mine = oracle.occ_ema_update(occs0.copy(), cell_ids, masked, decay)
    t = torch.from_numpy(mine)
    _occ_ops.all_reduce_max(t)
    ok_update = bool(np.array_equal(t.numpy(), want))

    # partition covers [0, n) exactly once, also when n is not divisible / smaller than world
    parts_ok = True
    for m in (0, 1, 2, 3, 7, 3000, 3001):
        spans = [_occ_ops.shard_range(m, world, r) for r in range(world)]
        covered = sum(h - l for l, h in spans)
        parts_ok &= covered == m and spans[0][0] == 0 and spans[-1][1] == m
        parts_ok &= all(spans[i][1] == spans[i + 1][0] for i in range(world - 1))

    # bench aggregation: MAX of times, SUM of units
    ms, units = aggregate_over_ranks(10.0 + rank, 1000 * (rank + 1), dist, torch.device("cpu"))
    # per-rank ray blocks differ (seeded by rank) — weak scaling, no shared rays
    o, _ = make_rays(64, 1234 + rank)
    gathered = [torch.zeros(64, 3) for _ in range(world)]
    dist.all_gather(gathered, torch.from_numpy(o))
    distinct = not torch.equal(gathered[0], gathered[1])

    if rank == 0:
        with open(out_path, "w") as f:
            json.dump({"update": ok_update, "parts": bool(parts_ok), "ms": ms, "units": units,
                       "distinct": bool(distinct), "world": world}, f)
    dist.barrier()
    dist.destroy_process_group()
    if not (ok_update and parts_ok):
        sys.exit(1)


if __name__ == "__main__":
    main()
