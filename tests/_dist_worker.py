"""Worker for tests/test_dist_gloo.py: one rank of a world_size-2 `gloo` group on CPU.

Exercises the host logic of the N>1 path: the shard partition, the MAX
all-reduce that completes the sharded occupancy update, bench.py's max/sum aggregation, and the
product's own `OccGridEstimator._update` (shard_updates opt-in: private generator, slice of the
evaluations, all-reduce on the chosen group) against the same updates done by one rank.
The per-rank "kernels" are CPU stand-ins (the oracle + torch; test infrastructure) because the
product kernels need a GPU — only the three device calls of _occ_ops are swapped, every line of
`_update` and `_sample_uniform_and_occupied_cells` is the product's.
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


def _cpu_standins():
    """CPU versions of the three device calls `_update` makes (nfb_occ_cell_points,
    nfb_occ_ema_update, nfb_occ_binarize), same semantics."""
    def cell_points(cells, u, resolution, aabb):
        resolution = torch.as_tensor(resolution)  # host ints from `_update`, or a tensor
        res = resolution.to(torch.long)
        ijk = torch.stack([cells // (res[1] * res[2]), (cells // res[2]) % res[1], cells % res[2]], -1)
        g = (ijk.float() + u) / resolution.float()
        return aabb[:3] + g * (aabb[3:] - aabb[:3])

    def ema_update(occs, cell_ids, occ, lo, hi, decay):
        n = cell_ids.shape[0]
        if n == 0:
            return
        field = np.full(n, -np.inf, np.float32)
        field[lo:hi] = occ.numpy()
        occs.copy_(torch.from_numpy(oracle.occ_ema_update(occs.numpy().copy(), cell_ids.numpy(), field, decay)))

    def binarize(occs, occ_thre):
        vis = occs[occs >= 0]
        return occs > torch.clamp(vis.double().mean().float(), max=occ_thre)

    return cell_points, ema_update, binarize


def estimator_update_matches_single_rank(world, rank):
    """OccGridEstimator._update through the world > 1 branch == the same updates on one rank."""
    import nerfacc_b200 as nf

    _occ_ops.cell_points, _occ_ops.ema_update, _occ_ops.binarize = _cpu_standins()

    def occ_eval_fn(x):  # smooth in x, so different jitter offsets give different values
        r = x.norm(dim=-1, keepdim=True)
        return 0.05 * torch.exp(-((r - 0.85) / 0.1) ** 2)

    def make():
        est = nf.OccGridEstimator([-1.0, -1.0, -1.0, 1.0, 1.0, 1.0], resolution=16, levels=2)
        est.occs[::7] = -1.0  # some invisible cells
        return est

    sharded = make().shard_updates(None, seed=5)
    single = make().shard_updates(None, seed=5, _world_rank=(1, 0))
    assert sharded._shard[1:3] == (world, rank)
    ok = True
    for step in (0, 16, 256, 272, 288):  # two warm-up updates (all cells), three steady-state ones
        sharded._update(step, occ_eval_fn, warmup_steps=256)
        single._update(step, occ_eval_fn, warmup_steps=256)
        ok &= torch.equal(sharded.occs, single.occs) and torch.equal(sharded.binaries, single.binaries)
    ok &= bool(single.binaries.any()) and not bool(single.binaries.all())
    # and every rank holds the same grid
    both = [torch.zeros_like(sharded.occs) for _ in range(world)]
    dist.all_gather(both, sharded.occs)
    ok &= all(torch.equal(both[0], b) for b in both[1:])
    # default (never opted in): rank-local update, no collective, global RNG
    local = make()
    torch.manual_seed(100 + rank)
    local._update(300, occ_eval_fn, warmup_steps=256)
    assert local._shard is None
    return bool(ok)


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

    ok_est = estimator_update_matches_single_rank(world, rank)

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
            json.dump({"update": ok_update, "estimator": ok_est, "parts": bool(parts_ok), "ms": ms, "units": units,
                       "distinct": bool(distinct), "world": world}, f)
    dist.barrier()
    dist.destroy_process_group()
    if not (ok_update and parts_ok and ok_est):
        sys.exit(1)


if __name__ == "__main__":
    main()
