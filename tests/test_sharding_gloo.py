"""World-size-2 gloo test of the point-sharding logic behind the multi-GPU path, on CPU.

Each rank shards the scene exactly as bench.py does (steviapp_b200.synthetic.shard_scene over
sba_partition_points), evaluates ITS shard with the oracle, and the partial camera blocks / operator
products are all-reduced over gloo -- the identities the library relies on when it all-reduces the same
quantities over NCCL: sum of shard partials == global quantity; point blocks need no communication."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import ROOT


def _worker(rank, world, port, out):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import oracle_py
        from steviapp_b200 import synthetic
        sc = synthetic.make_scene("small", seed=77)
        # a camera prior, replicated on every rank (must be counted once)
        sc["prior_kind"] = np.concatenate([sc["prior_kind"], [2]]).astype(np.int32)
        sc["prior_index"] = np.concatenate([sc["prior_index"], [3]]).astype(np.int32)
        sc["prior_A"] = np.concatenate([sc["prior_A"], (np.eye(3) * 10).reshape(1, 9)])
        sc["prior_b"] = np.concatenate([sc["prior_b"], sc["t_true"][3:4]])
        glob = oracle_py.Problem(sc)
        g_blocks = glob.normal_blocks()
        g_cost = glob.evaluate()["cost"]
        rng = np.random.default_rng(4)
        lam_c = rng.uniform(0.5, 2, (sc["n_cams"], 6)); lam_p = rng.uniform(0.5, 2, (sc["n_pts"], 3))
        x = rng.normal(size=(sc["n_cams"], 6))
        g_y, g_rhs = glob.schur_matvec(lam_c, lam_p, x)

        sh = synthetic.shard_scene(sc, rank, world)
        lo, hi = sh["point_range"]
        # the shard keeps every camera prior; only rank 0 counts them (the library adds them after the all-reduce)
        strip = dict(sh)
        if rank != 0:
            keep = sh["prior_kind"] == 0
            for k in ("prior_kind", "prior_index", "prior_A", "prior_b"):
                strip[k] = np.ascontiguousarray(sh[k][keep])
        loc = oracle_py.Problem(strip)
        b = loc.normal_blocks()
        cost = torch.tensor([loc.evaluate()["cost"]], dtype=torch.float64)
        gc = torch.from_numpy(b["gc"].copy()); Hcc = torch.from_numpy(b["Hcc"].copy())
        y, rhs = loc.schur_matvec(lam_c, lam_p[lo:hi], x)
        y_t = torch.from_numpy(y.copy()); rhs_t = torch.from_numpy(rhs.copy())
        n_obs = torch.tensor([sh["n_obs"]], dtype=torch.int64)
        for tns in (cost, gc, Hcc, y_t, rhs_t, n_obs):
            dist.all_reduce(tns)
        rel = lambda a, r: float(np.abs(a - r).max() / max(np.abs(r).max(), 1.0))
        # every rank's Hcc x + lam x term was summed `world` times; the library adds them once, after the all-reduce
        y_fix = y_t.numpy() - (world - 1) * lam_c * x
        res = {
            "n_obs": int(n_obs.item()) == sc["n_obs"],
            "cost": abs(cost.item() - g_cost) <= 1e-12 * g_cost,
            "gc": rel(gc.numpy(), g_blocks["gc"]), "Hcc": rel(Hcc.numpy(), g_blocks["Hcc"]),
            "gp": rel(b["gp"], g_blocks["gp"][lo:hi]), "Hpp": rel(b["Hpp"], g_blocks["Hpp"][lo:hi]),
            "y": rel(y_fix, g_y), "rhs": rel(rhs_t.numpy(), g_rhs),
            "range": (lo, hi),
        }
        out[rank] = res
    finally:
        dist.destroy_process_group()


def test_point_sharding_world2(oracle, sba_lib):
    world = 2
    mgr = mp.Manager()
    out = mgr.dict()
    port = 29500 + os.getpid() % 2000
    mp.spawn(_worker, args=(world, port, out), nprocs=world, join=True)
    assert len(out) == world
    ranges = sorted(out[r]["range"] for r in range(world))
    assert ranges[0][0] == 0 and ranges[0][1] == ranges[1][0]
    for r in range(world):
        o = out[r]
        assert o["n_obs"] and o["cost"], o
        for k in ("gc", "Hcc", "gp", "Hpp", "rhs"):
            assert o[k] < 1e-11, (k, o[k])
        assert o["y"] < 1e-10, o["y"]
