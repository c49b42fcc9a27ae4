"""Multi-GPU parity check, run under torchrun: sharded solve vs the same problem on one GPU, for every way the sharded
PCG can apply the reduced camera system:
  explicit tiles (reduce-scattered) or implicit operator  x  all-reduce over peer mailboxes (fused kernel) or NCCL.
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 \
      scripts/multigpu_check.py [shape] [iters]"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist
from steviapp_b200 import SBASolver, synthetic, comm_unique_id

shape = sys.argv[1] if len(sys.argv) > 1 else "ladybug"
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 4
only = sys.argv[3].split(",") if len(sys.argv) > 3 and sys.argv[3] != "all" else None
ext = len(sys.argv) > 4 and sys.argv[4] == "ext"   # free intrinsics + offsets + free lever arms + UV <-> UV factors (explicit system only)
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
fixed = dict(function_tolerance=0.0, parameter_tolerance=0.0, gradient_tolerance=0.0)
if shape.startswith("cams"):   # e.g. cams300: several tiles, band structure
    n = int(shape[4:])
    sc = synthetic.make_scene(None, n_cams=n, n_pts=100 * n, n_obs=500 * n, seed=7)
else:
    sc = synthetic.make_scene(shape)
if ext:
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
    import test_gpu_shared as tgs
    import test_pairs_host as tph
    rng = np.random.default_rng(11)
    pairs = tph.pair_scene(shape if not shape.startswith("cams") else "ladybug", n_pairs=400, outliers=0.05)
    sc = tgs.with_pose_priors(tgs.add_rig(sc, rng, offsets=True, n_mounts=2))
    sc["intrinsics"] = tgs.perturbed_intrinsics(sc, rng); sc["intr_fixed_mask"] = np.array([0x10], np.uint8)
    if not shape.startswith("cams"):
        sc["pair_cam"], sc["pair_uv"], sc["loss"] = pairs["pair_cam"], pairs["pair_uv"], "huber"
    only = [m for m in (only or ["explicit+peer", "explicit+nccl"]) if m.startswith("explicit")]
sh = synthetic.shard_scene(sc, rank, world)

r1 = None
if rank == 0:
    s1 = SBASolver(device=local, pcg_check_every=4, **fixed)
    synthetic.load_into(s1, sc)
    r1 = s1.solve(max_iterations=iters)
    s1.close()

results = []
all_ok = True
for name, op_mode, peer in (("explicit+peer", 3, "1"), ("explicit+nccl", 3, "0"), ("implicit+peer", 2, "1"), ("implicit+nccl", 2, "0")):
    if only and name not in only:
        continue
    os.environ["SBA_PEER_XCHG"] = peer
    s = SBASolver(device=local, pcg_check_every=4, operator_mode=op_mode, **fixed)
    uid = [comm_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(uid, src=0)
    s.comm_init(rank, world, uid[0])
    raxis, t, pts = synthetic.load_into(s, sh)
    s.finalize()
    info = dict(s.explicit_info(), **s.explicit_shard(), sparse=s.explicit_sparse_info()["enabled"])
    r = s.solve(max_iterations=iters)
    ra = (raxis.copy(), t.copy(), pts.copy())
    out = {"rank": rank, "shard_obs": sh["n_obs"], "cost_history": r["cost_history"], "pcg": r["pcg_iters_history"],
           "ms_total": r["ms_total"], "ms_pcg": r["ms_pcg"], "info": info}
    gathered = [None] * world
    dist.all_gather_object(gathered, out)
    # camera blocks must be bit-identical on every rank (replicated state)
    cams = torch.from_numpy(np.concatenate([ra[0].ravel(), ra[1].ravel()])).cuda()
    ref = cams.clone()
    dist.broadcast(ref, src=0)
    same = bool(torch.equal(cams, ref))
    flags = [None] * world
    dist.all_gather_object(flags, same)
    s.close()
    if rank == 0:
        rel = max(abs(a - b) / max(b, 1.0) for a, b in zip(r["cost_history"], r1["cost_history"]))
        ok = rel < 1e-8 and all(flags) and all(g["cost_history"] == r["cost_history"] for g in gathered)
        all_ok = all_ok and ok
        results.append({"mode": name, "ok": ok, "max_rel_cost_diff_vs_1gpu": rel, "replicated_cameras_bitwise_equal": flags,
                        "cost_sharded": r["cost_history"], "pcg_sharded": r["pcg_iters_history"],
                        "ms_sharded": [g["ms_total"] for g in gathered], "ms_pcg": [g["ms_pcg"] for g in gathered],
                        "info": [g["info"] for g in gathered], "shard_obs": [g["shard_obs"] for g in gathered]})
if rank == 0:
    print(json.dumps({"shape": shape, "world": world, "ok": all_ok, "cost_1gpu": r1["cost_history"], "pcg_1gpu": r1["pcg_iters_history"],
                      "ms_1gpu": r1["ms_total"], "ms_pcg_1gpu": r1["ms_pcg"], "modes": results}))
    assert all_ok
dist.destroy_process_group()
