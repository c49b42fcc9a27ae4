"""Deterministic synthetic bundle-adjustment scenes (SURVEY 8(d) generator).

Cameras on a ring looking inward with random roll, points in a cylinder around
the centre, every point observed by k cameras drawn from a window of the ring
nearest to its azimuth (banded co-visibility).  Pixels come from the FORWARD
lens model, i.e. the numerical inverse of PinholeUVProjector::dirFromUV
(sbamodules/pinholecameraprojectormodule.h:28-71, including the `B := ts`
quirk), plus Gaussian noise.  Pure numpy; used by tests, bench.py and smoke().
"""
import numpy as np

SHAPES = {
    # name: (n_cams, n_pts, n_obs)
    "tiny": (6, 60, 300),
    "small": (16, 800, 3600),
    "ladybug": (49, 7776, 31843),
    "venice": (1778, 993923, 5001946),
    "city": (10000, 10_000_000, 50_000_000),
}

DEFAULT_INTRINSICS = np.array([2000.0, 2000.0, 1500.0, -0.05, 0.01, 0.0, 1e-4, -1e-4, 0.0, 0.0])


def dir_from_uv(uv, intr):
    """numpy restatement of dirFromUV, only for the generator's forward-model inversion check."""
    f, ppx, ppy, k1, k2, k3, t1, t2 = intr[:8]
    y = (uv[:, 1] - ppy) / f
    x = (uv[:, 0] - ppx - t2 * y) / (f + t1)
    r2 = x * x + y * y
    rad = r2 * (k1 + r2 * (k2 + r2 * k3))
    xy = x * y
    cx = x + x * rad + 2 * xy * t1 + (r2 + 2 * x * x) * t2
    cy = y + y * rad + (r2 + 2 * y * y) * t1 + 2 * xy * t2
    return np.stack([cx * f, cy * f, np.full_like(cx, f)], axis=1)


def uv_from_normalized(xn, intr, iters=25):
    """Forward lens model: find (x,y) with corrected(x,y) = xn by Newton, then apply the skew (B := ts)."""
    f, ppx, ppy, k1, k2, k3, t1, t2 = intr[:8]
    x = xn[:, 0].copy()
    y = xn[:, 1].copy()
    for _ in range(iters):
        r2 = x * x + y * y
        rad = r2 * (k1 + r2 * (k2 + r2 * k3))
        drad = k1 + r2 * (2 * k2 + 3 * k3 * r2)          # d rad / d r2
        cx = x + x * rad + 2 * x * y * t1 + (r2 + 2 * x * x) * t2
        cy = y + y * rad + (r2 + 2 * y * y) * t1 + 2 * x * y * t2
        # Jacobian of (cx, cy) w.r.t. (x, y)
        a = 1 + rad + x * drad * 2 * x + 2 * y * t1 + 6 * x * t2
        b = x * drad * 2 * y + 2 * x * t1 + 2 * y * t2
        c = y * drad * 2 * x + 2 * x * t1 + 2 * y * t2
        d = 1 + rad + y * drad * 2 * y + 6 * y * t1 + 2 * x * t2
        e0 = cx - xn[:, 0]
        e1 = cy - xn[:, 1]
        det = a * d - b * c
        dx = (d * e0 - b * e1) / det
        dy = (-c * e0 + a * e1) / det
        x -= dx
        y -= dy
        if max(np.abs(dx).max(initial=0.0), np.abs(dy).max(initial=0.0)) < 1e-16:
            break
    u = (f + t1) * x + t2 * y + ppx
    v = f * y + ppy
    return np.stack([u, v], axis=1)


def rotmat_to_axis_angle(R):
    """(n,3,3) -> (n,3); valid away from theta = 0 and pi (callers re-draw those)."""
    tr = np.clip((np.trace(R, axis1=1, axis2=2) - 1.0) * 0.5, -1.0, 1.0)
    th = np.arccos(tr)
    w = np.stack([R[:, 2, 1] - R[:, 1, 2], R[:, 0, 2] - R[:, 2, 0], R[:, 1, 0] - R[:, 0, 1]], axis=1)
    s = 2.0 * np.sin(th)
    return w / s[:, None] * th[:, None], th


def axis_angle_to_rotmat(r):
    th = np.linalg.norm(r, axis=1)
    k = r / np.maximum(th, 1e-300)[:, None]
    K = np.zeros((r.shape[0], 3, 3))
    K[:, 0, 1] = -k[:, 2]; K[:, 0, 2] = k[:, 1]
    K[:, 1, 0] = k[:, 2]; K[:, 1, 2] = -k[:, 0]
    K[:, 2, 0] = -k[:, 1]; K[:, 2, 1] = k[:, 0]
    I = np.eye(3)[None]
    return I + np.sin(th)[:, None, None] * K + (1 - np.cos(th))[:, None, None] * (K @ K)


def make_scene(shape="small", seed=None, n_cams=None, n_pts=None, n_obs=None, noise_px=0.5,
               init_noise=(0.01, 0.05, 0.1), mixed=False, spherical=False, shuffle=None,
               ring_radius=30.0, cloud_radius=10.0, cloud_height=5.0, window=None, far_fraction=0.3,
               intrinsics=None):
    """Returns a dict of numpy arrays describing one BA problem (truth, initial values, observations, priors)."""
    if shape is not None:
        nc, npt, no = SHAPES[shape]
        n_cams = n_cams or nc
        n_pts = n_pts or npt
        n_obs = n_obs or no
    if seed is None:
        seed = 42 + (list(SHAPES).index(shape) if shape in SHAPES else 0)
    rng = np.random.Generator(np.random.PCG64(seed))
    intr = np.array(DEFAULT_INTRINSICS if intrinsics is None else intrinsics, dtype=np.float64).reshape(-1, 10)

    # ---- cameras: ring, looking at the centre, random roll; sensor->world pose
    ang = 2 * np.pi * np.arange(n_cams) / n_cams
    centre = np.stack([ring_radius * np.cos(ang), ring_radius * np.sin(ang),
                       rng.uniform(-1.0, 1.0, n_cams)], axis=1)
    fwd = -centre / np.linalg.norm(centre, axis=1)[:, None]
    up0 = np.array([0.0, 0.0, 1.0])
    right = np.cross(fwd, up0)
    right /= np.linalg.norm(right, axis=1)[:, None]
    down = np.cross(fwd, right)
    raxis = np.zeros((n_cams, 3))
    todo = np.arange(n_cams)
    for _ in range(64):
        roll = rng.uniform(0, 2 * np.pi, todo.size)
        cx = np.cos(roll)[:, None] * right[todo] + np.sin(roll)[:, None] * down[todo]
        cy = -np.sin(roll)[:, None] * right[todo] + np.cos(roll)[:, None] * down[todo]
        R = np.stack([cx, cy, fwd[todo]], axis=2)          # columns = camera axes in world
        r, th = rotmat_to_axis_angle(R)
        good = (th > 0.05) & (th < np.pi - 0.05)
        raxis[todo[good]] = r[good]
        todo = todo[~good]
        if todo.size == 0:
            break
    assert todo.size == 0, "could not draw rolls away from theta = 0 / pi"
    t = centre.copy()

    # ---- points: cylinder around the origin
    rho = cloud_radius * np.sqrt(rng.uniform(0, 1, n_pts))
    phi = rng.uniform(0, 2 * np.pi, n_pts)
    X = np.stack([rho * np.cos(phi), rho * np.sin(phi), rng.uniform(-cloud_height, cloud_height, n_pts)], axis=1)

    # ---- visibility: k_p >= 2 views, exact total n_obs, cameras from a ring window at the point's azimuth
    mean_k = n_obs / n_pts
    if window is None:
        window = int(min(n_cams, max(8, 4 * np.ceil(mean_k))))
    window = min(window, n_cams)
    k = 2 + rng.poisson(max(mean_k - 2.0, 0.0), n_pts)
    k = np.clip(k, 2, window)
    diff = int(n_obs - k.sum())
    while diff != 0:
        step = 1 if diff > 0 else -1
        cand = np.flatnonzero((k + step >= 2) & (k + step <= window))
        pick = rng.choice(cand, size=min(abs(diff), cand.size), replace=False)
        k[pick] += step
        diff = int(n_obs - k.sum())
    i0 = np.floor(phi / (2 * np.pi) * n_cams).astype(np.int64)
    order = np.argsort(rng.random((n_pts, window), dtype=np.float32), axis=1).astype(np.int32)
    take = np.arange(window)[None, :] < k[:, None]
    obs_pt = np.repeat(np.arange(n_pts, dtype=np.int32), k)
    off = order[take].astype(np.int64) - window // 2       # row-major: point-major order
    # long-range links: a fraction of the views comes from anywhere on the ring (every camera looks at
    # the centre, so they are geometrically valid).  View j of a point draws from the j-th stratum of
    # the ring outside its window, so the cameras of one point stay distinct.
    far_len = n_cams - window - 1
    strat = far_len // window if window > 0 else 0
    if far_fraction > 0 and strat >= 1:
        col = np.broadcast_to(np.arange(window, dtype=np.int64), (n_pts, window))[take]
        is_far = rng.random(off.size) < far_fraction
        far_off = (window - window // 2) + col * strat + np.floor(rng.random(off.size) * strat).astype(np.int64)
        off = np.where(is_far, far_off, off)
    obs_cam = ((i0[obs_pt] + off) % n_cams).astype(np.int32)
    del order, take

    # ---- pixels through the forward model
    Rw = axis_angle_to_rotmat(raxis)                       # sensor->world
    v = X[obs_pt] - t[obs_cam]
    Pbar = np.einsum("nji,nj->ni", Rw[obs_cam], v)         # R^T v
    assert (Pbar[:, 2] > 0).all()
    xn = Pbar[:, :2] / Pbar[:, 2:3]
    uv = uv_from_normalized(xn, intr[0])
    uv += rng.normal(0.0, noise_px, uv.shape)

    if shuffle is None:
        shuffle = n_obs <= 200_000
    if shuffle:
        perm = rng.permutation(n_obs)
        obs_pt, obs_cam, uv = obs_pt[perm], obs_cam[perm], uv[perm]

    # ---- initial values
    sr, st, sx = init_noise
    raxis0 = raxis + rng.normal(0, sr, raxis.shape)
    t0 = t + rng.normal(0, st, t.shape)
    X0 = X + rng.normal(0, sx, X.shape)
    cam_fixed = np.zeros(n_cams, dtype=np.uint8)
    cam_fixed[0] = 1                                       # gauge: first camera constant
    raxis0[0] = raxis[0]
    t0[0] = t[0]
    # 3-D priors sigma = 0.01 on 4 points (GCP-style, landmarkssbamodule.cpp:171-202: stiffness = 1/sigma)
    gcp = np.linspace(0, n_pts - 1, 4).astype(np.int32)
    prior_kind = np.zeros(4, dtype=np.int32)
    prior_A = np.tile((np.eye(3) / 0.01).reshape(1, 9), (4, 1))
    prior_b = X[gcp].copy()

    if mixed:
        cost_kind = (np.arange(n_cams) % 2).astype(np.uint8)
    else:
        cost_kind = np.full(n_cams, 1 if spherical else 0, dtype=np.uint8)

    return {
        "n_cams": n_cams, "n_pts": n_pts, "n_obs": n_obs, "seed": seed, "window": window,
        "intrinsics": intr.copy(), "cam_model": np.zeros(n_cams, dtype=np.int32),
        "cam_fixed": cam_fixed, "cost_kind": cost_kind, "pt_fixed": np.zeros(n_pts, dtype=np.uint8),
        "raxis_true": raxis, "t_true": t, "pts_true": X,
        "raxis": np.ascontiguousarray(raxis0), "t": np.ascontiguousarray(t0), "pts": np.ascontiguousarray(X0),
        "obs_cam": np.ascontiguousarray(obs_cam), "obs_pt": np.ascontiguousarray(obs_pt),
        "obs_uv": np.ascontiguousarray(uv), "obs_info": None,
        "prior_kind": prior_kind, "prior_index": gcp, "prior_A": prior_A, "prior_b": prior_b,
    }


def load_into(solver, scene, copy=True):
    """Registers a scene on an SBASolver.  Returns the (raxis, t, pts) arrays the solver updates in place."""
    raxis = scene["raxis"].copy() if copy else scene["raxis"]
    t = scene["t"].copy() if copy else scene["t"]
    pts = scene["pts"].copy() if copy else scene["pts"]
    if scene.get("intr_fixed_mask") is not None:      # free intrinsic blocks: the array is updated in place (solver.intrinsics)
        solver.intrinsics = np.ascontiguousarray(scene["intrinsics"], dtype=np.float64).reshape(-1, 10).copy()
        solver.set_intrinsics_blocks(solver.intrinsics, scene["intr_fixed_mask"])
        if scene.get("intr_prior_model") is not None:
            solver.add_intrinsics_priors(scene["intr_prior_model"], scene["intr_prior_block"], scene["intr_prior_A"], scene["intr_prior_b"])
    else:
        solver.set_intrinsics(scene["intrinsics"])
    solver.set_poses(raxis, t, scene["cam_model"], scene["cam_fixed"], scene.get("cost_kind"))
    if scene.get("cam_offset") is not None:
        solver.set_sensor_offsets(scene["cam_offset"])
    if scene.get("mount_b") is not None:
        solver.mount_b = np.ascontiguousarray(scene["mount_b"], dtype=np.float64).reshape(-1, 3).copy()
        solver.mount_l = np.ascontiguousarray(scene["mount_l"], dtype=np.float64).reshape(-1, 3).copy()
        solver.set_lever_arms(solver.mount_b, solver.mount_l, scene["cam_mount"], scene.get("mount_fixed"))
    solver.set_points(pts, scene["pt_fixed"])
    solver.add_observations(scene["obs_cam"], scene["obs_pt"], scene["obs_uv"], scene["obs_info"])
    if scene.get("pair_cam") is not None and len(scene["pair_cam"]):
        pc = np.asarray(scene["pair_cam"], dtype=np.int32).reshape(-1, 2); puv = np.asarray(scene["pair_uv"], dtype=np.float64).reshape(-1, 4)
        pinfo = None if scene.get("pair_info") is None else np.asarray(scene["pair_info"], dtype=np.float64).reshape(-1, 8)
        solver.add_uv_pairs(pc[:, 0], pc[:, 1], puv[:, :2], puv[:, 2:], None if pinfo is None else pinfo[:, :4], None if pinfo is None else pinfo[:, 4:])
        if scene.get("loss"):
            solver.set_pair_loss(scene["loss"], scene.get("loss_a", 1.0))
    if scene["prior_kind"].size:
        solver.add_priors(scene["prior_kind"], scene["prior_index"], scene["prior_A"], scene["prior_b"])
    return raxis, t, pts


def shard_scene(scene, rank, nranks, begin=None):
    """Point-sharded view of a scene for one rank: contiguous point range balanced by
    observation count, all its observations, its point priors; cameras and camera priors replicated."""
    from .solver import partition_points
    n_pts = scene["n_pts"]
    if begin is None:
        opp = np.bincount(scene["obs_pt"], minlength=n_pts).astype(np.int32)
        begin = partition_points(opp, nranks)
    lo, hi = int(begin[rank]), int(begin[rank + 1])
    sel = (scene["obs_pt"] >= lo) & (scene["obs_pt"] < hi)
    out = dict(scene)
    out["n_pts"] = hi - lo
    out["pts"] = np.ascontiguousarray(scene["pts"][lo:hi])
    if "pts_true" in scene:   # synthetic scenes only (a loaded BAL file has no ground truth)
        out["pts_true"] = scene["pts_true"][lo:hi]
    out["pt_fixed"] = np.ascontiguousarray(scene["pt_fixed"][lo:hi])
    out["obs_cam"] = np.ascontiguousarray(scene["obs_cam"][sel])
    out["obs_pt"] = np.ascontiguousarray(scene["obs_pt"][sel] - lo)
    out["obs_uv"] = np.ascontiguousarray(scene["obs_uv"][sel])
    out["obs_info"] = None if scene["obs_info"] is None else np.ascontiguousarray(scene["obs_info"][sel])
    out["n_obs"] = int(sel.sum())
    pk, pi = scene["prior_kind"], scene["prior_index"]
    keep = (pk != 0) | ((pi >= lo) & (pi < hi))
    out["prior_kind"] = np.ascontiguousarray(pk[keep])
    out["prior_index"] = np.ascontiguousarray(np.where(pk[keep] == 0, pi[keep] - lo, pi[keep]).astype(np.int32))
    out["prior_A"] = np.ascontiguousarray(scene["prior_A"][keep])
    out["prior_b"] = np.ascontiguousarray(scene["prior_b"][keep])
    out["point_range"] = (lo, hi)
    return out
