"""Seeded synthetic inputs of the shapes named in BASELINE.json `configs`.

No datasets or droid.pth exist offline, so tests and bench.py use these.  Pure numpy float64 geometry
written independently of both the oracle and the CUDA kernels (rotation matrices, not the quaternion
formulas of the reference), results cast to float32.  Poses are world->camera (tx,ty,tz,qx,qy,qz,qw).
"""
import numpy as np

TARTAN_INTR = np.array([32.0, 32.0, 32.0, 24.0], np.float32)  # 0.8*[320,320,320,240]/8, 48x64 grid
EUROC_INTR = np.array([435.2 * 512 / 752, 435.2 * 320 / 480, 367.45 * 512 / 752, 252.2 * 320 / 480],
                      np.float64) / 8.0                          # evaluation_scripts/test_euroc.py:51,70-74
EUROC_INTR = EUROC_INTR.astype(np.float32)


def _hat(w):
    return np.array([[0, -w[2], w[1]], [w[2], 0, -w[0]], [-w[1], w[0], 0]], np.float64)


def se3_exp(xi):
    """xi = (tau, phi) -> (R, t), float64."""
    tau, phi = np.asarray(xi[:3], np.float64), np.asarray(xi[3:], np.float64)
    th = np.linalg.norm(phi)
    W = _hat(phi)
    if th < 1e-9:
        R = np.eye(3) + W
        V = np.eye(3) + 0.5 * W
    else:
        R = np.eye(3) + np.sin(th) / th * W + (1 - np.cos(th)) / th ** 2 * W @ W
        V = np.eye(3) + (1 - np.cos(th)) / th ** 2 * W + (th - np.sin(th)) / th ** 3 * W @ W
    return R, V @ tau


def rot_to_quat(R):
    w = np.sqrt(max(0.0, 1 + R[0, 0] + R[1, 1] + R[2, 2])) / 2
    if w > 1e-6:
        x = (R[2, 1] - R[1, 2]) / (4 * w)
        y = (R[0, 2] - R[2, 0]) / (4 * w)
        z = (R[1, 0] - R[0, 1]) / (4 * w)
    else:  # 180-degree rotation; not reached by the trajectories below
        x = np.sqrt(max(0.0, 1 + R[0, 0] - R[1, 1] - R[2, 2])) / 2
        y = np.sqrt(max(0.0, 1 - R[0, 0] + R[1, 1] - R[2, 2])) / 2
        z = np.sqrt(max(0.0, 1 - R[0, 0] - R[1, 1] + R[2, 2])) / 2
    return np.array([x, y, z, w])


def pose_vec(R, t):
    return np.concatenate([t, rot_to_quat(R)])


def trajectory(n, loop=250, step=(0.05, 0.0, 0.01, 0.0, 0.01, 0.0)):
    """Smooth world->camera trajectory that returns to its start every `loop` keyframes, so that
    proximity (loop-closure) edges exist for the global-BA configuration."""
    Rs, ts = [], []
    step = np.asarray(step, np.float64)
    for k in range(n):
        R, t = se3_exp(step * (k % loop))
        Rs.append(R); ts.append(t)
    return np.stack(Rs), np.stack(ts)


def gt_disps(n, ht, wd, rng):
    y, x = np.meshgrid(np.arange(ht), np.arange(wd), indexing="ij")
    ph = rng.uniform(0, 2 * np.pi, size=n)
    return np.stack([0.5 + 0.3 * np.sin(x / 10.0 + p) * np.cos(y / 7.0) for p in ph])


def reproject(Rs, ts, disps, intr, ii, jj):
    """Pixel grid of frame ii[e] -> frame jj[e] (ii==jj: stereo baseline (-0.1,0,0)), float64 [E,2,ht,wd]."""
    fx, fy, cx, cy = [float(v) for v in intr]
    n, ht, wd = disps.shape
    y, x = np.meshgrid(np.arange(ht, dtype=np.float64), np.arange(wd, dtype=np.float64), indexing="ij")
    X = np.stack([(x - cx) / fx, (y - cy) / fy, np.ones_like(x)], 0).reshape(3, -1)
    out = np.empty((len(ii), 2, ht, wd), np.float64)
    for e, (i, j) in enumerate(zip(ii, jj)):
        if i == j:
            Rij, tij = np.eye(3), np.array([-0.1, 0.0, 0.0])
        else:
            Rij = Rs[j] @ Rs[i].T
            tij = ts[j] - Rij @ ts[i]
        Y = Rij @ X + tij[:, None] * disps[i].reshape(1, -1)
        z = np.maximum(Y[2], 1e-3)
        out[e, 0] = (fx * Y[0] / z + cx).reshape(ht, wd)
        out[e, 1] = (fy * Y[1] / z + cy).reshape(ht, wd)
    return out


def edges_radius(n, rad, first=0):
    ii, jj = [], []
    for i in range(first, n):
        for j in range(first, n):
            if i != j and abs(i - j) <= rad:
                ii.append(i); jj.append(j)
    return ii, jj


def edges_global(Rs, ts, rad=2, knn=8, max_factors=None):
    """Temporal radius-`rad` neighbours plus, for every keyframe, its `knn` nearest-in-space
    non-adjacent keyframes, both directions (what add_proximity_factors produces in shape)."""
    n = len(ts)
    centers = -np.einsum("nij,ni->nj", Rs, ts)
    pairs = set()
    for i in range(n):
        for j in range(max(0, i - rad), min(n, i + rad + 1)):
            if i != j:
                pairs.add((i, j))
    d = np.linalg.norm(centers[:, None] - centers[None], axis=-1)
    for i in range(n):
        dd = d[i].copy()
        dd[max(0, i - rad):i + rad + 1] = np.inf
        for j in np.argsort(dd, kind="stable")[:knn]:
            if np.isfinite(dd[j]):
                pairs.add((i, int(j))); pairs.add((int(j), i))
    pairs = sorted(pairs)
    if max_factors is not None:
        pairs = pairs[:max_factors]
    return [p[0] for p in pairs], [p[1] for p in pairs]


def make_ba_problem(name="C1", seed=1234, n_frames=None, stereo=False, sensor=False, nbuf=None, knn=8, loop=None,
                    shape=None):
    """Returns a dict of float32/int64 numpy arrays: poses, disps, intrinsics, disps_sens, targets,
    weights, eta, ii, jj, t0, t1 (+ gt_poses, gt_disps).  C4 with n_frames=1000, knn=16 is the
    BASELINE.json backend configuration: 1000 keyframes, 20 862 edges (~20k), 40x64."""
    rng = np.random.default_rng(seed)
    if name == "C1":      # 8 frames, 24 edges, 48x64
        n, ht, wd, intr = 8, 48, 64, TARTAN_INTR
        ii, jj = edges_radius(n, 2)
        keep = [(a, b) for a, b in zip(ii, jj) if (a, b) not in ((5, 7), (7, 5))]
        ii, jj = [k[0] for k in keep], [k[1] for k in keep]
        loop = 10 ** 6
    elif name == "C3":    # frontend window: 12 KF, 60 edges, 48x64
        n, ht, wd, intr = 12, 48, 64, TARTAN_INTR
        ii, jj = edges_radius(n, 2)
        extra = sorted([(a, b) for a in range(n) for b in range(n) if abs(a - b) >= 3],
                       key=lambda p: (-abs(p[0] - p[1]), p))[:18]
        ii += [p[0] for p in extra]; jj += [p[1] for p in extra]
        loop = 10 ** 6
    elif name == "C4":    # backend: n_frames KF (default 1000), ~20 edges per KF, 40x64
        n = n_frames or 1000
        ht, wd, intr = 40, 64, EUROC_INTR
        loop = loop or 250          # keyframes per lap (a larger value than n_frames: an open trajectory, no loop closures)
        ii = jj = None
    else:
        raise ValueError(name)
    if n_frames is not None and name != "C4":
        raise ValueError("n_frames only applies to C4")
    if shape is not None:      # another grid (e.g. 41x73: a 16:9 video at 1/8 resolution); intrinsics scaled along
        intr = intr * np.array([shape[1] / wd, shape[0] / ht, shape[1] / wd, shape[0] / ht])
        ht, wd = int(shape[0]), int(shape[1])
    Rs, ts = trajectory(n, loop=loop)
    if ii is None:
        ii, jj = edges_global(Rs, ts, rad=2, knn=knn, max_factors=None)
    if stereo:
        ii = list(ii) + list(range(n)); jj = list(jj) + list(range(n))
    ii = np.asarray(ii, np.int64); jj = np.asarray(jj, np.int64)
    nb = nbuf or n
    gd = gt_disps(n, ht, wd, rng)
    target = reproject(Rs, ts, gd, intr, ii, jj) + rng.normal(0, 0.25, size=(len(ii), 2, ht, wd))
    weight = rng.uniform(0, 1, size=(len(ii), 2, ht, wd))
    weight[rng.uniform(size=weight.shape) < 0.1] = 0.0
    gt_poses = np.stack([pose_vec(R, t) for R, t in zip(Rs, ts)])
    poses = np.zeros((nb, 7)); poses[:, 6] = 1.0
    for k in range(n):
        if k < 2:
            poses[k] = gt_poses[k]
        else:
            dR, dt = se3_exp(rng.normal(0, 0.01, size=6))
            poses[k] = pose_vec(dR @ Rs[k], dR @ ts[k] + dt)
    disps = np.ones((nb, ht, wd))
    disps[:n] = np.clip(gd * (1 + rng.normal(0, 0.05, size=gd.shape)), 0.05, None)
    disps_sens = np.zeros((nb, ht, wd))
    if sensor:
        m = rng.uniform(size=(n, ht, wd)) < 0.5
        disps_sens[:n] = np.where(m, gd * (1 + rng.normal(0, 0.02, size=gd.shape)), 0.0)
    t0, t1 = max(1, int(ii.min()) + 1), int(max(ii.max(), jj.max())) + 1
    kx = np.unique(np.concatenate([np.arange(t0, t1), ii]))
    eta = rng.uniform(1e-4, 1e-2, size=(len(kx), ht, wd))
    f = np.float32
    return dict(poses=poses.astype(f), disps=disps.astype(f), intrinsics=intr.astype(f),
                disps_sens=disps_sens.astype(f), targets=target.astype(f), weights=weight.astype(f),
                eta=eta.astype(f), ii=ii, jj=jj, t0=t0, t1=t1, gt_poses=gt_poses.astype(f),
                gt_disps=gd.astype(f), ht=ht, wd=wd, n=n)


def make_corr_problem(n_edges=24, ht=48, wd=64, dim=128, seed=1234, oob_frac=0.0, dtype=np.float16):
    """Feature maps (smooth random fields, unit-ish variance like instance-normed fnet output) and
    lookup coordinates = pixel grid + smooth flow in [-8, 8] px (+ optionally a fraction far outside)."""
    rng = np.random.default_rng(seed)
    n_frames = max(2, n_edges // 4 + 3)
    base = rng.normal(size=(n_frames, dim, (ht + 3) // 4, (wd + 3) // 4))
    fm = np.repeat(np.repeat(base, 4, axis=2), 4, axis=3)[:, :, :ht, :wd] + 0.5 * rng.normal(size=(n_frames, dim, ht, wd))
    ii = np.array([e % (n_frames - 2) for e in range(n_edges)], np.int64)
    jj = np.array([ii[e] + 1 + (e % 2) for e in range(n_edges)], np.int64)
    y, x = np.meshgrid(np.arange(ht, dtype=np.float64), np.arange(wd, dtype=np.float64), indexing="ij")
    lo = rng.uniform(-8, 8, size=(n_edges, 2, (ht + 7) // 8, (wd + 7) // 8))
    flow = np.repeat(np.repeat(lo, 8, axis=2), 8, axis=3)[:, :, :ht, :wd] + rng.uniform(-0.5, 0.5, size=(n_edges, 2, ht, wd))
    coords = np.stack([x, y])[None] + flow
    if oob_frac > 0:
        m = rng.uniform(size=(n_edges, 1, ht, wd)) < oob_frac
        coords = np.where(m, coords + rng.choice([-1, 1], size=coords.shape) * rng.uniform(40, 90, size=coords.shape), coords)
    return dict(fmaps=fm.astype(dtype), ii=ii, jj=jj, coords=coords.astype(np.float32), ht=ht, wd=wd)
