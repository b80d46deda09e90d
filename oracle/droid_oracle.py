"""TEST INFRASTRUCTURE ONLY -- CPU restatement (numpy) of the reference `droid_backends` operators.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this
module, and only as the checker / the timed CPU baseline; the product (droid-slam_b200/) never does.

Every function restates the algorithm of the reference CUDA source and cites it
(paths relative to /root/reference).  Arithmetic is IEEE float32 (numpy, no FMA contraction) wherever the
reference computes in float, float16 with per-operation rounding where the reference runs on c10::Half,
and float64 where the reference does (host assembly + Cholesky, depth_filter comparison).

Parity pinning: the reference ships NO golden vectors or tests for these ops.  This
oracle is pinned against outputs of the reference itself: the unmodified reference extension compiled by
oracle/build_ref.sh (oracle/_ref/droid_backends_ref*.so) is run on the GPU box by
tests/golden/make_golden_gpu.py and its outputs are committed under tests/golden/; tests/test_oracle.py checks this
oracle against them on CPU.  The functions that restate the reference's PYTHON (reproject, proximity_edges,
filter_repeated_edges, cvx_upsample) are pinned against outputs of the unmodified Python itself, run in the build
container: tests/golden/make_golden_pyref.py (on the reference's own lietorch, oracle/build_lietorch.sh),
make_golden_proximity.py, make_golden_upsample.py.  scatter_mean restates a third-party package that is absent
(torch_scatter): parity unpinned.
"""
import numpy as np

F32 = np.float32
MIN_DEPTH = F32(0.25)  # src/droid_kernels.cu:26


# ------------------------------------------------------------------------------------------------
# SE3 helpers on arrays [...,3] / [...,4]              src/droid_kernels.cu:58-175
# ------------------------------------------------------------------------------------------------
def act_so3(q, X):
    """src/droid_kernels.cu:58-68.  q [...,4] (x,y,z,w), X [...,3]."""
    q0, q1, q2, q3 = q[..., 0], q[..., 1], q[..., 2], q[..., 3]
    X0, X1, X2 = X[..., 0], X[..., 1], X[..., 2]
    uv0 = F32(2.0) * (q1 * X2 - q2 * X1)
    uv1 = F32(2.0) * (q2 * X0 - q0 * X2)
    uv2 = F32(2.0) * (q0 * X1 - q1 * X0)
    Y0 = X0 + q3 * uv0 + (q1 * uv2 - q2 * uv1)
    Y1 = X1 + q3 * uv1 + (q2 * uv0 - q0 * uv2)
    Y2 = X2 + q3 * uv2 + (q0 * uv1 - q1 * uv0)
    return np.stack(np.broadcast_arrays(Y0, Y1, Y2), -1).astype(F32)


def act_se3(t, q, X):
    """src/droid_kernels.cu:70-77.  X [...,4] homogeneous (x,y,z,d)."""
    Y = act_so3(q, X[..., :3])
    d = X[..., 3:4]
    return np.concatenate([Y + d * t, np.broadcast_to(d, Y.shape[:-1] + (1,))], -1).astype(F32)


def adj_se3(t, q, X):
    """src/droid_kernels.cu:79-94.  X [...,6] covector -> Ad(T)^T X."""
    qinv = q * np.array([-1, -1, -1, 1], F32)
    Y0 = act_so3(qinv, X[..., 0:3])
    Y1 = act_so3(qinv, X[..., 3:6])
    u = np.stack(np.broadcast_arrays(t[..., 2] * X[..., 1] - t[..., 1] * X[..., 2],
                                     t[..., 0] * X[..., 2] - t[..., 2] * X[..., 0],
                                     t[..., 1] * X[..., 0] - t[..., 0] * X[..., 1]), -1).astype(F32)
    v = act_so3(qinv, u)
    return np.concatenate([Y0, Y1 + v], -1).astype(F32)


def rel_se3(ti, qi, tj, qj):
    """src/droid_kernels.cu:96-107.  Tij = Tj * Ti^-1."""
    q = np.empty(np.broadcast(qi, qj).shape, F32)
    q[..., 0] = -qj[..., 3] * qi[..., 0] + qj[..., 0] * qi[..., 3] - qj[..., 1] * qi[..., 2] + qj[..., 2] * qi[..., 1]
    q[..., 1] = -qj[..., 3] * qi[..., 1] + qj[..., 1] * qi[..., 3] - qj[..., 2] * qi[..., 0] + qj[..., 0] * qi[..., 2]
    q[..., 2] = -qj[..., 3] * qi[..., 2] + qj[..., 2] * qi[..., 3] - qj[..., 0] * qi[..., 1] + qj[..., 1] * qi[..., 0]
    q[..., 3] = qj[..., 3] * qi[..., 3] + qj[..., 0] * qi[..., 0] + qj[..., 1] * qi[..., 1] + qj[..., 2] * qi[..., 2]
    t = tj - act_so3(q, ti)
    return t.astype(F32), q


def exp_so3(phi):
    """src/droid_kernels.cu:110-132 (single vector)."""
    phi = phi.astype(F32)
    theta_sq = F32(phi[0] * phi[0] + phi[1] * phi[1] + phi[2] * phi[2])
    theta_p4 = F32(theta_sq * theta_sq)
    theta = F32(np.sqrt(theta_sq))
    if float(theta_sq) < 1e-8:
        imag = F32(0.5 - (1.0 / 48.0) * float(theta_sq) + (1.0 / 3840.0) * float(theta_p4))
        real = F32(1.0 - (1.0 / 8.0) * float(theta_sq) + (1.0 / 384.0) * float(theta_p4))
    else:
        imag = F32(np.sin(F32(0.5 * float(theta)), dtype=F32) / theta)
        real = F32(np.cos(F32(0.5 * float(theta)), dtype=F32))
    return np.array([imag * phi[0], imag * phi[1], imag * phi[2], real], F32)


def _cross(a, b):
    return np.array([a[1] * b[2] - a[2] * b[1], a[2] * b[0] - a[0] * b[2], a[0] * b[1] - a[1] * b[0]], F32)


def exp_se3(xi):
    """src/droid_kernels.cu:147-175."""
    xi = xi.astype(F32)
    q = exp_so3(xi[3:])
    tau, phi = xi[:3].copy(), xi[3:].copy()
    theta_sq = F32(phi[0] * phi[0] + phi[1] * phi[1] + phi[2] * phi[2])
    theta = F32(np.sqrt(theta_sq))
    t = tau.copy()
    if float(theta) > 1e-4:
        a = F32((F32(1) - np.cos(theta, dtype=F32)) / theta_sq)
        tau = _cross(phi, tau)
        t = t + a * tau
        b = F32((theta - np.sin(theta, dtype=F32)) / F32(theta * theta_sq))
        tau = _cross(phi, tau)
        t = t + b * tau
    return t.astype(F32), q


def retr_se3(xi, t, q):
    """src/droid_kernels.cu:877-895.  T1 = Exp(xi) * T."""
    dt, dq = exp_se3(xi)
    q1 = np.array([
        dq[3] * q[0] + dq[0] * q[3] + dq[1] * q[2] - dq[2] * q[1],
        dq[3] * q[1] + dq[1] * q[3] + dq[2] * q[0] - dq[0] * q[2],
        dq[3] * q[2] + dq[2] * q[3] + dq[0] * q[1] - dq[1] * q[0],
        dq[3] * q[3] - dq[0] * q[0] - dq[1] * q[1] - dq[2] * q[2]], F32)
    t1 = act_so3(dq, t) + dt
    return t1.astype(F32), q1


def _grid(ht, wd, intr):
    fx, fy, cx, cy = [F32(v) for v in intr]
    v, u = np.meshgrid(np.arange(ht, dtype=F32), np.arange(wd, dtype=F32), indexing="ij")
    return u, v, ((u - cx) / fx).astype(F32), ((v - cy) / fy).astype(F32)


def _edge_pose(poses, ix, jx, stereo_rule):
    if stereo_rule and ix == jx:  # src/droid_kernels.cu:219-229
        return np.array([-0.1, 0, 0], F32), np.array([0, 0, 0, 1], F32)
    return rel_se3(poses[ix, :3], poses[ix, 3:], poses[jx, :3], poses[jx, 3:])


def _transform(poses, disps, intr, ix, jx, stereo_rule=False):
    ht, wd = disps.shape[1:]
    u, v, xn, yn = _grid(ht, wd, intr)
    tij, qij = _edge_pose(poses, ix, jx, stereo_rule)
    Xi = np.stack([xn, yn, np.ones_like(xn), disps[ix].astype(F32)], -1)
    return u, v, Xi, act_se3(tij, qij, Xi), tij, qij


# ------------------------------------------------------------------------------------------------
# geometry ops
# ------------------------------------------------------------------------------------------------
def projmap(poses, disps, intr, ii, jj):
    """src/droid_kernels.cu:427-516.  -> coords [n,ht,wd,3], valid [n,ht,wd,1]."""
    n = len(ii)
    ht, wd = disps.shape[1:]
    fx, fy, cx, cy = [F32(v) for v in intr]
    coords = np.zeros((n, ht, wd, 3), F32)
    valid = np.zeros((n, ht, wd, 1), F32)
    for e in range(n):
        u, v, Xi, Xj, _, _ = _transform(poses, disps, intr, int(ii[e]), int(jj[e]))
        ok = Xj[..., 2] > 0.01
        with np.errstate(all="ignore"):
            pu = fx * (Xj[..., 0] / Xj[..., 2]) + cx
            pv = fy * (Xj[..., 1] / Xj[..., 2]) + cy
        coords[e, ..., 0] = np.where(ok, pu, u)
        coords[e, ..., 1] = np.where(ok, pv, v)
        valid[e, ..., 0] = (Xj[..., 2] > MIN_DEPTH).astype(F32)
    return coords, valid


def _quat_mul(a, b):
    """Eigen quaternion product a * b, coefficients (x, y, z, w)."""
    ax, ay, az, aw = [a[..., k] for k in range(4)]
    bx, by, bz, bw = [b[..., k] for k in range(4)]
    return np.stack([aw * bx + ax * bw + ay * bz - az * by,
                     aw * by + ay * bw + az * bx - ax * bz,
                     aw * bz + az * bw + ax * by - ay * bx,
                     aw * bw - ax * bx - ay * by - az * bz], -1).astype(F32)


def reproject(poses, disps, intr, ii, jj, target=None):
    """DepthVideo.reproject (droid_slam/depth_video.py:139-147) = projective_transform with jacobian=False
    (droid_slam/geom/projective_ops.py:96-125), group arithmetic as in lietorch (thirdparty/lietorch/lietorch/include/
    se3.h:36-52, so3.h:43-60).  intr [nbuf,4] per frame.  -> coords [n,ht,wd,2], valid [n,ht,wd,1]
    (+ motn [n,4,ht,wd], the motion features of droid_slam/factor_graph.py:203-204, when target is given)."""
    n = len(ii)
    ht, wd = disps.shape[1:]
    coords = np.zeros((n, ht, wd, 2), F32)
    valid = np.zeros((n, ht, wd, 1), F32)
    motn = np.zeros((n, 4, ht, wd), F32) if target is not None else None
    for e in range(n):
        ix, jx = int(ii[e]), int(jj[e])
        u, v, xn, yn = _grid(ht, wd, intr[ix])                                   # iproj, projective_ops.py:18-37
        X0 = np.stack([xn, yn, np.ones_like(xn), disps[ix].astype(F32)], -1)
        if ix == jx:                                                             # projective_ops.py:105
            t, q = np.array([-0.1, 0, 0], F32), np.array([0, 0, 0, 1], F32)
        else:                                                                    # Gj * Gi.inv()
            ti, qi, tj, qj = poses[ix, :3].astype(F32), poses[ix, 3:].astype(F32), poses[jx, :3].astype(F32), poses[jx, 3:].astype(F32)
            qinv = qi * np.array([-1, -1, -1, 1], F32)
            tinv = -act_so3(qinv, ti)
            q = _quat_mul(qj, qinv)
            t = (tj + act_so3(qj, tinv)).astype(F32)
        X1 = act_se3(t, q, X0)                                                   # actp: act4
        fx, fy, cx, cy = [F32(a) for a in intr[jx]]
        Z = np.where(X1[..., 2] < F32(0.1), F32(1.0), X1[..., 2]).astype(F32)   # proj, projective_ops.py:39-53
        d = (F32(1.0) / Z).astype(F32)
        coords[e, ..., 0] = fx * (X1[..., 0] * d) + cx
        coords[e, ..., 1] = fy * (X1[..., 1] * d) + cy
        valid[e, ..., 0] = (X1[..., 2] > F32(0.2)).astype(F32)                   # X0[...,2] = 1 > 0.2 always
        if motn is not None:
            m = np.concatenate([coords[e] - np.stack([u, v], -1), target[e].astype(F32) - coords[e]], -1)
            motn[e] = np.clip(m, -64.0, 64.0).transpose(2, 0, 1)
    return (coords, valid) if motn is None else (coords, valid, motn)


def iproj(poses, disps, intr):
    """src/droid_kernels.cu:779-850.  -> points [nm,ht,wd,3]."""
    nm, ht, wd = disps.shape
    _, _, xn, yn = _grid(ht, wd, intr)
    out = np.zeros((nm, ht, wd, 3), F32)
    for n in range(nm):
        Xi = np.stack([xn, yn, np.ones_like(xn), disps[n].astype(F32)], -1)
        Xj = act_se3(poses[n, :3], poses[n, 3:], Xi)
        with np.errstate(all="ignore"):
            out[n] = Xj[..., :3] / Xj[..., 3:4]
    return out


def _tree256(x):
    """blockReduce + warpReduce pairing order, src/droid_kernels.cu:36-55.  x [256, ...] float32."""
    x = x.astype(F32).copy()
    for s in (128, 64, 32, 16, 8, 4, 2, 1):
        x[:s] = x[:s] + x[s:2 * s]
    return x[0]


def _lanes(a):
    """[HW, ...] -> [iterations, 256, ...]: element k is visited by thread k % 256 in iteration k // 256
    (GPU_1D_KERNEL_LOOP, src/droid_kernels.cu:32-33); the tail is zero padded."""
    hw = a.shape[0]
    nit = (hw + 255) // 256
    pad = np.zeros((nit * 256 - hw,) + a.shape[1:], a.dtype)
    return np.concatenate([a, pad], 0).reshape((nit, 256) + a.shape[1:])


def _cta_gram(A, B):
    """sum_p A[p,:,None] * B[p,None,:] the way a 256-thread CTA does it (strided FFMA partials + tree)."""
    la, lb = _lanes(A), _lanes(B)
    acc = np.zeros((256, A.shape[1], B.shape[1]), F32)
    for it in range(la.shape[0]):
        acc = _fma(la[it][:, :, None], lb[it][:, None, :], acc)
    return _tree256(acc)


def frame_distance(poses, disps, intr, ii, jj, beta):
    """src/droid_kernels.cu:518-657 (one direction; thread-strided partial sums + 256->1 tree)."""
    n = len(ii)
    ht, wd = disps.shape[1:]
    fx, fy, cx, cy = [F32(v) for v in intr]
    beta = F32(beta)
    omb = F32(F32(1) - beta)
    out = np.zeros(n, F32)
    hw = ht * wd
    nit = (hw + 255) // 256
    for e in range(n):
        u, v, Xi, Xj, tij, qij = _transform(poses, disps, intr, int(ii[e]), int(jj[e]))
        with np.errstate(all="ignore"):
            du = fx * (Xj[..., 0] / Xj[..., 2]) + cx - u
            dv = fy * (Xj[..., 1] / Xj[..., 2]) + cy - v
            d1 = np.sqrt(du * du + dv * dv, dtype=F32)
            ok1 = Xj[..., 2] > MIN_DEPTH
            X2 = Xi[..., :3] + Xi[..., 3:4] * tij
            du = fx * (X2[..., 0] / X2[..., 2]) + cx - u
            dv = fy * (X2[..., 1] / X2[..., 2]) + cy - v
            d2 = np.sqrt(du * du + dv * dv, dtype=F32)
            ok2 = X2[..., 2] > MIN_DEPTH
        pad = nit * 256 - hw

        def lanes(a):
            return np.concatenate([a.reshape(-1), np.zeros(pad, a.dtype)]).reshape(nit, 256)
        d1l, d2l, o1, o2 = lanes(d1), lanes(d2), lanes(ok1), lanes(ok2)
        live = lanes(np.ones(hw, bool))
        accum = np.zeros(256, F32); valid = np.zeros(256, F32); total = np.zeros(256, F32)
        for it in range(nit):
            lv = live[it]
            total = np.where(lv, total + beta, total)
            accum = np.where(lv & o1[it], accum + beta * d1l[it], accum)
            valid = np.where(lv & o1[it], valid + beta, valid)
            total = np.where(lv, total + omb, total)
            accum = np.where(lv & o2[it], accum + omb * d2l[it], accum)
            valid = np.where(lv & o2[it], valid + omb, valid)
        a, t, w = _tree256(accum), _tree256(total), _tree256(valid)
        out[e] = F32(1000.0) if (float(w) / (float(t) + 1e-8) < 0.75) else F32(a / w)
    return out


def depth_filter(poses, disps, intr, ix, thresh):
    """src/droid_kernels.cu:661-775.  -> counter [num,ht,wd] (integers 0..6 as float32)."""
    num = len(ix)
    nbuf, ht, wd = disps.shape
    fx, fy, cx, cy = [F32(v) for v in intr]
    counter = np.zeros((num, ht, wd), F32)
    for b in range(num):
        i = int(ix[b])
        for nb in range(6):
            j = i - nb - 1 if nb < 3 else i + nb   # :695
            if j < 0 or j >= nbuf:
                continue
            _, _, Xi, Xj, _, _ = _transform(poses, disps, intr, i, j)
            with np.errstate(all="ignore"):
                uj = fx * (Xj[..., 0] / Xj[..., 2]) + cx
                vj = fy * (Xj[..., 1] / Xj[..., 2]) + cy
                dj = (Xj[..., 3] / Xj[..., 2]).astype(F32)
                u0 = np.floor(uj); v0 = np.floor(vj)
            inb = (u0 >= 0) & (v0 >= 0) & (u0 < wd - 1) & (v0 < ht - 1) & np.isfinite(u0) & np.isfinite(v0)
            u0i = np.where(inb, u0, 0).astype(np.int64); v0i = np.where(inb, v0, 0).astype(np.int64)
            dm = disps[j]
            hit = np.zeros((ht, wd), bool)
            with np.errstate(all="ignore"):
                idj = 1.0 / dj.astype(np.float64)
                for dy, dx in ((0, 0), (0, 1), (1, 0), (1, 1)):
                    dn = dm[np.minimum(v0i + dy, ht - 1), np.minimum(u0i + dx, wd - 1)].astype(np.float64)
                    hit |= np.abs(idj - 1.0 / dn) < float(thresh[b])
            counter[b] += (hit & inb).astype(F32)
    return counter


# ------------------------------------------------------------------------------------------------
# correlation lookup                                    src/correlation_kernels.cu:19-124
# ------------------------------------------------------------------------------------------------
def _mad(acc, s, w, dtype):
    """acc += s * scalar_t(w) with the reference's per-op rounding (c10::Half rounds product and sum)."""
    if dtype == np.float16:
        p = (s.astype(F32) * w.astype(np.float16).astype(F32)).astype(np.float16)
        return (acc.astype(F32) + p.astype(F32)).astype(np.float16)
    if dtype == np.float64:
        return acc + s * w.astype(np.float64)
    return (acc + s * w.astype(F32)).astype(F32)


def corr_index_forward(volume, coords, r):
    """src/correlation_kernels.cu:19-70.  volume [B,H1,W1,H2,W2], coords [B,2,H1,W1] f32 ->
    corr [B,2r+1,2r+1,H1,W1] (dtype of volume).  Tap loop order i (x) outer, j (y) inner, 4 scatter
    targets per tap, accumulation in the volume dtype."""
    B, H1, W1, H2, W2 = volume.shape
    dt = volume.dtype.type
    rd = 2 * r + 1
    x0 = coords[:, 0].astype(F32); y0 = coords[:, 1].astype(F32)
    fx0 = np.floor(x0); fy0 = np.floor(y0)
    dx = (x0 - fx0).astype(F32); dy = (y0 - fy0).astype(F32)
    one = F32(1.0)
    w11 = (dx * dy).astype(F32); w10 = (dx * (one - dy)).astype(F32)
    w01 = ((one - dx) * dy).astype(F32); w00 = ((one - dx) * (one - dy)).astype(F32)
    corr = np.zeros((B, rd, rd, H1, W1), dt)
    bi, yi, xi = np.meshgrid(np.arange(B), np.arange(H1), np.arange(W1), indexing="ij")
    for i in range(rd + 1):
        for j in range(rd + 1):
            x1 = fx0.astype(np.int64) - r + i
            y1 = fy0.astype(np.int64) - r + j
            inb = (y1 >= 0) & (y1 < H2) & (x1 >= 0) & (x1 < W2)
            s = volume[bi, yi, xi, np.clip(y1, 0, H2 - 1), np.clip(x1, 0, W2 - 1)]
            for (ci, cj, w) in ((i - 1, j - 1, w11), (i - 1, j, w10), (i, j - 1, w01), (i, j, w00)):
                if 0 <= ci < rd and 0 <= cj < rd:
                    new = _mad(corr[:, ci, cj], s, w, dt)
                    corr[:, ci, cj] = np.where(inb, new, corr[:, ci, cj])
    return corr


def corr_index_backward(volume_shape, dtype, coords, corr_grad, r):
    """src/correlation_kernels.cu:73-124.  -> volume_grad [B,H1,W1,H2,W2]."""
    B, H1, W1, H2, W2 = volume_shape
    dt = np.dtype(dtype).type
    rd = 2 * r + 1
    x0 = coords[:, 0].astype(F32); y0 = coords[:, 1].astype(F32)
    fx0 = np.floor(x0); fy0 = np.floor(y0)
    dx = (x0 - fx0).astype(F32); dy = (y0 - fy0).astype(F32)
    one = F32(1.0)
    w11 = (dx * dy).astype(F32); w10 = (dx * (one - dy)).astype(F32)
    w01 = ((one - dx) * dy).astype(F32); w00 = ((one - dx) * (one - dy)).astype(F32)
    vg = np.zeros(volume_shape, dt)
    bi, yi, xi = np.meshgrid(np.arange(B), np.arange(H1), np.arange(W1), indexing="ij")
    for i in range(rd + 1):
        for j in range(rd + 1):
            x1 = fx0.astype(np.int64) - r + i
            y1 = fy0.astype(np.int64) - r + j
            inb = (y1 >= 0) & (y1 < H2) & (x1 >= 0) & (x1 < W2)
            g = np.zeros((B, H1, W1), dt)
            for (ci, cj, w) in ((i - 1, j - 1, w11), (i - 1, j, w10), (i, j - 1, w01), (i, j, w00)):
                if 0 <= ci < rd and 0 <= cj < rd:
                    g = _mad(g, corr_grad[:, ci, cj], w, dt)
            sel = inb
            vg[bi[sel], yi[sel], xi[sel], y1[sel], x1[sel]] = g[sel]
    return vg


def corr_build(fmap1, fmap2, num_levels=4):
    """droid_slam/modules/corr.py:24-38,63-71 for f16 inputs [B,C,H,W]: each fmap is divided by 4 in
    f16, the product accumulates in f32 (cuBLAS HGEMM) and is rounded to f16; every pyramid level is the
    2x2 mean (f32 accumulate, f16 result) of the previous f16 level (F.avg_pool2d on half)."""
    B, C, H, W = fmap1.shape
    a = (fmap1.astype(np.float16) / np.float16(4.0)).astype(np.float16).reshape(B, C, H * W).astype(F32)
    b = (fmap2.astype(np.float16) / np.float16(4.0)).astype(np.float16).reshape(B, C, H * W).astype(F32)
    corr = np.einsum("bci,bcj->bij", a, b, optimize=True).astype(np.float16)
    lv = [corr.reshape(B, H, W, H, W)]
    cur = lv[0]
    for _ in range(1, num_levels):
        h, w = cur.shape[3] // 2, cur.shape[4] // 2
        c = cur[:, :, :, :2 * h, :2 * w].astype(F32).reshape(B, H, W, h, 2, w, 2)
        cur = ((c[..., 0, :, 0] + c[..., 0, :, 1] + c[..., 1, :, 0] + c[..., 1, :, 1]) * F32(0.25)).astype(np.float16)
        lv.append(cur)
    return lv


# ------------------------------------------------------------------------------------------------
# alt correlation                                       src/altcorr_kernel.cu:27-286
# ------------------------------------------------------------------------------------------------
def altcorr_forward(fmap1, fmap2, coords, r):
    """src/altcorr_kernel.cu:27-149.  fmap1 [B,H1,W1,C], fmap2 [B,H2,W2,C], coords [B,N,H1,W1,2] ->
    corr [B,N,(2r+1)^2,H1,W1]; channel = iy + rd*ix.  Dot products over C in float32 (32-channel chunks
    like the reference, each chunk weighted and added separately)."""
    B, H1, W1, C = fmap1.shape
    _, H2, W2, _ = fmap2.shape
    N = coords.shape[1]
    rd = 2 * r + 1
    out = np.zeros((B, N, rd * rd, H1, W1), F32)
    one = F32(1.0)
    for b in range(B):
        f2 = fmap2[b].astype(F32)
        f1 = fmap1[b].astype(F32)
        for n in range(N):
            x = coords[b, n, ..., 0].astype(F32); y = coords[b, n, ..., 1].astype(F32)
            fx0 = np.floor(x); fy0 = np.floor(y)
            dx = (x - fx0).astype(F32); dy = (y - fy0).astype(F32)
            for c0 in range(0, C, 32):
                f1c = f1[..., c0:c0 + 32]
                for iy in range(rd + 1):
                    for ix in range(rd + 1):
                        h2 = fy0.astype(np.int64) - r + iy
                        w2 = fx0.astype(np.int64) - r + ix
                        inb = (h2 >= 0) & (h2 < H2) & (w2 >= 0) & (w2 < W2)
                        g = f2[np.clip(h2, 0, H2 - 1), np.clip(w2, 0, W2 - 1), c0:c0 + 32]
                        s = np.where(inb, np.einsum("hwc,hwc->hw", f1c, g).astype(F32), F32(0))
                        for (oy, ox, w) in ((iy - 1, ix - 1, dy * dx), (iy - 1, ix, dy * (one - dx)),
                                            (iy, ix - 1, (one - dy) * dx), (iy, ix, (one - dy) * (one - dx))):
                            if 0 <= oy < rd and 0 <= ox < rd:
                                out[b, n, oy + rd * ox] += (s * w.astype(F32)).astype(F32)
    return out


def altcorr_backward(fmap1, fmap2, coords, corr_grad, r):
    """src/altcorr_kernel.cu:152-286.  -> fmap1_grad, fmap2_grad, coords_grad (zeros)."""
    B, H1, W1, C = fmap1.shape
    _, H2, W2, _ = fmap2.shape
    N = coords.shape[1]
    rd = 2 * r + 1
    g1 = np.zeros_like(fmap1, dtype=np.float64)
    g2 = np.zeros_like(fmap2, dtype=np.float64)
    one = F32(1.0)
    for b in range(B):
        for n in range(N):
            x = coords[b, n, ..., 0].astype(F32); y = coords[b, n, ..., 1].astype(F32)
            fx0 = np.floor(x); fy0 = np.floor(y)
            dx = (x - fx0).astype(F32); dy = (y - fy0).astype(F32)
            for iy in range(rd + 1):
                for ix in range(rd + 1):
                    h2 = fy0.astype(np.int64) - r + iy
                    w2 = fx0.astype(np.int64) - r + ix
                    inb = (h2 >= 0) & (h2 < H2) & (w2 >= 0) & (w2 < W2)
                    g = np.zeros((H1, W1), F32)
                    for (oy, ox, w) in ((iy - 1, ix - 1, dy * dx), (iy - 1, ix, dy * (one - dx)),
                                        (iy, ix - 1, (one - dy) * dx), (iy, ix, (one - dy) * (one - dx))):
                        if 0 <= oy < rd and 0 <= ox < rd:
                            g = g + corr_grad[b, n, oy + rd * ox] * w
                    g = np.where(inb, g, 0).astype(np.float64)
                    hc, wc = np.clip(h2, 0, H2 - 1), np.clip(w2, 0, W2 - 1)
                    g1[b] += g[..., None] * fmap2[b][hc, wc].astype(np.float64)
                    np.add.at(g2[b], (hc, wc), g[..., None] * fmap1[b].astype(np.float64))
    return g1.astype(F32), g2.astype(F32), np.zeros(coords.shape, F32)


# ------------------------------------------------------------------------------------------------
# dense bundle adjustment                               src/droid_kernels.cu:176-424, :854-1434
# ------------------------------------------------------------------------------------------------
def _fma(a, b, c):
    """float32 fused multiply-add (one rounding), emulated through float64."""
    return (np.asarray(a, np.float64) * np.asarray(b, np.float64) + np.asarray(c, np.float64)).astype(F32)


def _edge_linearise(poses, disps, intr, target, weight, ix, jx):
    """projective_transform_kernel for ONE edge (src/droid_kernels.cu:176-424).
    target, weight [2,ht,wd].  Returns Hs [4,6,6], vs [2,6], Eii [6,HW], Eij [6,HW], Cii [HW], bz [HW]."""
    fx, fy, cx, cy = [F32(v) for v in intr]
    u, v, Xi, Xj, tij, qij = _transform(poses, disps, intr, ix, jx, stereo_rule=True)
    x, y, z, h = Xj[..., 0], Xj[..., 1], Xj[..., 2], Xj[..., 3]
    behind = z < MIN_DEPTH
    with np.errstate(all="ignore"):
        d = np.where(behind, F32(0), (1.0 / z.astype(np.float64)).astype(F32)).astype(F32)   # :302
    d2 = d * d
    wu = np.where(behind, F32(0), (.001 * weight[0].astype(np.float64)).astype(F32)).astype(F32)  # :305
    wv = np.where(behind, F32(0), (.001 * weight[1].astype(np.float64)).astype(F32)).astype(F32)
    # nvcc contracts (fx*d)*x + cx into one FFMA (single rounding); the residual is a difference of
    # nearly equal numbers, so this rounding matters at the 1e-5 level of r and is emulated here
    ru = (target[0] - _fma((fx * d).astype(F32), x, cx)).astype(F32)                     # :307
    rv = (target[1] - _fma((fy * d).astype(F32), y, cy)).astype(F32)                     # :308
    zero = np.zeros_like(x)
    one = F32(1)
    Jj_u = np.stack([fx * (h * d), fx * zero, fx * (-x * h * d2), fx * (-x * y * d2),
                     fx * (one + x * x * d2), fx * (-y * d)], -1).astype(F32)           # :312-317
    Jz_u = (fx * (tij[0] * d - tij[2] * (x * d2))).astype(F32)                           # :319
    Jj_v = np.stack([fy * zero, fy * (h * d), fy * (-y * h * d2), fy * (-one - y * y * d2),
                     fy * (x * y * d2), fy * (x * d)], -1).astype(F32)                   # :345-350
    Jz_v = (fy * (tij[1] * d - tij[2] * (y * d2))).astype(F32)                           # :352
    Ji_u = (-adj_se3(tij, qij, Jj_u)).astype(F32)                                        # :325-326
    Ji_v = (-adj_se3(tij, qij, Jj_v)).astype(F32)
    Cii = (wu * Jz_u * Jz_u).astype(F32)
    bz = (wu * ru * Jz_u).astype(F32)
    Cii = (Cii + wv * Jz_v * Jz_v).astype(F32)
    bz = (bz + wv * rv * Jz_v).astype(F32)
    if ix == jx:                                                                         # :323,:356
        wu = np.zeros_like(wu); wv = np.zeros_like(wv)
    Ju = np.concatenate([Ji_u, Jj_u], -1).reshape(-1, 12)
    Jv = np.concatenate([Ji_v, Jj_v], -1).reshape(-1, 12)
    wuf, wvf, ruf, rvf = wu.reshape(-1), wv.reshape(-1), ru.reshape(-1), rv.reshape(-1)
    # Per-thread partial sums over the thread-strided pixels (k = tid, tid+256, ...), u row then v row
    # per pixel (:328-339, :361-371), then the 256 -> 1 shared-memory tree (:382-423).  The product
    # (w*J[n])*J[m] is added with one rounding (nvcc contracts it to FFMA).
    WJu = (wuf[:, None] * Ju).astype(F32); WJv = (wvf[:, None] * Jv).astype(F32)
    wru = (wuf * ruf).astype(F32); wrv = (wvf * rvf).astype(F32)
    lu, lv_, lwju, lwjv = _lanes(Ju), _lanes(Jv), _lanes(WJu), _lanes(WJv)
    lwru, lwrv = _lanes(wru), _lanes(wrv)
    accH = np.zeros((256, 12, 12), F32); accV = np.zeros((256, 12), F32)
    for it in range(lu.shape[0]):
        accH = _fma(lwju[it][:, :, None], lu[it][:, None, :], accH)
        accV = _fma(lwru[it][:, None], lu[it], accV)
        accH = _fma(lwjv[it][:, :, None], lv_[it][:, None, :], accH)
        accV = _fma(lwrv[it][:, None], lv_[it], accV)
    H = _tree256(accH); vv = _tree256(accV)
    H = np.tril(H) + np.tril(H, -1).T          # the kernel accumulates n >= m and mirrors (:400-423)
    Hs = np.stack([H[:6, :6], H[:6, 6:], H[6:, :6], H[6:, 6:]]).astype(F32)              # :400-423
    vs = np.stack([vv[:6], vv[6:]]).astype(F32)
    au = (wuf * Jz_u.reshape(-1)).astype(F32)[:, None]
    av = (wvf * Jz_v.reshape(-1)).astype(F32)[:, None]
    Eii = ((au * Ju[:, :6]).astype(F32) + (av * Jv[:, :6]).astype(F32)).astype(F32).T    # :340-341,373-374
    Eij = ((au * Ju[:, 6:]).astype(F32) + (av * Jv[:, 6:]).astype(F32)).astype(F32).T
    return Hs, vs, np.ascontiguousarray(Eii), np.ascontiguousarray(Eij), Cii.reshape(-1), bz.reshape(-1)


def _accum(data, ix, jx):
    """accum_cuda (src/droid_kernels.cu:948-998): out[j] = sum_{i: ix[i]==jx[j]} data[i], float32."""
    out = np.zeros((len(jx),) + data.shape[1:], F32)
    for j, key in enumerate(jx):
        rows = np.nonzero(ix == key)[0]
        acc = np.zeros(data.shape[1:], F32)
        for i in rows:
            acc = (acc + data[i]).astype(F32)
        out[j] = acc
    return out


def _dense_solve(A, b, lm, ep):
    """SparseBlock::solve (src/droid_kernels.cu:1192-1213): diag += ep + lm*diag, LLT (lower triangle),
    zeros on failure.  Eigen's sparse LLT is an exact Cholesky of the same matrix; dense float64 here."""
    L = A.copy()
    dg = np.arange(L.shape[0])
    L[dg, dg] += float(F32(ep)) + float(F32(lm)) * L[dg, dg]
    L = np.tril(L) + np.tril(L, -1).T
    try:
        c = np.linalg.cholesky(L)
    except np.linalg.LinAlgError:
        return None
    y = np.linalg.solve(c, b) if c.shape[0] < 64 else __import__("scipy.linalg").linalg.solve_triangular(c, b, lower=True)
    x = np.linalg.solve(c.T, y) if c.shape[0] < 64 else __import__("scipy.linalg").linalg.solve_triangular(c.T, y, lower=False)
    return x


def ba(poses, disps, intr, disps_sens, targets, weights, eta, ii, jj, t0, t1, iterations, lm, ep,
       motion_only, return_debug=False):
    """ba_cuda (src/droid_kernels.cu:1314-1434).  poses [nbuf,7], disps [nbuf,ht,wd] are updated IN PLACE
    (numpy arrays, float32).  Returns (dx [P,6], dz [M,HW] or None)."""
    ii = np.asarray(ii, np.int64); jj = np.asarray(jj, np.int64)
    num = len(ii)
    ht, wd = disps.shape[1:]
    HW = ht * wd
    P = t1 - t0
    ts = np.arange(t0, t1, dtype=np.int64)
    ii_exp = np.concatenate([ts, ii]); jj_exp = np.concatenate([ts, jj])
    kx, kk_exp = np.unique(ii_exp, return_inverse=True)                                   # :1340-1344
    M = len(kx)
    dx = dz = None
    dbg = {}
    for _ in range(iterations):
        Hs = np.zeros((4, num, 6, 6), F32); vs = np.zeros((2, num, 6), F32)
        Eii = np.zeros((num, 6, HW), F32); Eij = np.zeros((num, 6, HW), F32)
        Cii = np.zeros((num, HW), F32); wi = np.zeros((num, HW), F32)
        for e in range(num):
            h, v, ei, ej, c, b = _edge_linearise(poses, disps, intr, targets[e], weights[e], int(ii[e]), int(jj[e]))
            Hs[:, e] = h; vs[:, e] = v; Eii[e] = ei; Eij[e] = ej; Cii[e] = c; wi[e] = b
        # pose x pose block in double (SparseBlock::update_lhs/rhs :1131-1173): negative indices dropped
        n6 = 6 * P
        A = np.zeros((n6, n6)); bvec = np.zeros(n6)
        for blk, (ri, ci) in enumerate(((ii, ii), (ii, jj), (jj, ii), (jj, jj))):         # :1378-1380
            for e in range(num):
                i, j = int(ri[e]) - t0, int(ci[e]) - t0
                if i >= 0 and j >= 0:
                    A[6 * i:6 * i + 6, 6 * j:6 * j + 6] += Hs[blk, e].astype(np.float64)
        for blk, ri in enumerate((ii, jj)):                                                # :1382-1383
            for e in range(num):
                i = int(ri[e]) - t0
                if i >= 0:
                    bvec[6 * i:6 * i + 6] += vs[blk, e].astype(np.float64)
        if motion_only:
            x = _dense_solve(A, bvec, lm, ep)
            dx = np.zeros((P, 6), F32) if x is None else x.reshape(P, 6).astype(F32)
        else:
            alpha = F32(0.05)
            m = (disps_sens[kx] > 0).astype(F32).reshape(-1, HW)                           # :1397
            C = ((_accum(Cii, ii, kx) + m * alpha).astype(F32) + ((F32(1) - m) * eta.reshape(-1, HW)).astype(F32)).astype(F32)
            w = (_accum(wi, ii, kx) - (m * alpha * (disps[kx] - disps_sens[kx]).reshape(-1, HW)).astype(F32)).astype(F32)
            Q = (F32(1.0) / C).astype(F32)                                                 # :1400
            Ei = _accum(Eii.reshape(num, 6 * HW), ii, ts).reshape(P, 6, HW)                # :1402
            E = np.concatenate([Ei, Eij], 0)                                               # :1403
            # schur_block (:1222-1311): rows n with t0 <= jj_exp[n] <= t1 grouped by pose; triples with
            # equal depth index kk; S blocks float32 (EEt6x6), summed in double
            S = np.zeros((n6, n6)); vS = np.zeros(n6)
            rows_pose = jj_exp - t0
            for k in range(M):
                rows = np.nonzero((kk_exp == k) & (rows_pose >= 0) & (rows_pose < P))[0]
                if len(rows) == 0:
                    continue
                Ek = E[rows].reshape(len(rows) * 6, HW)
                Sk = _cta_gram((Ek * Q[k][None, :]).astype(F32).T, Ek.T)                   # :1025-1055
                vk = _cta_gram((Q[k] * w[k]).astype(F32)[:, None], Ek.T)[0]                # :1074-1091
                for a, ra in enumerate(rows):
                    pa = int(rows_pose[ra])
                    vS[6 * pa:6 * pa + 6] += vk[6 * a:6 * a + 6].astype(np.float64)
                    for b_, rb in enumerate(rows):
                        pb = int(rows_pose[rb])
                        S[6 * pa:6 * pa + 6, 6 * pb:6 * pb + 6] += Sk[6 * a:6 * a + 6, 6 * b_:6 * b_ + 6].astype(np.float64)
            if return_debug:
                dbg = dict(A=A.copy(), b=bvec.copy(), S=S.copy(), vS=vS.copy(), Q=Q, w=w, kx=kx)
            x = _dense_solve(A - S, bvec - vS, lm, ep)                                     # :1406
            dx = np.zeros((P, 6), F32) if x is None else x.reshape(P, 6).astype(F32)
            # back-substitution (:1408-1417): rows with pose index <= 0 or >= P are skipped (:1105)
            dw = np.zeros((len(jj_exp), HW), F32)
            for n_ in range(len(jj_exp)):
                pz = int(rows_pose[n_])
                if pz <= 0 or pz >= P:
                    continue
                acc = np.zeros(HW, F32)
                for c_ in range(6):
                    acc = (acc + E[n_, c_] * dx[pz, c_]).astype(F32)
                dw[n_] = acc
            dz = (Q * (w - _accum(dw, ii_exp, kx))).astype(F32)
        for k in range(t0, t1):                                                            # :898-931
            t1_, q1_ = retr_se3(dx[k - t0], poses[k, :3].copy(), poses[k, 3:].copy())
            poses[k, :3] = t1_; poses[k, 3:] = q1_
        if not motion_only:
            disps[kx] = (disps[kx] + dz.reshape(-1, ht, wd)).astype(F32)                   # :933-946
    if return_debug:
        return dx, dz, dbg
    return dx, dz


# ------------------------------------------------------------------------------------------------
# edge selection
# ------------------------------------------------------------------------------------------------
def proximity_edges(d, t, t0, t1, rad, nms, thresh, max_factors, stereo, ii_exist, jj_exist):
    """FactorGraph.add_proximity_factors (droid_slam/factor_graph.py:305-369) from the point where the distances are
    known: d = video.distance(ii, jj, beta) over meshgrid(arange(t0, t), arange(t1, t)), row-major.  ii_exist /
    jj_exist = cat(active, bad, inactive) edges (:322-323).  Returns the edge list handed to add_factors, in order:
    the mandatory temporal (and stereo) edges, then the accepted pairs (both directions) in ascending distance.
    Ties in d are taken in index order (the reference's argsort leaves them unspecified)."""
    d = np.array(d, np.float32, copy=True)
    W = t - t1
    gi, gj = np.meshgrid(np.arange(t0, t), np.arange(t1, t), indexing="ij")
    gi, gj = gi.reshape(-1), gj.reshape(-1)
    d[gi - rad < gj] = np.inf                                                     # :318
    d[d > 100] = np.inf                                                           # :319

    def suppress(i, j):                                                           # :324-333, :359-367
        r = max(min(abs(i - j) - 2, nms), 0)
        for di in range(-nms, nms + 1):
            for dj in range(-nms, nms + 1):
                if abs(di) + abs(dj) <= r:
                    i1, j1 = i + di, j + dj
                    if t0 <= i1 < t and t1 <= j1 < t:
                        d[(i1 - t0) * W + (j1 - t1)] = np.inf

    for i, j in zip(np.asarray(ii_exist, np.int64), np.asarray(jj_exist, np.int64)):
        suppress(int(i), int(j))
    es = []
    for i in range(t0, t):                                                        # :336-345
        if stereo:
            es.append((i, i))
            d[(i - t0) * W + (i - t1)] = np.inf
        for j in range(max(i - rad - 1, 0), i):
            es.append((i, j))
            es.append((j, i))
            d[(i - t0) * W + (j - t1)] = np.inf
    for k in np.argsort(d, kind="stable"):                                        # :347-367
        if d[k] > thresh:
            continue
        if len(es) > max_factors:
            break
        i, j = int(gi[k]), int(gj[k])
        es.append((i, j))
        es.append((j, i))
        suppress(i, j)
    e = np.asarray(es, np.int64).reshape(-1, 2)
    return e[:, 0].copy(), e[:, 1].copy()


# ------------------------------------------------------------------------------------------------
# convex upsampling
# ------------------------------------------------------------------------------------------------
def filter_repeated_edges(ii, jj, ii_exist, jj_exist):
    """FactorGraph.__filter_repeated_edges (droid_slam/factor_graph.py:43-54): candidates that are not among the
    existing (active + inactive) edges, original order; duplicates inside the candidate list are all kept."""
    eset = set(zip(np.asarray(ii_exist).tolist(), np.asarray(jj_exist).tolist()))
    keep = np.array([(int(i), int(j)) not in eset for i, j in zip(ii, jj)], bool).reshape(-1)
    return np.asarray(ii, np.int64)[keep], np.asarray(jj, np.int64)[keep]


def cvx_upsample(data, mask):
    """droid_slam/droid_net.py:21-35.  data [batch,ht,wd,dim], mask [batch,9*8*8,ht,wd] logits -> [batch,8ht,8wd,dim]:
    out[b, 8y+i, 8x+j, :] = sum_k softmax_k(mask[b, k, i, j, y, x]) * data[b, y + k//3 - 1, x + k%3 - 1, :] with zeros
    outside the image (F.unfold(..., padding=1))."""
    data = np.asarray(data, F32)
    batch, ht, wd, dim = data.shape
    m = np.asarray(mask, F32).reshape(batch, 9, 8, 8, ht, wd)
    m = m - m.max(axis=1, keepdims=True)
    e = np.exp(m).astype(F32)
    w = (e / e.sum(axis=1, keepdims=True, dtype=F32)).astype(F32)                 # softmax over the 9 neighbours
    pad = np.zeros((batch, ht + 2, wd + 2, dim), F32)
    pad[:, 1:-1, 1:-1] = data
    out = np.zeros((batch, ht, 8, wd, 8, dim), F32)
    for k in range(9):
        nb = pad[:, k // 3:k // 3 + ht, k % 3:k % 3 + wd]                         # [batch,ht,wd,dim]
        out += w[:, k].transpose(0, 3, 1, 4, 2)[..., None] * nb[:, :, None, :, None, :]
    return out.reshape(batch, 8 * ht, 8 * wd, dim)


# ------------------------------------------------------------------------------------------------
# segmented mean over edges
# ------------------------------------------------------------------------------------------------
def scatter_mean(src, index, dim_size=None):
    """torch_scatter.scatter_mean(src, index, dim=1) as GraphAgg.forward uses it (droid_slam/droid_net.py:59-75).
    torch_scatter is not vendored in the reference (conda `pytorch-scatter`, unpinned, environment.yaml:10), so this
    restates its published algorithm: out = scatter_sum(src) / clamp(scatter_sum(ones), min=1).  src [1, n, ...],
    index [n] -> [1, M, ...]; sums in float64 here (the checker), result in src's dtype."""
    src = np.asarray(src)
    index = np.asarray(index, np.int64)
    M = int(dim_size) if dim_size is not None else (int(index.max()) + 1 if len(index) else 0)
    acc = np.zeros((M,) + src.shape[2:], np.float64)
    cnt = np.zeros(M, np.float64)
    np.add.at(acc, index, src[0].astype(np.float64))
    np.add.at(cnt, index, 1.0)
    out = acc / np.maximum(cnt, 1.0).reshape((M,) + (1,) * (src.ndim - 2))
    return out.astype(src.dtype)[None]
