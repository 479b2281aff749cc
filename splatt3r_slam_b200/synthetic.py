"""Seeded synthetic inputs of MASt3R shape for the backend hot path.

Everything is generated on the CPU with a torch.Generator so that the same seed gives the same
tensors here and on the GPU box; callers move the result to the device.

Pointmap noise: sigma = 0.001 scene units (~0.25 px of ray angle at f = 500).  MASt3R pointmaps are
smooth at pixel scale; with sigma = 0.005 (2.5x the angular pixel pitch) the LM iteration of
iter_proj diverges on 72% of the pixels, which is neither realistic nor what the tracker sees
(~91% of the pixels converge at sigma <= 0.001; the rest project outside the image).

Scene: a smooth height field z = f(x, y) in the world frame, observed by pinhole cameras
(fx = fy = 500*w/512, principal point at the image centre) placed at seeded Sim3 poses
[t(3), q xyzw(4), s(1)] (the lietorch layout the reference stores in `T_WC.data`).
"""
import math

import torch


def _quat_from_axis_angle(v):
    th = v.norm(dim=-1, keepdim=True).clamp_min(1e-12)
    return torch.cat([torch.sin(0.5 * th) * v / th, torch.cos(0.5 * th)], dim=-1)


def quat_rotate(q, X):
    qv, qw = q[..., :3], q[..., 3:]
    uv = 2.0 * torch.cross(qv.expand_as(X), X, dim=-1)
    return X + qw * uv + torch.cross(qv.expand_as(X), uv, dim=-1)


def sim3_act(T, X):
    """T (...,8) acting on X (...,3): s*R*X + t"""
    return T[..., 7:8] * quat_rotate(T[..., 3:7], X) + T[..., 0:3]


def sim3_inv_act(T, X):
    q = T[..., 3:7] * torch.tensor([-1.0, -1.0, -1.0, 1.0], dtype=T.dtype, device=T.device)
    return quat_rotate(q, X - T[..., 0:3]) / T[..., 7:8]


def random_pose(gen, rot_deg=5.0, trans=0.1, scale=0.05):
    axis = torch.randn(3, generator=gen, dtype=torch.float64)
    axis = axis / axis.norm()
    ang = math.radians(rot_deg) * (2 * torch.rand(1, generator=gen, dtype=torch.float64) - 1)
    q = _quat_from_axis_angle(axis * ang)
    t = trans * (2 * torch.rand(3, generator=gen, dtype=torch.float64) - 1)
    s = 1.0 + scale * (2 * torch.rand(1, generator=gen, dtype=torch.float64) - 1)
    return torch.cat([t, q, s])


class Scene:
    def __init__(self, seed, h=384, w=512):
        self.gen = torch.Generator().manual_seed(seed)
        self.h, self.w = h, w
        f = 500.0 * w / 512.0
        self.K = torch.tensor([[f, 0, w / 2.0], [0, f, h / 2.0], [0, 0, 1]], dtype=torch.float64)
        g = self.gen
        self.freq = 0.5 + 2.5 * torch.rand(4, 2, generator=g, dtype=torch.float64)
        self.phase = 2 * math.pi * torch.rand(4, generator=g, dtype=torch.float64)

    def surf(self, x, y):
        z = torch.full_like(x, 2.0)
        for k in range(4):
            z = z + 0.125 * torch.sin(self.freq[k, 0] * x + self.freq[k, 1] * y + self.phase[k])
        return z

    def raycast(self, T):
        """World points (h,w,3) hit by the pixel rays of the camera at Sim3 pose T (fp64)."""
        h, w, K = self.h, self.w, self.K
        dev = K.device
        v, u = torch.meshgrid(torch.arange(h, dtype=torch.float64, device=dev), torch.arange(w, dtype=torch.float64, device=dev),
                              indexing="ij")
        d_cam = torch.stack([(u - K[0, 2]) / K[0, 0], (v - K[1, 2]) / K[1, 1], torch.ones_like(u)], -1)
        d = quat_rotate(T[3:7], d_cam)  # world direction (scale folded into the ray parameter)
        o = T[0:3]
        t = torch.full((h, w), 2.0, dtype=torch.float64, device=dev)
        for _ in range(8):
            P = o + t[..., None] * d
            t = t + (self.surf(P[..., 0], P[..., 1]) - P[..., 2]) / d[..., 2]
        return o + t[..., None] * d

    def project(self, T, Pw):
        """World points -> (u, v, z) in the camera at pose T."""
        Pc = sim3_inv_act(T, Pw)
        z = Pc[..., 2]
        u = self.K[0, 0] * Pc[..., 0] / z + self.K[0, 2]
        v = self.K[1, 1] * Pc[..., 1] / z + self.K[1, 2]
        return u, v, z


def _descriptors(Pw, fdim, W, noise, gen):
    feat = torch.cat([torch.sin(3.0 * Pw), torch.cos(3.0 * Pw), Pw], dim=-1)  # (...,9)
    D = feat @ W + noise * torch.randn(*Pw.shape[:-1], fdim, generator=gen, dtype=torch.float64)
    return D / D.norm(dim=-1, keepdim=True)


def make_pair(seed=0, h=384, w=512, fdim=24, b=1, noise=0.001):
    """One matcher call's inputs: X11, X21 (b,h,w,3), D11, D21 (b,h,w,fdim), all fp32.
    View 1 is the identity camera; X21 are the points seen by a second camera at a seeded
    Sim3 offset, expressed in camera 1 (what the asymmetric MASt3R head emits)."""
    X11, X21, D11, D21 = [], [], [], []
    for i in range(b):
        sc = Scene(seed * 1000 + i, h, w)
        g = sc.gen
        W = torch.randn(9, fdim, generator=g, dtype=torch.float64)
        T1 = torch.tensor([0, 0, 0, 0, 0, 0, 1, 1], dtype=torch.float64)
        T2 = random_pose(g)
        P1, P2 = sc.raycast(T1), sc.raycast(T2)
        X11.append(P1 + noise * torch.randn(h, w, 3, generator=g, dtype=torch.float64))
        X21.append(P2 + noise * torch.randn(h, w, 3, generator=g, dtype=torch.float64))
        D11.append(_descriptors(P1, fdim, W, 0.1, g))
        D21.append(_descriptors(P2, fdim, W, 0.1, g))
    f = lambda L: torch.stack(L).float().contiguous()
    return f(X11), f(X21), f(D11), f(D21)


def make_graph(seed=0, N=2, E_undirected=1, h=384, w=512, noise=0.001, pose_noise=0.01,
               extra_edges="random"):
    """A keyframe graph in the layout `global_opt.py:104-139` hands to the backend.

    Returns dict with Twc (N,8) [perturbed], Twc_gt, Xs (N,n,3), Cs (N,n,1), K (3,3),
    ii, jj (E,) int64 keyframe ids, idx_ii2jj (E,n) int64, valid_match (E,n,1) bool,
    Q (E,n,1) fp32, with E = 2*E_undirected directed edges (both directions).
    Edges: ring (k,k+1) first, then seeded random chords."""
    sc = Scene(seed, h, w)
    g = sc.gen
    n = h * w
    poses = [torch.tensor([0, 0, 0, 0, 0, 0, 1, 1], dtype=torch.float64)]
    for k in range(1, N):
        poses.append(random_pose(g, rot_deg=6.0, trans=0.15, scale=0.05))
    Tgt = torch.stack(poses)
    Pw = [sc.raycast(Tgt[k]) for k in range(N)]
    Xs = torch.stack([sim3_inv_act(Tgt[k], Pw[k]).reshape(n, 3) for k in range(N)])
    Xs = Xs + noise * torch.randn(N, n, 3, generator=g, dtype=torch.float64)
    Cs = 1.0 + torch.exp(0.5 * torch.randn(N, n, 1, generator=g, dtype=torch.float64))
    und = [(k, (k + 1) % N) for k in range(min(N - 1 if N == 2 else N, E_undirected))]
    while len(und) < E_undirected:
        a, b_ = torch.randint(0, N, (2,), generator=g).tolist()
        if a != b_:
            und.append((a, b_))
    ii = torch.tensor([a for a, _ in und] + [b_ for _, b_ in und], dtype=torch.int64)
    jj = torch.tensor([b_ for _, b_ in und] + [a for a, _ in und], dtype=torch.int64)
    E = ii.numel()
    idx = torch.zeros(E, n, dtype=torch.int64)
    valid = torch.zeros(E, n, 1, dtype=torch.bool)
    for e in range(E):
        i, j = int(ii[e]), int(jj[e])
        u, v, z = sc.project(Tgt[i], Pw[j].reshape(n, 3))
        ui, vi = torch.round(u).long(), torch.round(v).long()
        ok = (ui >= 0) & (ui < w) & (vi >= 0) & (vi < h) & (z > 0)
        idx[e] = torch.where(ok, ui.clamp(0, w - 1) + w * vi.clamp(0, h - 1), torch.zeros_like(ui))
        valid[e, :, 0] = ok
    Q = 1.0 + torch.exp(0.5 * torch.randn(E, n, 1, generator=g, dtype=torch.float64))
    # perturb the free poses (left-multiplicative small Sim3 step; pose 0 is the gauge)
    T = Tgt.clone()
    for k in range(1, N):
        d = random_pose(g, rot_deg=math.degrees(pose_noise), trans=pose_noise, scale=pose_noise)
        t = d[7] * quat_rotate(d[3:7], T[k, 0:3]) + d[0:3]
        qd, q = d[3:7], T[k, 3:7]
        qn = torch.stack([
            qd[3] * q[0] + qd[0] * q[3] + qd[1] * q[2] - qd[2] * q[1],
            qd[3] * q[1] - qd[0] * q[2] + qd[1] * q[3] + qd[2] * q[0],
            qd[3] * q[2] + qd[0] * q[1] - qd[1] * q[0] + qd[2] * q[3],
            qd[3] * q[3] - qd[0] * q[0] - qd[1] * q[1] - qd[2] * q[2]])
        T[k] = torch.cat([t, qn, (d[7] * T[k, 7]).reshape(1)])
    return dict(Twc=T.float().contiguous(), Twc_gt=Tgt.float(), Xs=Xs.float().contiguous(),
                Cs=Cs.float().contiguous(), K=sc.K.float().contiguous(), ii=ii, jj=jj,
                idx_ii2jj=idx.contiguous(), valid_match=valid.contiguous(),
                Q=Q.float().contiguous(), h=h, w=w)


def make_graph_device(seed, N, E_undirected, h, w, device, edge_range=None, noise=0.001, pose_noise=0.01):
    """Same construction as make_graph, for graphs too large to build on the host (BASELINE configs[3]:
    N = 1024, E = 8192 at 512x384 is 21 GB of per-edge arrays): topology and poses come from the CPU
    generator (identical on every rank), the pointmaps / confidences / per-edge arrays are generated ON
    `device` with a device generator, and only the directed edges [e0, e1) of `edge_range` are
    materialised (a rank's shard).  Values differ from make_graph's for the same seed (different random
    streams); statistics and structure are the same."""
    sc = Scene(seed, h, w)
    g = sc.gen
    n = h * w
    poses = [torch.tensor([0, 0, 0, 0, 0, 0, 1, 1], dtype=torch.float64)]
    for k in range(1, N):
        poses.append(random_pose(g, rot_deg=6.0, trans=0.15, scale=0.05))
    Tgt = torch.stack(poses)
    und = [(k, (k + 1) % N) for k in range(min(N - 1 if N == 2 else N, E_undirected))]
    while len(und) < E_undirected:
        a, b_ = torch.randint(0, N, (2,), generator=g).tolist()
        if a != b_:
            und.append((a, b_))
    ii = torch.tensor([a for a, _ in und] + [b_ for _, b_ in und], dtype=torch.int64)
    jj = torch.tensor([b_ for _, b_ in und] + [a for a, _ in und], dtype=torch.int64)
    E = ii.numel()
    T = Tgt.clone()
    for k in range(1, N):
        d = random_pose(g, rot_deg=math.degrees(pose_noise), trans=pose_noise, scale=pose_noise)
        t = d[7] * quat_rotate(d[3:7], T[k, 0:3]) + d[0:3]
        qd, q = d[3:7], T[k, 3:7]
        qn = torch.stack([
            qd[3] * q[0] + qd[0] * q[3] + qd[1] * q[2] - qd[2] * q[1],
            qd[3] * q[1] - qd[0] * q[2] + qd[1] * q[3] + qd[2] * q[0],
            qd[3] * q[2] + qd[0] * q[1] - qd[1] * q[0] + qd[2] * q[3],
            qd[3] * q[3] - qd[0] * q[0] - qd[1] * q[1] - qd[2] * q[2]])
        T[k] = torch.cat([t, qn, (d[7] * T[k, 7]).reshape(1)])
    # heavy arrays on the device
    sc.K, sc.freq, sc.phase = sc.K.to(device), sc.freq.to(device), sc.phase.to(device)
    dg = torch.Generator(device=device).manual_seed(seed * 7919 + 13)
    Tg = Tgt.to(device)
    Xs = torch.empty((N, n, 3), dtype=torch.float32, device=device)
    Cs = torch.empty((N, n, 1), dtype=torch.float32, device=device)
    for k in range(N):
        Pw = sc.raycast(Tg[k])
        X = sim3_inv_act(Tg[k], Pw).reshape(n, 3)
        Xs[k] = (X + noise * torch.randn(n, 3, generator=dg, dtype=torch.float64, device=device)).float()
        Cs[k] = (1.0 + torch.exp(0.5 * torch.randn(n, 1, generator=dg, dtype=torch.float32, device=device)))
    e0, e1 = (0, E) if edge_range is None else edge_range
    Ec = e1 - e0
    idx = torch.zeros((Ec, n), dtype=torch.int64, device=device)
    valid = torch.zeros((Ec, n, 1), dtype=torch.bool, device=device)
    Q = torch.empty((Ec, n, 1), dtype=torch.float32, device=device)
    for e in range(e0, e1):
        i, j = int(ii[e]), int(jj[e])
        Pw_j = sim3_act(Tg[j], Xs[j].double())  # world points of keyframe j (noise included: sub-pixel)
        u, v, z = sc.project(Tg[i], Pw_j)
        ui, vi = torch.round(u).long(), torch.round(v).long()
        ok = (ui >= 0) & (ui < w) & (vi >= 0) & (vi < h) & (z > 0)
        idx[e - e0] = torch.where(ok, ui.clamp(0, w - 1) + w * vi.clamp(0, h - 1), torch.zeros_like(ui))
        valid[e - e0, :, 0] = ok
        eg = torch.Generator(device=device).manual_seed(seed * 104729 + e)  # per edge: independent of the sharding
        Q[e - e0] = 1.0 + torch.exp(0.5 * torch.randn(n, 1, generator=eg, dtype=torch.float32, device=device))
    return dict(Twc=T.float().contiguous().to(device), Twc_gt=Tgt.float(), Xs=Xs, Cs=Cs, K=sc.K.float().contiguous(),
                ii=ii.to(device), jj=jj.to(device), idx_ii2jj=idx, valid_match=valid, Q=Q, h=h, w=w)


def make_track_problem(seed=0, n=4096, noise=0.001, invalid=0.2, device="cpu"):
    """Frame-to-keyframe tracking problem in the shape `FrameTracker.opt_pose_ray_dist_sim3`
    (reference tracker.py:174) receives it: matched frame points Xf (n,3), keyframe points Xk (n,3), poses
    T_WCf (start), T_WCk as Sim3 data (8,), match confidence Qk (n,1), valid (n,1) bool — plus the true relative
    pose T_CkCf_gt the optimiser should recover.  A tenth of the valid points are gross outliers (Huber)."""
    gen = torch.Generator().manual_seed(seed)
    Xk = torch.randn(n, 3, generator=gen) * 0.5 + torch.tensor([0.0, 0.0, 2.5])
    T_gt = random_pose(gen, rot_deg=4.0, trans=0.08, scale=0.04).float()
    Tk = random_pose(gen, rot_deg=20.0, trans=0.5, scale=0.1).float()
    Xf = sim3_inv_act(T_gt, Xk) + noise * torch.randn(n, 3, generator=gen)
    out = torch.rand(n, generator=gen) < 0.1
    Xf[out] += 0.2 * torch.randn(int(out.sum()), 3, generator=gen)
    Qk = torch.rand(n, 1, generator=gen) * 3.0 + 0.5
    valid = (torch.rand(n, 1, generator=gen) > invalid) & (Qk > 1.5)
    d = dict(Xf=Xf.contiguous(), Xk=Xk.contiguous(), T_WCf=Tk.clone(), T_WCk=Tk, Qk=Qk, valid=valid, T_CkCf_gt=T_gt)
    return {k: v.to(device) for k, v in d.items()}


def make_track_problem_calib(seed=0, h=48, w=64, noise=0.001, invalid=0.2, device="cpu"):
    """Calibrated tracking problem (`FrameTracker.opt_pose_calib_sim3`, reference tracker.py:216): the
    keyframe points lie on the rays of the h x w pixel grid of a pinhole camera K (what
    constrain_points_to_ray produces, tracker.py:143-144); a few have non-positive depth."""
    gen = torch.Generator().manual_seed(seed)
    n = h * w
    K = torch.tensor([[0.9 * w, 0.0, 0.5 * w], [0.0, 0.9 * w, 0.5 * h], [0.0, 0.0, 1.0]])
    v, u = torch.meshgrid(torch.arange(h, dtype=torch.float32), torch.arange(w, dtype=torch.float32), indexing="ij")
    z = 2.0 + torch.rand(h, w, generator=gen)
    Xk = torch.stack(((u - K[0, 2]) / K[0, 0] * z, (v - K[1, 2]) / K[1, 1] * z, z), dim=-1).view(n, 3)
    T_gt = random_pose(gen, rot_deg=3.0, trans=0.06, scale=0.03).float()
    Tk = random_pose(gen, rot_deg=20.0, trans=0.5, scale=0.1).float()
    Xf = sim3_inv_act(T_gt, Xk) + noise * torch.randn(n, 3, generator=gen)
    out = torch.rand(n, generator=gen) < 0.1
    Xf[out] += 0.2 * torch.randn(int(out.sum()), 3, generator=gen)
    Xk[::97, 2] = -1.0  # invalid keyframe depth (valid_meas_k false, tracker.py:150)
    Qk = torch.rand(n, 1, generator=gen) * 3.0 + 0.5
    valid = (torch.rand(n, 1, generator=gen) > invalid) & (Qk > 1.5)
    d = dict(Xf=Xf.contiguous(), Xk=Xk.contiguous(), T_WCf=Tk.clone(), T_WCk=Tk, Qk=Qk, valid=valid, K=K, T_CkCf_gt=T_gt)
    d = {k: t.to(device) for k, t in d.items()}
    d["img_size"] = (h, w)
    return d


def make_splat_scene(seed=0, P=2000, H=96, W=128, fov_deg=60.0, sh_coeffs=16, identity_camera=False):
    """A synthetic 3D-Gaussian scene + camera in the conventions of the reference's rasteriser calls
    (splatt3r_core/src/pixelsplat_src/cuda_splatting.py:77-124): `viewmatrix` = transposed world->camera matrix,
    `projmatrix` = viewmatrix @ transposed perspective matrix, `campos` = camera centre; splats in front of, behind and
    beside the camera, a few very large and very small ones.  `identity_camera`: camera frame == world frame, so the view
    depth of a splat is its z coordinate EXACTLY in every implementation, and a tenth of the splats are given the depth of
    another one (order ties, resolved by splat index).  (With a rotated camera, near-equal depths would be ordered by the
    last bit of a 3-term fp32 sum — a property of the rounding, not of the algorithm.)
    Returns CPU float32 tensors: means (P,3), cov (P,6), scales (P,3), rotations (P,4) [r,x,y,z; cov == cov3D(scales,
    rotations)], opacities (P,1), colors (P,3), shs (P,sh_coeffs,3), view (4,4), proj (4,4), campos (3), bg (3), and the
    floats tanfovx, tanfovy."""
    g = torch.Generator().manual_seed(seed)
    r = lambda *s: torch.rand(*s, generator=g, dtype=torch.float64)
    rn = lambda *s: torch.randn(*s, generator=g, dtype=torch.float64)
    tanx = math.tan(math.radians(fov_deg) / 2)
    tany = tanx * H / W
    # camera-to-world: small random rotation + translation
    ax = rn(3); ax = ax / ax.norm(); ang = 0.3
    Kx = torch.tensor([[0, -ax[2], ax[1]], [ax[2], 0, -ax[0]], [-ax[1], ax[0], 0]], dtype=torch.float64)
    R = torch.eye(3, dtype=torch.float64) + math.sin(ang) * Kx + (1 - math.cos(ang)) * (Kx @ Kx)
    c2w = torch.eye(4, dtype=torch.float64); c2w[:3, :3] = R; c2w[:3, 3] = rn(3) * 0.5
    if identity_camera:
        R = torch.eye(3, dtype=torch.float64)
        c2w = torch.eye(4, dtype=torch.float64)
    # points in the camera frame
    z = 0.6 + 5.0 * r(P)
    z[: P // 20] = -1.0 - r(P // 20)                    # behind the camera
    z[P // 20: P // 10] = 0.05 + 0.2 * r(P // 10 - P // 20)   # around the 0.2 near cull
    x = (2.6 * r(P) - 1.3) * tanx * z.abs()             # up to 1.3x outside the frustum
    y = (2.6 * r(P) - 1.3) * tany * z.abs()
    pc = torch.stack([x, y, z], 1)
    if identity_camera:
        pc = pc.float().double()
        pc[P // 2: P // 2 + P // 10, 2] = pc[P // 2 - P // 10: P // 2, 2]  # exactly equal view depths
    means = (pc @ R.T + c2w[:3, 3]).float()
    scales = (0.01 + 0.06 * r(P, 3)) * z.abs()[:, None] * 0.5
    scales[-8:] *= 12.0                                 # a few huge splats (hundreds of tiles)
    scales[-16:-8] *= 0.01                              # and tiny ones (low-pass filter dominates)
    q = rn(P, 4); q = q / q.norm(dim=1, keepdim=True)
    rr, xx, yy, zz = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    Rq = torch.stack([torch.stack([1 - 2 * (yy * yy + zz * zz), 2 * (xx * yy - rr * zz), 2 * (xx * zz + rr * yy)], 1),
                      torch.stack([2 * (xx * yy + rr * zz), 1 - 2 * (xx * xx + zz * zz), 2 * (yy * zz - rr * xx)], 1),
                      torch.stack([2 * (xx * zz - rr * yy), 2 * (yy * zz + rr * xx), 1 - 2 * (xx * xx + yy * yy)], 1)], 1)
    Sigma = Rq @ torch.diag_embed(scales ** 2) @ Rq.transpose(1, 2)
    cov = torch.stack([Sigma[:, 0, 0], Sigma[:, 0, 1], Sigma[:, 0, 2], Sigma[:, 1, 1], Sigma[:, 1, 2], Sigma[:, 2, 2]], 1).float()
    near, far = 0.1, 100.0
    Pj = torch.zeros(4, 4, dtype=torch.float64)
    Pj[0, 0], Pj[1, 1], Pj[3, 2] = 1 / tanx, 1 / tany, 1.0
    Pj[2, 2], Pj[2, 3] = far / (far - near), -(far * near) / (far - near)
    view = torch.linalg.inv(c2w).T.contiguous()
    proj = (view @ Pj.T).contiguous()
    return dict(means=means, cov=cov, scales=scales.float(), rotations=q.float(), opacities=(0.05 + 0.95 * r(P, 1)).float(),
                colors=r(P, 3).float(), shs=(rn(P, sh_coeffs, 3) * 0.3).float(), view=view.float(), proj=proj.float(),
                campos=c2w[:3, 3].float(), bg=torch.tensor([0.1, 0.2, 0.3]), tanfovx=tanx, tanfovy=tany, H=H, W=W)
