"""TEST INFRASTRUCTURE ONLY — CPU restatement of the reference's frame tracker optimisation,
`FrameTracker.opt_pose_ray_dist_sim3` + `FrameTracker.solve` (splatt3r_slam/tracker.py:156-214) with
`check_convergence` / `huber` (nonlinear_optimizer.py:5-33), `point_to_ray_dist` / `act_Sim3` /
`skew_sym` (geometry.py:5-52), in the same PyTorch fp32 operations and the same order, statement for
statement.  lietorch (the reference's Sim3 class) is not installed here, so the four Sim3 methods the
function uses are restated on the (t, q_xyzw, s) 8-vector: inv, mul, act (lietorch sim3.h:43-58,
156-165) and retr = Exp(tau) * T through the C oracle's retraction (oracle_gn_solve.c, pinned against
the reference CUDA build's pose_retr_kernel).

PARITY PINNED against the reference's own tracker.py: oracle/build_lietorch.py builds the reference's
lietorch and byte-compiles its drivers; tests/golden/make_golden_drivers.py ran
FrameTracker.opt_pose_ray_dist_sim3 / opt_pose_calib_sim3 with them and committed the poses
(tests/golden/drivers_64x48.npz); tests/test_drivers_golden.py holds this restatement to 1e-5 of those
(measured <= 1.7e-6) and, where oracle/_ref is present, to the live reference tracker on fresh seeds.
Only tests/, __graft_entry__.smoke() and bench.py's CPU baselines may import this module."""
import math

import torch

import oracle


def quat_mul(a, b):
    ax, ay, az, aw = a.unbind(-1)
    bx, by, bz, bw = b.unbind(-1)
    return torch.stack((aw * bx + ax * bw + ay * bz - az * by, aw * by - ax * bz + ay * bw + az * bx,
                        aw * bz + ax * by - ay * bx + az * bw, aw * bw - ax * bx - ay * by - az * bz), dim=-1)


def quat_rot(q, X):
    qv, qw = q[..., :3], q[..., 3:4]
    uv = 2.0 * torch.linalg.cross(qv.expand_as(X), X)
    return X + qw * uv + torch.linalg.cross(qv.expand_as(X), uv)


def sim3_act(T, X):
    return T[7] * quat_rot(T[3:7], X) + T[:3]


def sim3_inv(T):
    qi = torch.cat((-T[3:6], T[6:7]))
    si = 1.0 / T[7]
    return torch.cat((-si * quat_rot(qi, T[:3]), qi, si.view(1)))


def sim3_mul(A, B):
    return torch.cat((A[7] * quat_rot(A[3:7], B[:3]) + A[:3], quat_mul(A[3:7], B[3:7]), (A[7] * B[7]).view(1)))


def sim3_retr(T, tau):
    out = T.clone().view(1, 8).contiguous()
    oracle.pose_retr(out, tau.view(1, 7).contiguous(), num_fix=0)
    return out.view(8)


def skew_sym(x):
    b = x.shape[:-1]
    x, y, z = x.unbind(dim=-1)
    o = torch.zeros_like(x)
    return torch.stack([o, -z, y, z, o, -x, -y, x, o], dim=-1).view(*b, 3, 3)


def point_to_ray_dist(X, jacobian=False):  # geometry.py:17-34
    b = X.shape[:-1]
    d = torch.linalg.norm(X, dim=-1, keepdim=True)
    d_inv = 1.0 / d
    r = d_inv * X
    rd = torch.cat((r, d), dim=-1)
    if not jacobian:
        return rd
    d_inv_2 = d_inv ** 2
    I = torch.eye(3, dtype=X.dtype).repeat(*b, 1, 1)
    dr_dX = d_inv.unsqueeze(-1) * (I - d_inv_2.unsqueeze(-1) * (X.unsqueeze(-1) @ X.unsqueeze(-2)))
    return rd, torch.cat((dr_dX, r.unsqueeze(-2)), dim=-2)


def act_Sim3(T, pC):  # geometry.py:45-52
    pW = sim3_act(T, pC)
    dpC_dt = torch.eye(3).repeat(*pW.shape[:-1], 1, 1)
    return pW, torch.cat([dpC_dt, -skew_sym(pW), pW.reshape(*pW.shape[:-1], -1, 1)], dim=-1)


def huber(r, k=1.345):  # nonlinear_optimizer.py:28-33
    r_abs = torch.abs(r)
    return torch.where(r_abs < k, torch.ones(1, dtype=r.dtype), k / r_abs)


def check_convergence(rel_error_threshold, delta_norm_threshold, old_cost, new_cost, delta):
    cost_diff = old_cost - new_cost
    rel_dec = math.fabs(cost_diff / old_cost)
    return rel_dec < rel_error_threshold or torch.linalg.norm(delta) < delta_norm_threshold


def solve(sqrt_info, r, J, huber_k):  # tracker.py:156-172
    whitened_r = sqrt_info * r
    robust_sqrt_info = sqrt_info * torch.sqrt(huber(whitened_r, k=huber_k))
    mdim = J.shape[-1]
    A = (robust_sqrt_info[..., None] * J).view(-1, mdim)
    b = (robust_sqrt_info * r).view(-1, 1)
    H = A.T @ A
    g = -A.T @ b
    cost = 0.5 * (b.T @ b).item()
    L = torch.linalg.cholesky(H, upper=False)  # raises on a non-SPD system (tracker.py:88-90 catches it)
    return torch.cholesky_solve(g, L, upper=False).view(1, -1), cost


def opt_pose_ray_dist_sim3(Xf, Xk, T_WCf, T_WCk, Qk, valid, sigma_ray=0.003, sigma_dist=10.0, huber_k=1.345,
                           max_iters=50, rel_error=1e-3, delta_norm=1e-3):
    """Xf (n,3) matched frame points, Xk (n,3), T_* (8,), Qk (n,1), valid (n,1) bool -> T_WCf, T_CkCf, iters."""
    sqrt_info_ray = 1 / sigma_ray * valid * torch.sqrt(Qk)
    sqrt_info_dist = 1 / sigma_dist * valid * torch.sqrt(Qk)
    sqrt_info = torch.cat((sqrt_info_ray.repeat(1, 3), sqrt_info_dist), dim=1)
    T_CkCf = sim3_mul(sim3_inv(T_WCk), T_WCf)
    rd_k = point_to_ray_dist(Xk, jacobian=False)
    old_cost = float("inf")
    steps = 0
    for step in range(max_iters):
        Xf_Ck, dXf_Ck_dT_CkCf = act_Sim3(T_CkCf, Xf)
        rd_f_Ck, drd_f_Ck_dXf_Ck = point_to_ray_dist(Xf_Ck, jacobian=True)
        r = rd_k - rd_f_Ck
        J = -drd_f_Ck_dXf_Ck @ dXf_Ck_dT_CkCf
        tau, new_cost = solve(sqrt_info, r, J, huber_k)
        T_CkCf = sim3_retr(T_CkCf, tau)
        steps = step + 1
        if check_convergence(rel_error, delta_norm, old_cost, new_cost, tau):
            break
        old_cost = new_cost
    return sim3_mul(T_WCk, T_CkCf), T_CkCf, steps


# ---- calibrated variant (tracker.py:216-266, geometry.py:63-105) ------------------------------------
def project_calib(P, K, img_size, border=0, z_eps=0.0):
    b = P.shape[:-1]
    K_rep = K.repeat(*b, 1, 1)
    p = (K_rep @ P[..., None]).squeeze(-1)
    p = p / p[..., 2:3]
    p = p[..., :2]
    u, v = p.split([1, 1], dim=-1)
    x, y, z = P.split([1, 1, 1], dim=-1)
    valid_u = (u > border) & (u < img_size[1] - 1 - border)
    valid_v = (v > border) & (v < img_size[0] - 1 - border)
    valid_z = z > z_eps
    valid = valid_u & valid_v & valid_z
    logz = torch.log(z)
    logz[torch.logical_not(valid_z)] = 0.0
    pz = torch.cat((p, logz), dim=-1)
    fx, fy = K[0, 0], K[1, 1]
    z_inv = 1.0 / z[..., 0]
    dpz_dP = torch.zeros(*b + (3, 3), dtype=P.dtype)
    dpz_dP[..., 0, 0] = fx
    dpz_dP[..., 1, 1] = fy
    dpz_dP[..., 0, 2] = -fx * x[..., 0] * z_inv
    dpz_dP[..., 1, 2] = -fy * y[..., 0] * z_inv
    dpz_dP *= z_inv[..., None, None]
    dpz_dP[..., 2, 2] = z_inv
    return pz, dpz_dP, valid


def calib_measurements(Xk, img_size, depth_eps=1e-6):
    """meas_k / valid_meas_k as FrameTracker.get_points_poses builds them (tracker.py:143-152) from
    keyframe points that already lie on their pixel rays (constrain_points_to_ray)."""
    h, w = img_size
    u, v = torch.meshgrid(torch.arange(w), torch.arange(h), indexing="xy")
    uv_k = torch.stack((u, v), dim=-1).to(Xk.dtype).view(-1, 2)
    meas_k = torch.cat((uv_k, torch.log(Xk[..., 2:3])), dim=-1)
    valid_meas_k = Xk[..., 2:3] > depth_eps
    meas_k[~valid_meas_k.repeat(1, 3)] = 0.0
    return meas_k.contiguous(), valid_meas_k


def opt_pose_calib_sim3(Xf, Xk, T_WCf, T_WCk, Qk, valid, meas_k, valid_meas_k, K, img_size, sigma_pixel=1.0,
                        sigma_depth=10.0, huber_k=1.345, pixel_border=-10, depth_eps=1e-6, max_iters=50,
                        rel_error=1e-3, delta_norm=1e-3):
    sqrt_info_pixel = 1 / sigma_pixel * valid * torch.sqrt(Qk)
    sqrt_info_depth = 1 / sigma_depth * valid * torch.sqrt(Qk)
    sqrt_info = torch.cat((sqrt_info_pixel.repeat(1, 2), sqrt_info_depth), dim=1)
    T_CkCf = sim3_mul(sim3_inv(T_WCk), T_WCf)
    old_cost = float("inf")
    steps = 0
    for step in range(max_iters):
        Xf_Ck, dXf_Ck_dT_CkCf = act_Sim3(T_CkCf, Xf)
        pzf_Ck, dpzf_Ck_dXf_Ck, valid_proj = project_calib(Xf_Ck, K, img_size, border=pixel_border, z_eps=depth_eps)
        valid2 = valid_proj & valid_meas_k
        sqrt_info2 = valid2 * sqrt_info
        r = meas_k - pzf_Ck
        J = -dpzf_Ck_dXf_Ck @ dXf_Ck_dT_CkCf
        tau, new_cost = solve(sqrt_info2, r, J, huber_k)
        T_CkCf = sim3_retr(T_CkCf, tau)
        steps = step + 1
        if check_convergence(rel_error, delta_norm, old_cost, new_cost, tau):
            break
        old_cost = new_cost
    return sim3_mul(T_WCk, T_CkCf), T_CkCf, steps
