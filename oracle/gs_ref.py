"""CPU oracle (numpy, fp32) of the FORWARD pass of the reference's 3DGS tile rasteriser.  TEST INFRASTRUCTURE ONLY:
imported by tests/ and bench.py's cpu legs, never by the product path.

Restates, function by function, thirdparty/diff-gaussian-rasterization-modified/cuda_rasterizer:
  preprocess()  forward.cu:157-260 (preprocessCUDA), :75-113 (computeCov2D), :118-153 (computeCov3D),
                :20-72 (computeColorFromSH, e3nn axis convention), auxiliary.h:63-78 (ndc2Pix, getRect), :161-186 (near cull)
  render()      rasterizer_impl.cu:78-150 (instances per 16x16 tile, (tile | depth) order, ties in splat order) and
                forward.cu:262-370 (renderCUDA: front-to-back alpha blending with the 0.99 / 1/255 / 1e-4 cut-offs)
Pinned against the reference's own CUDA build (oracle/build_gs_ref.py -> oracle/_ref/diffgauss_ref.so) by
tests/test_gpu_gs.py on the GPU box; the floating-point bar is written there.
"""
import numpy as np

TILE = 16
F = np.float32
SH_C0 = F(0.28209479177387814)
SH_C1 = F(0.4886025119029199)
SH_C2 = np.array([1.0925484305920792, -1.0925484305920792, 0.31539156525252005, -1.0925484305920792, 0.5462742152960396], dtype=F)
SH_C3 = np.array([-0.5900435899266435, 2.890611442640554, -0.4570457994644658, 0.3731763325901154, -0.4570457994644658,
                  1.445305721320277, -0.5900435899266435], dtype=F)


def _sh_color(means, campos, shs, deg):
    d = means - campos[None]
    d = d / np.linalg.norm(d, axis=1, keepdims=True).astype(F)
    x, y, z = d[:, 0:1], d[:, 1:2], d[:, 2:3]
    res = SH_C0 * shs[:, 0]
    if deg > 0:
        res = res - SH_C1 * x * shs[:, 1] + SH_C1 * y * shs[:, 2] - SH_C1 * z * shs[:, 3]
        if deg > 1:
            zz, xx, yy, zx, xy, zy = z * z, x * x, y * y, z * x, x * y, z * y
            res = (res + SH_C2[0] * zx * shs[:, 4] + SH_C2[1] * xy * shs[:, 5] + SH_C2[2] * (F(2) * yy - zz - xx) * shs[:, 6]
                   + SH_C2[3] * zy * shs[:, 7] + SH_C2[4] * (zz - xx) * shs[:, 8])
            if deg > 2:
                res = (res + SH_C3[0] * x * (F(3) * zz - xx) * shs[:, 9] + SH_C3[1] * zx * y * shs[:, 10]
                       + SH_C3[2] * x * (F(4) * yy - zz - xx) * shs[:, 11]
                       + SH_C3[3] * y * (F(2) * yy - F(3) * zz - F(3) * xx) * shs[:, 12]
                       + SH_C3[4] * z * (F(4) * yy - zz - xx) * shs[:, 13] + SH_C3[5] * z * (zz - xx) * shs[:, 14]
                       + SH_C3[6] * z * (zz - F(3) * xx) * shs[:, 15])
    return np.maximum(res + F(0.5), F(0)).astype(F)


def _cov3d(scales, rotations, mod):
    s = (F(mod) * scales).astype(F)
    r, x, y, z = (rotations[:, k] for k in range(4))
    # glm columns of R as written at forward.cu:132-136; M = S * R, Sigma = M^T M
    Rc = np.stack([np.stack([1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y)], 1),
                   np.stack([2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x)], 1),
                   np.stack([2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)], 1)], 1).astype(F)  # [P, col, row]
    M = Rc * s[:, None, :]                                  # M[col][row] = s[row] * R[col][row]
    Sig = np.einsum("pak,pbk->pab", M, M).astype(F)          # Sigma[a][b] = sum_k M[a][k] M[b][k]
    return np.stack([Sig[:, 0, 0], Sig[:, 0, 1], Sig[:, 0, 2], Sig[:, 1, 1], Sig[:, 1, 2], Sig[:, 2, 2]], 1).astype(F)


def preprocess(means, cov3d, scales, rotations, scale_modifier, opacities, colors, shs, sh_degree, campos, view, proj,
               tanfovx, tanfovy, H, W):
    """Per splat: radius (0 = culled), tile rectangle, view depth, pixel centre, conic, colour."""
    means = np.asarray(means, F); view = np.asarray(view, F).reshape(16); proj = np.asarray(proj, F).reshape(16)
    P = means.shape[0]
    tx_, ty_ = (W + TILE - 1) // TILE, (H + TILE - 1) // TILE
    px, py, pz = means[:, 0], means[:, 1], means[:, 2]
    V, Pm = view, proj
    vx = V[0] * px + V[4] * py + V[8] * pz + V[12]
    vy = V[1] * px + V[5] * py + V[9] * pz + V[13]
    vz = V[2] * px + V[6] * py + V[10] * pz + V[14]
    ok = vz > F(0.2)
    hx = Pm[0] * px + Pm[4] * py + Pm[8] * pz + Pm[12]
    hy = Pm[1] * px + Pm[5] * py + Pm[9] * pz + Pm[13]
    hw = Pm[3] * px + Pm[7] * py + Pm[11] * pz + Pm[15]
    with np.errstate(all="ignore"):
        pw = F(1) / (hw + F(0.0000001))
        projx, projy = hx * pw, hy * pw
        c3 = np.asarray(cov3d, F) if cov3d is not None else _cov3d(np.asarray(scales, F), np.asarray(rotations, F), scale_modifier)
        fy, fx = F(H / (2.0 * tanfovy)), F(W / (2.0 * tanfovx))
        limx, limy = F(1.3 * tanfovx), F(1.3 * tanfovy)
        tz = vz
        tx = np.minimum(limx, np.maximum(-limx, vx / tz)) * tz
        ty = np.minimum(limy, np.maximum(-limy, vy / tz)) * tz
        j00, j02, j11, j12 = fx / tz, -(fx * tx) / (tz * tz), fy / tz, -(fy * ty) / (tz * tz)
        t00, t01, t02 = V[0] * j00 + V[2] * j02, V[4] * j00 + V[6] * j02, V[8] * j00 + V[10] * j02
        t10, t11, t12 = V[1] * j11 + V[2] * j12, V[5] * j11 + V[6] * j12, V[9] * j11 + V[10] * j12
        a0 = c3[:, 0] * t00 + c3[:, 1] * t01 + c3[:, 2] * t02
        a1 = c3[:, 1] * t00 + c3[:, 3] * t01 + c3[:, 4] * t02
        a2 = c3[:, 2] * t00 + c3[:, 4] * t01 + c3[:, 5] * t02
        b0 = c3[:, 0] * t10 + c3[:, 1] * t11 + c3[:, 2] * t12
        b1 = c3[:, 1] * t10 + c3[:, 3] * t11 + c3[:, 4] * t12
        b2 = c3[:, 2] * t10 + c3[:, 4] * t11 + c3[:, 5] * t12
        cxx = t00 * a0 + t01 * a1 + t02 * a2 + F(0.3)
        cxy = t00 * b0 + t01 * b1 + t02 * b2
        cyy = t10 * b0 + t11 * b1 + t12 * b2 + F(0.3)
        det = cxx * cyy - cxy * cxy
        ok &= det != 0
        dinv = F(1) / det
        conic = np.stack([cyy * dinv, -cxy * dinv, cxx * dinv], 1).astype(F)
        mid = F(0.5) * (cxx + cyy)
        root = np.sqrt(np.maximum(F(0.1), mid * mid - det))
        rad = np.ceil(F(3) * np.sqrt(np.maximum(mid + root, mid - root)))
        ix = ((projx + F(1)) * F(W) - F(1)) * F(0.5)
        iy = ((projy + F(1)) * F(H) - F(1)) * F(0.5)
        rad = np.where(np.isfinite(rad), rad, 0).astype(np.int64)
        ix64, iy64 = np.where(np.isfinite(ix), ix, 0), np.where(np.isfinite(iy), iy, 0)
        trunc = lambda v: np.trunc(v).astype(np.int64)
        x0 = np.clip(trunc((ix64 - rad.astype(F)) / F(TILE)), 0, tx_); y0 = np.clip(trunc((iy64 - rad.astype(F)) / F(TILE)), 0, ty_)
        x1 = np.clip(trunc((ix64 + rad.astype(F) + F(TILE - 1)) / F(TILE)), 0, tx_)
        y1 = np.clip(trunc((iy64 + rad.astype(F) + F(TILE - 1)) / F(TILE)), 0, ty_)
    ok &= (x1 - x0) * (y1 - y0) != 0
    if colors is not None:
        rgb = np.asarray(colors, F)
    else:
        rgb = _sh_color(means, np.asarray(campos, F), np.asarray(shs, F), sh_degree)
    radii = np.where(ok, rad, 0).astype(np.int32)
    return dict(ok=ok, radii=radii, rect=np.stack([x0, y0, x1, y1], 1), depth=vz.astype(F), xy=np.stack([ix, iy], 1).astype(F),
                conic=conic, rgb=rgb, opacity=np.asarray(opacities, F).reshape(-1))


def render(pp, bg, H, W):
    """(3,H,W) image from preprocess() output."""
    bg = np.asarray(bg, F)
    out = np.zeros((3, H, W), F)
    tx_, ty_ = (W + TILE - 1) // TILE, (H + TILE - 1) // TILE
    ids = np.nonzero(pp["ok"])[0]
    order = ids[np.argsort(pp["depth"][ids], kind="stable")]  # depth > 0.2: value order == bit order; ties in index order
    per_tile = [[] for _ in range(tx_ * ty_)]
    for g in order:
        x0, y0, x1, y1 = pp["rect"][g]
        for y in range(y0, y1):
            for x in range(x0, x1):
                per_tile[y * tx_ + x].append(g)
    for tyi in range(ty_):
        for txi in range(tx_):
            ys, xs = np.meshgrid(np.arange(tyi * TILE, min(H, (tyi + 1) * TILE)), np.arange(txi * TILE, min(W, (txi + 1) * TILE)), indexing="ij")
            pxf, pyf = xs.astype(F), ys.astype(F)
            T = np.ones(xs.shape, F); C = np.zeros((3,) + xs.shape, F); done = np.zeros(xs.shape, bool)
            for g in per_tile[tyi * tx_ + txi]:
                if done.all():
                    break
                dx, dy = pp["xy"][g, 0] - pxf, pp["xy"][g, 1] - pyf
                a, b, c = pp["conic"][g]
                power = F(-0.5) * (a * dx * dx + c * dy * dy) - b * dx * dy
                with np.errstate(all="ignore"):
                    alpha = np.minimum(F(0.99), pp["opacity"][g] * np.exp(power).astype(F))
                use = (~done) & ~(power > 0) & ~(alpha < F(1.0 / 255.0))
                test_T = T * (F(1) - alpha)
                stop = use & (test_T < F(0.0001))
                done |= stop
                use &= ~stop
                w = np.where(use, alpha * T, F(0))
                C += pp["rgb"][g][:, None, None] * w[None]
                T = np.where(use, test_T, T)
            out[:, ys, xs] = C + T[None] * bg[:, None, None]
    return out


def rasterize(means, cov3d, scales, rotations, scale_modifier, opacities, colors, shs, sh_degree, campos, view, proj, tanfovx,
              tanfovy, bg, H, W):
    pp = preprocess(means, cov3d, scales, rotations, scale_modifier, opacities, colors, shs, sh_degree, campos, view, proj,
                    tanfovx, tanfovy, H, W)
    return render(pp, bg, H, W), pp["radii"]
