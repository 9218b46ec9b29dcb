"""CPU ORACLE (TEST INFRASTRUCTURE ONLY) -- PyTorch-CPU float32 restatement of the reference
rasterizer.  Nothing in the product path (splatam_b200/, the import-name shims) may import this
file; only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs do.

Parity pin: the reference repo ships NO golden vectors or tests for this path.
The pins are (a) tests/golden/*.npz -- outputs of the UNMODIFIED reference CUDA extensions
(oracle/_ref, built by oracle/build_ref.py) on seeded inputs, generated on a B200 by
tests/golden/make_golden.py -- and (b) the live reference extensions in the GPU tests.  Until
those fixtures exist in the tree this oracle is "parity unpinned".

Every function cites the reference lines it restates (HX = /root/reference/
hessian-diff-gaussian-rasterization-w-depth).  Arithmetic is float32 with the reference's
operation order where that is cheap; the CPU cannot reproduce nvcc's FMA contraction, so quantities
that sit exactly on a discrete threshold (ceil of the radius, tile-rect truncation, alpha <
1/255, T < 1e-4) can flip -- `preprocess()` therefore also returns a `borderline` mask.
"""
import math

import numpy as np
import torch

TILE = 16


# --------------------------------------------------------------------------------------------
# preprocess forward  (HX/cuda_rasterizer/forward.cu:74-152,155-256; auxiliary.h:41-77,139-163)
# --------------------------------------------------------------------------------------------
def _rot_matrix(q):
    """Row-major math matrix R^T of forward.cu:133-137 is built column-major in glm; we return the
    glm matrix as M[c][r] semantics translated to a math matrix G with G[r, c] = glm[c][r]."""
    r, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    # glm columns: col0 = (1-2(yy+zz), 2(xy-rz), 2(xz+ry)), ...
    c0 = torch.stack([1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y)], 1)
    c1 = torch.stack([2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x)], 1)
    c2 = torch.stack([2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)], 1)
    return torch.stack([c0, c1, c2], 2)  # [P, row, col]


def cov3d(scales, mod, rots):
    """Sigma = M^T M, M = S R (forward.cu:118-152); returns [P,6] (xx,xy,xz,yy,yz,zz)."""
    R = _rot_matrix(rots)
    S = torch.diag_embed(scales * mod)
    M = S @ R
    Sg = M.transpose(1, 2) @ M
    return torch.stack([Sg[:, 0, 0], Sg[:, 0, 1], Sg[:, 0, 2], Sg[:, 1, 1], Sg[:, 1, 2], Sg[:, 2, 2]], 1)


def _sym(c6):
    return torch.stack([torch.stack([c6[:, 0], c6[:, 1], c6[:, 2]], 1),
                        torch.stack([c6[:, 1], c6[:, 3], c6[:, 4]], 1),
                        torch.stack([c6[:, 2], c6[:, 4], c6[:, 5]], 1)], 1)


def _view_parts(means, V, fx, fy, tanx, tany):
    """t (clamped), J (2x3), W (3x3 rotation of the view matrix) as in forward.cu:80-99."""
    Vm = V.reshape(4, 4)  # "transposed" storage: math element (r,c) = Vm[c, r]
    Wm = Vm[:3, :3].t().contiguous()       # math rotation
    t = means @ Wm.t() + Vm[3, :3]
    limx, limy = 1.3 * tanx, 1.3 * tany
    txtz, tytz = t[:, 0] / t[:, 2], t[:, 1] / t[:, 2]
    tx = torch.clamp(txtz, -limx, limx) * t[:, 2]
    ty = torch.clamp(tytz, -limy, limy) * t[:, 2]
    tz = t[:, 2]
    zero = torch.zeros_like(tz)
    J = torch.stack([torch.stack([fx / tz, zero, -(fx * tx) / (tz * tz)], 1),
                     torch.stack([zero, fy / tz, -(fy * ty) / (tz * tz)], 1)], 1)  # [P,2,3]
    return t, tx, ty, tz, txtz, tytz, J, Wm


def cov2d(means, cov3, V, fx, fy, tanx, tany):
    """(a,b,c) with +0.3 dilation (forward.cu:74-113).  cov = (J W) Sigma (J W)^T."""
    _, _, _, _, _, _, J, Wm = _view_parts(means, V, fx, fy, tanx, tany)
    A = J @ Wm                                # [P,2,3]
    c = A @ _sym(cov3) @ A.transpose(1, 2)    # [P,2,2]
    return c[:, 0, 0] + 0.3, c[:, 0, 1], c[:, 1, 1] + 0.3


def preprocess(means, scales, rots, opac, V, Pm, W, H, tanx, tany, near, mod=1.0, cov3_pre=None):
    """Returns dict(radii int32[P], xy [P,2], depth [P], conic_opacity [P,4], rect [P,4] int
    (x0,y0,x1,y1), tiles_touched [P], borderline bool[P])."""
    f32 = torch.float32
    P = means.shape[0]
    fx, fy = W / (2.0 * tanx), H / (2.0 * tany)
    gx, gy = (W + TILE - 1) // TILE, (H + TILE - 1) // TILE
    Vm, Pmm = V.reshape(4, 4), Pm.reshape(4, 4)
    pview_z = means @ Vm[:3, 2] + Vm[3, 2]
    ph = means @ Pmm[:3, :] + Pmm[3, :]
    pw = 1.0 / (ph[:, 3] + 0.0000001)
    ppx, ppy = ph[:, 0] * pw, ph[:, 1] * pw
    c3 = cov3_pre if cov3_pre is not None else cov3d(scales, mod, rots)
    zsafe = pview_z > near
    a, b, c = cov2d(means, c3, V, fx, fy, tanx, tany)
    det = a * c - b * b
    det_inv = 1.0 / det
    conic = torch.stack([c * det_inv, -b * det_inv, a * det_inv], 1)
    mid = 0.5 * (a + c)
    disc = torch.sqrt(torch.clamp(mid * mid - det, min=0.1))
    lam = torch.maximum(mid + disc, mid - disc)
    rad_f = 3.0 * torch.sqrt(lam)
    rad = torch.ceil(rad_f)
    # ndc2Pix in double (auxiliary.h:41-44)
    px = (((ppx.double() + 1.0) * W - 1.0) * 0.5).to(f32)
    py = (((ppy.double() + 1.0) * H - 1.0) * 0.5).to(f32)

    def trunc_div(v):
        return torch.trunc(v / TILE).to(torch.int64)
    x0 = torch.clamp(trunc_div(px - rad), 0, gx); y0 = torch.clamp(trunc_div(py - rad), 0, gy)
    x1 = torch.clamp(trunc_div(px + rad + TILE - 1), 0, gx); y1 = torch.clamp(trunc_div(py + rad + TILE - 1), 0, gy)
    ok = zsafe & (det != 0) & torch.isfinite(rad) & (((x1 - x0) * (y1 - y0)) != 0)
    radii = torch.where(ok, rad, torch.zeros_like(rad)).to(torch.int32)
    tiles = torch.where(ok, (x1 - x0) * (y1 - y0), torch.zeros_like(x0))
    # borderline: anything within float noise of a discrete decision
    eps = 2e-4
    frac = rad_f - torch.floor(rad_f)
    bl = (frac < eps * rad_f.clamp(min=1)) | (frac > 1 - eps * rad_f.clamp(min=1))
    for vv in (px - rad, px + rad + TILE - 1, py - rad, py + rad + TILE - 1):
        q = vv / TILE
        bl |= (q - torch.round(q)).abs() < 1e-4
    bl |= (pview_z - near).abs() < 1e-6
    return dict(radii=radii, xy=torch.stack([px, py], 1), depth=pview_z.to(f32),
                conic_opacity=torch.cat([conic, opac.reshape(P, 1)], 1),
                rect=torch.stack([x0, y0, x1, y1], 1), tiles_touched=tiles, borderline=bl & zsafe,
                cov3D=c3, gx=gx, gy=gy)


# --------------------------------------------------------------------------------------------
# binning  (HX/cuda_rasterizer/rasterizer_impl.cu:70-138, 277-319)
# --------------------------------------------------------------------------------------------
def binning(pre):
    """Returns (point_list int64[R], ranges int64[T,2]).  Order = stable sort of keys
    (tile << 32 | depth bits) over instances emitted in ascending Gaussian index."""
    radii = pre["radii"].numpy()
    rect = pre["rect"].numpy()
    gx, gy = pre["gx"], pre["gy"]
    depth_bits = pre["depth"].numpy().view(np.uint32).astype(np.uint64)
    vis = np.nonzero(radii > 0)[0]
    ids, tiles = [], []
    w = (rect[vis, 2] - rect[vis, 0]); hgt = (rect[vis, 3] - rect[vis, 1])
    cnt = (w * hgt).astype(np.int64)
    rep = np.repeat(np.arange(vis.size), cnt)
    start = np.cumsum(cnt) - cnt
    local = np.arange(cnt.sum()) - np.repeat(start, cnt)
    ww = np.repeat(w, cnt)
    ty = np.repeat(rect[vis, 1], cnt) + local // np.maximum(ww, 1)
    tx = np.repeat(rect[vis, 0], cnt) + local % np.maximum(ww, 1)
    gid = vis[rep]
    keys = ((ty * gx + tx).astype(np.uint64) << np.uint64(32)) | depth_bits[gid]
    order = np.argsort(keys, kind="stable")
    keys, point_list = keys[order], gid[order]
    T = gx * gy
    ranges = np.zeros((T, 2), dtype=np.int64)
    if keys.size:
        tile_of = (keys >> np.uint64(32)).astype(np.int64)
        first = np.nonzero(np.diff(tile_of, prepend=-1))[0]
        last = np.append(first[1:], keys.size)
        ranges[tile_of[first], 0] = first
        ranges[tile_of[first], 1] = last
    return torch.from_numpy(point_list.astype(np.int64)), torch.from_numpy(ranges)


# --------------------------------------------------------------------------------------------
# blend forward  (HX/cuda_rasterizer/forward.cu:303-392)
# --------------------------------------------------------------------------------------------
def _tile_pixels(W, H, gx, gy):
    t = torch.arange(gx * gy)
    tx, ty = (t % gx) * TILE, (t // gx) * TILE
    l = torch.arange(TILE * TILE)
    px = tx[:, None] + (l % TILE)[None, :]
    py = ty[:, None] + (l // TILE)[None, :]
    return px, py, (px < W) & (py < H)


def render_forward(pre, point_list, ranges, colors, bg, W, H):
    """Tile-list compositing, vectorised over all tiles; one Python step per list position.
    Returns dict(color [3,H,W], depth [1,H,W], final_T [H,W], n_contrib [H,W] int64,
    pairs_tested, pairs_blended)."""
    gx, gy = pre["gx"], pre["gy"]
    T = gx * gy
    px, py, inside = _tile_pixels(W, H, gx, gy)
    pxf, pyf = px.float(), py.float()
    lens = (ranges[:, 1] - ranges[:, 0])
    order = torch.argsort(lens, descending=True, stable=True)
    lens_s = lens[order]
    start_s = ranges[order, 0]
    pxs, pys, ins = pxf[order], pyf[order], inside[order]
    Tr = torch.ones(T, 256); C = torch.zeros(T, 256, 3); D = torch.full((T, 256), 15.0)
    last = torch.zeros(T, 256, dtype=torch.int64); done = ~ins
    xy, co, depth = pre["xy"], pre["conic_opacity"], pre["depth"]
    tested = 0; blended = 0
    maxlen = int(lens_s[0]) if T else 0
    for j in range(maxlen):
        k = int((lens_s > j).sum())
        if k == 0 or bool(done[:k].all()):
            break
        gid = point_list[start_s[:k] + j]
        g_xy, g_co = xy[gid], co[gid]
        dx = g_xy[:, 0:1] - pxs[:k]; dy = g_xy[:, 1:2] - pys[:k]
        power = -0.5 * (g_co[:, 0:1] * dx * dx + g_co[:, 2:3] * dy * dy) - g_co[:, 1:2] * dx * dy
        alpha = torch.clamp(g_co[:, 3:4] * torch.exp(power), max=0.99)
        live = ~done[:k]
        tested += int(live.sum())
        ok = live & ~(power > 0) & ~(alpha < 1.0 / 255.0)
        Tk = Tr[:k]
        test_T = Tk * (1 - alpha)
        term = ok & (test_T < 0.0001)
        done[:k] |= term
        ok = ok & ~term
        blended += int(ok.sum())
        w = torch.where(ok, alpha * Tk, torch.zeros(()))
        C[:k] += colors[gid][:, None, :] * w[:, :, None]
        med = ok & (Tk > 0.5) & (test_T < 0.5)
        D[:k] = torch.where(med, depth[gid][:, None].expand(-1, 256), D[:k])
        Tr[:k] = torch.where(ok, test_T, Tk)
        last[:k] = torch.where(ok, torch.full((), j + 1, dtype=torch.int64), last[:k])
    inv = torch.empty_like(order); inv[order] = torch.arange(T)
    Tr, C, D, last = Tr[inv], C[inv], D[inv], last[inv]
    out_c = torch.zeros(3, H, W); out_d = torch.zeros(1, H, W)
    fT = torch.zeros(H, W); nc = torch.zeros(H, W, dtype=torch.int64)
    m = inside
    out = C + Tr[:, :, None] * bg.reshape(1, 1, 3)
    for ch in range(3):
        out_c[ch][py[m], px[m]] = out[:, :, ch][m]
    out_d[0][py[m], px[m]] = D[m]; fT[py[m], px[m]] = Tr[m]; nc[py[m], px[m]] = last[m]
    return dict(color=out_c, depth=out_d, final_T=fT, n_contrib=nc, pairs_tested=tested, pairs_blended=blended)


# --------------------------------------------------------------------------------------------
# per-Gaussian backward chain
#   cov2D part: HX/cuda_rasterizer/backward.cu:144-274 (explicit formulas, incl. 1/(denom^2+1e-7)
#               and the clamp masks); projection part: backward.cu:504-516;
#   cov3D -> scale/rotation: backward.cu:412-475, restated through autograd of cov3d() above
#               (same un-normalised quaternion convention; no `mod` factor on dL_dscale).
# Linear map (dconic.x,.y,.w, dmean2D.x,.y) -> (dmean3D[3], dcov3D[6], dscale[3], drot[4]).
# --------------------------------------------------------------------------------------------
def chain_backward(means, scales, rots, cov3, V, Pm, W, H, tanx, tany, dconic, dmean2d, mod=1.0):
    fx, fy = W / (2.0 * tanx), H / (2.0 * tany)
    t, tx, ty, tz, txtz, tytz, J, Rv = _view_parts(means, V, fx, fy, tanx, tany)
    limx, limy = 1.3 * tanx, 1.3 * tany
    xmul = ((txtz >= -limx) & (txtz <= limx)).float()
    ymul = ((tytz >= -limy) & (tytz <= limy)).float()
    A = J @ Rv
    Vrk = _sym(cov3)
    c2 = A @ Vrk @ A.transpose(1, 2)
    a, b, c = c2[:, 0, 0] + 0.3, c2[:, 0, 1], c2[:, 1, 1] + 0.3
    denom = a * c - b * b
    d2 = 1.0 / (denom * denom + 0.0000001)
    dcx, dcy, dcw = dconic[:, 0], dconic[:, 1], dconic[:, 2]
    da = d2 * (-c * c * dcx + 2 * b * c * dcy + (denom - a * c) * dcw)
    dc = d2 * (-a * a * dcw + 2 * a * b * dcy + (denom - a * c) * dcx)
    db = d2 * 2 * (b * c * dcx - (denom + 2 * b * b) * dcy + a * b * dcw)
    A0, A1 = A[:, 0, :], A[:, 1, :]
    G = torch.stack([torch.stack([da, 0.5 * db], 1), torch.stack([0.5 * db, dc], 1)], 1)
    dS = A.transpose(1, 2) @ G @ A                      # symmetric [P,3,3]
    dcov = torch.stack([dS[:, 0, 0], 2 * dS[:, 0, 1], 2 * dS[:, 0, 2], dS[:, 1, 1], 2 * dS[:, 1, 2], dS[:, 2, 2]], 1)
    r0 = (A0[:, None, :] @ Vrk)[:, 0, :]; r1 = (A1[:, None, :] @ Vrk)[:, 0, :]
    dT0 = 2 * r0 * da[:, None] + r1 * db[:, None]
    dT1 = 2 * r1 * dc[:, None] + r0 * db[:, None]
    dJ00 = dT0 @ Rv[0]; dJ02 = dT0 @ Rv[2]; dJ11 = dT1 @ Rv[1]; dJ12 = dT1 @ Rv[2]
    iz = 1.0 / tz; iz2 = iz * iz; iz3 = iz2 * iz
    dtx = xmul * -fx * iz2 * dJ02
    dty = ymul * -fy * iz2 * dJ12
    dtz = -fx * iz2 * dJ00 - fy * iz2 * dJ11 + (2 * fx * tx) * iz3 * dJ02 + (2 * fy * ty) * iz3 * dJ12
    dmean = torch.stack([dtx, dty, dtz], 1) @ Rv        # Rv^T dt
    Pmm = Pm.reshape(4, 4)
    mh = means @ Pmm[:3, :] + Pmm[3, :]
    mw = 1.0 / (mh[:, 3] + 0.0000001)
    mul1 = mh[:, 0] * mw * mw; mul2 = mh[:, 1] * mw * mw
    dmx, dmy = dmean2d[:, 0], dmean2d[:, 1]
    proj = torch.stack([(Pmm[k, 0] * mw - Pmm[k, 3] * mul1) * dmx + (Pmm[k, 1] * mw - Pmm[k, 3] * mul2) * dmy
                        for k in range(3)], 1)
    dmean = dmean + proj
    dscale = torch.zeros_like(means); drot = torch.zeros(means.shape[0], 4)
    if scales is not None:
        with torch.enable_grad():
            s_ = (scales * mod).detach().clone().requires_grad_(True)
            q_ = rots.detach().clone().requires_grad_(True)
            c6 = cov3d(s_, 1.0, q_)
            (c6 * dcov.detach()).sum().backward()
        dscale, drot = s_.grad, q_.grad
    return dmean, dcov, dscale, drot


# --------------------------------------------------------------------------------------------
# blend backward  (HX/cuda_rasterizer/backward.cu:626-746; power semantics backward.cu:1041-1137)
# --------------------------------------------------------------------------------------------
def render_backward(pre, point_list, ranges, colors, bg, W, H, fwd, dL_dpix, power=1, chain=None):
    """power==1: returns sums (dcolor [P,3], dmean2D [P,2], dconic [P,3] (x,y,w), dopacity [P]).
    power!=1 (HX fused): `chain(v5, gid)` must map per-pair (dconic xyw, dmean2D xy) [n,5] and
    Gaussian ids [n] to the 16 chained outputs [n,16]; returns the 25 powered sums:
    dict(dcolor, dmean2D, dconic, dopacity, dmean3D, dcov3D, dscale, drot)."""
    gx, gy = pre["gx"], pre["gy"]
    T = gx * gy
    P = pre["xy"].shape[0]
    px, py, inside = _tile_pixels(W, H, gx, gy)
    lens = ranges[:, 1] - ranges[:, 0]
    order = torch.argsort(lens, descending=True, stable=True)
    lens_s, start_s = lens[order], ranges[order, 0]
    pxs, pys, ins = px[order], py[order], inside[order]
    pxc, pyc = pxs.clamp(max=W - 1), pys.clamp(max=H - 1)
    T_final = torch.where(ins, fwd["final_T"][pyc, pxc], torch.zeros(()))
    last = torch.where(ins, fwd["n_contrib"][pyc, pxc], torch.zeros((), dtype=torch.int64))
    dp = torch.where(ins[:, :, None], dL_dpix[:, pyc, pxc].permute(1, 2, 0), torch.zeros(()))  # [T,256,3]
    bg_dot = (dp * bg.reshape(1, 1, 3)).sum(2)
    Tr = T_final.clone()
    accum = torch.zeros(T, 256, 3); last_alpha = torch.zeros(T, 256); last_color = torch.zeros(T, 256, 3)
    xy, co = pre["xy"], pre["conic_opacity"]
    pw = float(power)
    out = dict(dcolor=torch.zeros(P, 3), dmean2D=torch.zeros(P, 2), dconic=torch.zeros(P, 3), dopacity=torch.zeros(P))
    if chain is not None:
        out.update(dmean3D=torch.zeros(P, 3), dcov3D=torch.zeros(P, 6), dscale=torch.zeros(P, 3), drot=torch.zeros(P, 4))
    maxlen = int(min(int(lens_s[0]), int(last.max()))) if T else 0
    for j in range(maxlen - 1, -1, -1):
        k = int((lens_s > j).sum())
        if k == 0:
            continue
        gid = point_list[start_s[:k] + j]
        g_xy, g_co = xy[gid], co[gid]
        dx = g_xy[:, 0:1] - pxs[:k].float(); dy = g_xy[:, 1:2] - pys[:k].float()
        A_, B_, C_, o_ = g_co[:, 0:1], g_co[:, 1:2], g_co[:, 2:3], g_co[:, 3:4]
        power_ = -0.5 * (A_ * dx * dx + C_ * dy * dy) - B_ * dx * dy
        Gs = torch.exp(power_)
        alpha = torch.clamp(o_ * Gs, max=0.99)
        ok = (j < last[:k]) & ~(power_ > 0) & ~(alpha < 1.0 / 255.0)
        if not bool(ok.any()):
            continue
        Tk = torch.where(ok, Tr[:k] / (1 - alpha), Tr[:k])
        Tr[:k] = Tk
        dchan = alpha * Tk
        col = colors[gid][:, None, :]                                   # [k,1,3]
        acc_new = last_alpha[:k, :, None] * last_color[:k] + (1 - last_alpha[:k, :, None]) * accum[:k]
        accum[:k] = torch.where(ok[:, :, None], acc_new, accum[:k])
        last_color[:k] = torch.where(ok[:, :, None], col.expand(-1, 256, -1), last_color[:k])
        dL_dalpha = ((col - accum[:k]) * dp[:k]).sum(2) * Tk
        last_alpha[:k] = torch.where(ok, alpha, last_alpha[:k])
        dL_dalpha = dL_dalpha + (-T_final[:k] / (1 - alpha)) * bg_dot[:k]
        dL_dG = o_ * dL_dalpha
        gdx, gdy = Gs * dx, Gs * dy
        dG_dx = -gdx * A_ - gdy * B_
        dG_dy = -gdy * C_ - gdx * B_
        z = torch.zeros(())
        okf = ok
        c_ = torch.where(okf[:, :, None], dchan[:, :, None] * dp[:k], z)
        mx = torch.where(okf, dL_dG * dG_dx * (0.5 * W), z); my = torch.where(okf, dL_dG * dG_dy * (0.5 * H), z)
        qx = torch.where(okf, -0.5 * gdx * dx * dL_dG, z); qy = torch.where(okf, -0.5 * gdx * dy * dL_dG, z)
        qw = torch.where(okf, -0.5 * gdy * dy * dL_dG, z)
        op = torch.where(okf, Gs * dL_dalpha, z)
        if chain is None:
            out["dcolor"].index_add_(0, gid, c_.sum(1))
            out["dmean2D"].index_add_(0, gid, torch.stack([mx.sum(1), my.sum(1)], 1))
            out["dconic"].index_add_(0, gid, torch.stack([qx.sum(1), qy.sum(1), qw.sum(1)], 1))
            out["dopacity"].index_add_(0, gid, op.sum(1))
        else:
            def powsum(v):
                return torch.where(okf if v.dim() == 2 else okf[:, :, None], torch.pow(v, pw), z).sum(1)
            out["dcolor"].index_add_(0, gid, powsum(c_))
            out["dmean2D"].index_add_(0, gid, torch.stack([powsum(mx), powsum(my)], 1))
            out["dconic"].index_add_(0, gid, torch.stack([powsum(qx), powsum(qy), powsum(qw)], 1))
            out["dopacity"].index_add_(0, gid, powsum(op))
            v5 = torch.stack([qx, qy, qw, mx, my], 2).reshape(-1, 5)
            g16 = chain(v5, gid.repeat_interleave(256)).reshape(k, 256, 16)
            g16 = torch.where(okf[:, :, None], torch.pow(g16, pw), z).sum(1)
            out["dmean3D"].index_add_(0, gid, g16[:, 0:3]); out["dcov3D"].index_add_(0, gid, g16[:, 3:9])
            out["dscale"].index_add_(0, gid, g16[:, 9:12]); out["drot"].index_add_(0, gid, g16[:, 12:16])
    return out


# --------------------------------------------------------------------------------------------
# top level: the reference's forward / backward entry points
#   (HX/cuda_rasterizer/rasterizer_impl.cu:198-339, 343-433)
# --------------------------------------------------------------------------------------------
def forward(rs, near, means3D, colors_precomp, opacities, scales, rotations):
    """rs: GaussianRasterizationSettings with CPU tensors.  Returns a ctx dict with the outputs
    color/radii/depth and everything the backward needs."""
    W, H = int(rs.image_width), int(rs.image_height)
    V, Pm = rs.viewmatrix.reshape(16).float(), rs.projmatrix.reshape(16).float()
    pre = preprocess(means3D, scales, rotations, opacities, V, Pm, W, H, rs.tanfovx, rs.tanfovy, near,
                     mod=rs.scale_modifier)
    point_list, ranges = binning(pre)
    fwd = render_forward(pre, point_list, ranges, colors_precomp, rs.bg.float(), W, H)
    return dict(rs=rs, near=near, pre=pre, point_list=point_list, ranges=ranges, fwd=fwd,
                means3D=means3D, colors=colors_precomp, scales=scales, rotations=rotations,
                color=fwd["color"], radii=pre["radii"], depth=fwd["depth"],
                num_rendered=int(point_list.numel()))


def backward(ctx, dL_dpix, power=1):
    """Gradients in the reference's return layout (HX/rasterize_points.cu:151-159,195):
    dict(means3D [P,3], means2D [P,3], colors [P,3], opacities [P,1], scales [P,3], rotations [P,4],
    cov3D [P,6], conic [P,2,2]).  power != 1 follows renderCUDAFused (per-pair terms ** power)."""
    rs, pre = ctx["rs"], ctx["pre"]
    W, H = int(rs.image_width), int(rs.image_height)
    V, Pm = rs.viewmatrix.reshape(16).float(), rs.projmatrix.reshape(16).float()
    means, scales, rots = ctx["means3D"], ctx["scales"], ctx["rotations"]
    P = means.shape[0]
    vis = pre["radii"] > 0
    safe_means = torch.where(vis[:, None], means, torch.tensor([0.0, 0.0, 1.0]).expand(P, 3))

    def chain_all(dconic3, dmean2):
        dm, dcv, ds, dr = chain_backward(safe_means, scales, rots, pre["cov3D"], V, Pm, W, H, rs.tanfovx,
                                         rs.tanfovy, dconic3, dmean2, mod=rs.scale_modifier)
        return dm, dcv, ds, dr

    g = dict(means2D=torch.zeros(P, 3), conic=torch.zeros(P, 2, 2))
    if power == 1:
        r = render_backward(pre, ctx["point_list"], ctx["ranges"], ctx["colors"], rs.bg.float(), W, H,
                            ctx["fwd"], dL_dpix)
        dm, dcv, ds, dr = chain_all(r["dconic"], r["dmean2D"])
    else:
        # Jacobian of the (linear) chain per Gaussian: push the 5 unit vectors through it
        cols = []
        for p in range(5):
            e = torch.zeros(P, 5); e[:, p] = 1.0
            dm, dcv, ds, dr = chain_all(e[:, 0:3], e[:, 3:5])
            cols.append(torch.cat([dm, dcv, ds, dr], 1))
        Jm = torch.stack(cols, 2)                       # [P,16,5]

        def chain(v5, gid):
            return (Jm[gid] @ v5[:, :, None])[:, :, 0]
        r = render_backward(pre, ctx["point_list"], ctx["ranges"], ctx["colors"], rs.bg.float(), W, H,
                            ctx["fwd"], dL_dpix, power=power, chain=chain)
        dm, dcv, ds, dr = r["dmean3D"], r["dcov3D"], r["dscale"], r["drot"]
    z = torch.zeros(())
    g["means3D"] = torch.where(vis[:, None], dm, z)
    g["cov3D"] = torch.where(vis[:, None], dcv, z)
    g["scales"] = torch.where(vis[:, None], ds, z)
    g["rotations"] = torch.where(vis[:, None], dr, z)
    g["colors"] = r["dcolor"]
    g["opacities"] = r["dopacity"].reshape(P, 1)
    g["means2D"][:, :2] = r["dmean2D"]
    g["conic"][:, 0, 0] = r["dconic"][:, 0]; g["conic"][:, 0, 1] = r["dconic"][:, 1]; g["conic"][:, 1, 1] = r["dconic"][:, 2]
    return g
