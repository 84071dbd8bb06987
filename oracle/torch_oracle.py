"""PyTorch-CPU restatement of the rasterizer equations (differentiable).

TEST INFRASTRUCTURE ONLY (same rule as gsr_oracle.c): used by tests/ as an
independent cross-check of the C oracle -- autograd through this forward gives
gradients that were not hand-derived -- and by bench.py as the "PyTorch-CPU
evaluation of the same rasterization equations" baseline that BASELINE.json's
north_star asks to be reported.  Never imported by the product package.

Follows /root/reference/submodules/diff-gaussian-rasterization/cuda_rasterizer/
forward.cu:20-381 and rasterizer_impl.cu:70-138.  Plain tensor ops: no FMA
contraction, so key-deciding values can differ from the CUDA reference by an
ulp (SURVEY.md section 8c caveat); use gsr_oracle.c for bit-exact keys.
"""
from __future__ import annotations

import math
from typing import Dict, Optional

import torch

SH_C0 = 0.28209479177387814
SH_C1 = 0.4886025119029199
SH_C2 = [1.0925484305920792, -1.0925484305920792, 0.31539156525252005, -1.0925484305920792, 0.5462742152960396]
SH_C3 = [-0.5900435899266435, 2.890611442640554, -0.4570457994644658, 0.3731763325901154,
         -0.4570457994644658, 1.445305721320277, -0.5900435899266435]


def eval_sh_rgb(deg: int, sh: torch.Tensor, dirs: torch.Tensor) -> torch.Tensor:
    """sh (P,M,3), dirs (P,3) normalised.  forward.cu:20-71 without the +0.5/clamp."""
    res = SH_C0 * sh[:, 0]
    if deg > 0:
        x, y, z = dirs[:, 0:1], dirs[:, 1:2], dirs[:, 2:3]
        res = res - SH_C1 * y * sh[:, 1] + SH_C1 * z * sh[:, 2] - SH_C1 * x * sh[:, 3]
        if deg > 1:
            xx, yy, zz, xy, yz, xz = x * x, y * y, z * z, x * y, y * z, x * z
            res = res + SH_C2[0] * xy * sh[:, 4] + SH_C2[1] * yz * sh[:, 5] + SH_C2[2] * (2.0 * zz - xx - yy) * sh[:, 6] \
                + SH_C2[3] * xz * sh[:, 7] + SH_C2[4] * (xx - yy) * sh[:, 8]
            if deg > 2:
                res = res + SH_C3[0] * y * (3.0 * xx - yy) * sh[:, 9] + SH_C3[1] * xy * z * sh[:, 10] \
                    + SH_C3[2] * y * (4.0 * zz - xx - yy) * sh[:, 11] \
                    + SH_C3[3] * z * (2.0 * zz - 3.0 * xx - 3.0 * yy) * sh[:, 12] \
                    + SH_C3[4] * x * (4.0 * zz - xx - yy) * sh[:, 13] + SH_C3[5] * z * (xx - yy) * sh[:, 14] \
                    + SH_C3[6] * x * (xx - 3.0 * yy) * sh[:, 15]
    return res


def _cov3d(scales, rot, mod):
    r, x, y, z = rot[:, 0], rot[:, 1], rot[:, 2], rot[:, 3]
    # rows of the rotation matrix (glm columns transposed): Rm[i][j]
    Rm = torch.stack([
        torch.stack([1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y)], -1),
        torch.stack([2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x)], -1),
        torch.stack([2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)], -1)], 1)
    # glm M = S*R with R given column-wise == (Rm^T scaled per row) ; Sigma = M^T M = Rm diag(s^2) Rm^T
    s = mod * scales
    L = Rm * s[:, None, :]
    Sigma = L @ L.transpose(1, 2)
    return torch.stack([Sigma[:, 0, 0], Sigma[:, 0, 1], Sigma[:, 0, 2], Sigma[:, 1, 1], Sigma[:, 1, 2], Sigma[:, 2, 2]], -1)


def forward(means3D, opacities, *, shs=None, colors_precomp=None, scales=None, rotations=None,
            cov3D_precomp=None, viewmatrix, projmatrix, campos, bg, W, H, tanfovx, tanfovy,
            sh_degree=0, scale_modifier=1.0, tile_subset=None) -> Dict[str, torch.Tensor]:
    """Differentiable forward.  All tensors on CPU, any float dtype (float64 for gradient checks).
    tile_subset: optional iterable of tile ids to blend (bounded CPU-baseline sample)."""
    dt = means3D.dtype
    V = viewmatrix.to(dt).reshape(4, 4)      # V[i][j] = flat[4i+j]; column-major W2C
    Pm = projmatrix.to(dt).reshape(4, 4)
    P = means3D.shape[0]
    ones = torch.ones(P, 1, dtype=dt)
    hom = torch.cat([means3D, ones], 1)
    p_view = hom @ V                          # row-vector times W2C^T
    p_hom = hom @ Pm
    depth = p_view[:, 2]
    front = depth > 0.2
    p_w = 1.0 / (p_hom[:, 3] + 0.0000001)
    proj = p_hom[:, :2] * p_w[:, None]
    fx = W / (2.0 * tanfovx)
    fy = H / (2.0 * tanfovy)
    cov3D = cov3D_precomp if cov3D_precomp is not None else _cov3d(scales, rotations, scale_modifier)
    tz = p_view[:, 2]
    tz_safe = torch.where(front, tz, torch.ones_like(tz))
    limx, limy = 1.3 * tanfovx, 1.3 * tanfovy
    txc = torch.clamp(p_view[:, 0] / tz_safe, -limx, limx) * tz_safe
    tyc = torch.clamp(p_view[:, 1] / tz_safe, -limy, limy) * tz_safe
    zero = torch.zeros_like(tz)
    # J rows (2x3), T = J @ W2C[:3,:3] in row-major maths
    J = torch.stack([torch.stack([fx / tz_safe, zero, -(fx * txc) / (tz_safe * tz_safe)], -1),
                     torch.stack([zero, fy / tz_safe, -(fy * tyc) / (tz_safe * tz_safe)], -1)], 1)
    Wrot = V[:3, :3].t()                       # W2C rotation, row-major
    Tm = J @ Wrot                              # (P,2,3)
    Sig = torch.stack([torch.stack([cov3D[:, 0], cov3D[:, 1], cov3D[:, 2]], -1),
                       torch.stack([cov3D[:, 1], cov3D[:, 3], cov3D[:, 4]], -1),
                       torch.stack([cov3D[:, 2], cov3D[:, 4], cov3D[:, 5]], -1)], 1)
    cov2 = Tm @ Sig @ Tm.transpose(1, 2)
    a = cov2[:, 0, 0] + 0.3
    b = cov2[:, 0, 1]
    c = cov2[:, 1, 1] + 0.3
    det = a * c - b * b
    ok = front & (det != 0)
    det_safe = torch.where(ok, det, torch.ones_like(det))
    conic = torch.stack([c / det_safe, -b / det_safe, a / det_safe], -1)
    mid = 0.5 * (a + c)
    lam = mid + torch.sqrt(torch.clamp(mid * mid - det, min=0.1))
    radius = torch.ceil(3.0 * torch.sqrt(lam.detach().clamp(min=0))).to(torch.int64)
    px = ((proj[:, 0].double() + 1.0) * W - 1.0) * 0.5
    py = ((proj[:, 1].double() + 1.0) * H - 1.0) * 0.5
    px, py = px.to(dt), py.to(dt)
    gx, gy = (W + 15) // 16, (H + 15) // 16
    rf = radius.to(dt)
    pxd, pyd = px.detach(), py.detach()
    x0 = torch.clamp(((pxd - rf) / 16).trunc().long(), 0, gx)
    y0 = torch.clamp(((pyd - rf) / 16).trunc().long(), 0, gy)
    x1 = torch.clamp(((pxd + rf + 15) / 16).trunc().long(), 0, gx)
    y1 = torch.clamp(((pyd + rf + 15) / 16).trunc().long(), 0, gy)
    ntile = (x1 - x0) * (y1 - y0)
    vis = ok & (ntile > 0)
    ntile = torch.where(vis, ntile, torch.zeros_like(ntile))
    radius = torch.where(vis, radius, torch.zeros_like(radius))
    if colors_precomp is not None:
        feat = colors_precomp
        clamped = None
    else:
        d = means3D - campos.to(dt)[None]
        d = d / d.norm(dim=1, keepdim=True)
        raw = eval_sh_rgb(sh_degree, shs, d) + 0.5
        clamped = raw < 0
        feat = torch.clamp(raw, min=0.0)
    # ---- binning: emit in Gaussian order, stable sort by (tile, depth bits) ----
    ids = torch.nonzero(vis).squeeze(1)
    wv = (x1 - x0)[ids]
    cnt = ntile[ids]
    R = int(cnt.sum())
    rep = torch.repeat_interleave(ids, cnt)
    start = torch.cumsum(cnt, 0) - cnt
    local = torch.arange(R) - torch.repeat_interleave(start, cnt)
    wrep = torch.repeat_interleave(wv, cnt)
    ty = y0[rep] + local // wrep
    tx = x0[rep] + local % wrep
    tile = ty * gx + tx
    dbits = depth.detach().float().view(torch.int32).long()[rep]
    keys = (tile << 32) | dbits
    keys_sorted, order = torch.sort(keys, stable=True)
    point_list = rep[order]
    tile_sorted = keys_sorted >> 32
    tiles = gx * gy
    starts = torch.searchsorted(tile_sorted, torch.arange(tiles))
    ends = torch.searchsorted(tile_sorted, torch.arange(tiles), right=True)
    ranges = torch.stack([starts, ends], 1)
    ranges = torch.where((ends > starts)[:, None], ranges, torch.zeros_like(ranges))
    # ---- blending ----
    bgv = bg.to(dt)
    color = torch.zeros(3, H, W, dtype=dt)
    depth_img = torch.zeros(1, H, W, dtype=dt)
    alpha_img = torch.zeros(1, H, W, dtype=dt)
    n_contrib = torch.zeros(H, W, dtype=torch.int64)
    color_rows, tile_list = [], (range(tiles) if tile_subset is None else list(tile_subset))
    xy = torch.stack([px, py], -1)
    for t in tile_list:
        tyi, txi = divmod(t, gx)
        ys = torch.arange(tyi * 16, min(tyi * 16 + 16, H))
        xs = torch.arange(txi * 16, min(txi * 16 + 16, W))
        yy, xx = torch.meshgrid(ys, xs, indexing="ij")
        pixx = xx.reshape(-1).to(dt)
        pixy = yy.reshape(-1).to(dt)
        r0, r1 = int(ranges[t, 0]), int(ranges[t, 1])
        if r1 <= r0:
            for ch in range(3):
                color[ch, yy, xx] = bgv[ch]
            continue
        g = point_list[r0:r1]
        dx = xy[g, 0][None, :] - pixx[:, None]
        dy = xy[g, 1][None, :] - pixy[:, None]
        con = conic[g]
        power = -0.5 * (con[:, 0][None] * dx * dx + con[:, 2][None] * dy * dy) - con[:, 1][None] * dx * dy
        alpha = torch.clamp(opacities.reshape(-1)[g][None] * torch.exp(power), max=0.99)
        valid = (power <= 0) & (alpha >= 1.0 / 255.0)
        a_eff = torch.where(valid, alpha, torch.zeros_like(alpha))
        one_m = 1.0 - a_eff
        T_incl = torch.cumprod(one_m, 1)
        T_excl = torch.cat([torch.ones_like(T_incl[:, :1]), T_incl[:, :-1]], 1)
        stop = valid & (T_incl < 0.0001)
        dead = torch.cumsum(stop.long(), 1) > 0
        contrib = valid & ~dead
        wgt = torch.where(contrib, a_eff * T_excl, torch.zeros_like(a_eff))
        Cc = wgt @ feat[g]
        Dd = wgt @ p_view[:, 2][g]
        Ww = wgt.sum(1)
        T_final = torch.where(contrib, one_m, torch.ones_like(one_m)).prod(1)
        idxs = torch.arange(1, r1 - r0 + 1)[None].expand_as(contrib)
        last = torch.where(contrib, idxs, torch.zeros_like(idxs)).max(1).values
        for ch in range(3):
            color[ch, yy, xx] = (Cc[:, ch] + T_final * bgv[ch]).reshape(yy.shape)
        depth_img[0, yy, xx] = Dd.reshape(yy.shape)
        alpha_img[0, yy, xx] = Ww.reshape(yy.shape)
        n_contrib[yy, xx] = last.reshape(yy.shape)
    return dict(color=color, depth=depth_img, alpha=alpha_img, radii=radius.to(torch.int32), xy=xy,
                depths=depth, conic=conic, rgb=feat, tiles_touched=ntile, keys_sorted=keys_sorted,
                point_list=point_list, ranges=ranges, n_contrib=n_contrib, num_rendered=R,
                clamped=clamped)
