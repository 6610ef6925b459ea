"""CPU oracle of the reference differentiable Gaussian rasterizer  --  TEST INFRASTRUCTURE ONLY.

This file restates, in eager PyTorch on the CPU, the algorithm of the reference CUDA rasterizer
(`R/` = /root/reference/submodules_fsgs/diff-gaussian-rasterization-confidence).  It is the checker for
`tests/`, `__graft_entry__.smoke()` and `bench.py`'s `cpu_baseline` leg.  Nothing in the product package
(`dynamic3dgaussians_b200/`) may import it; the product path has no CPU fallback.

Parity pin status: the reference ships no tests, fixtures or golden vectors for this path, so
the oracle is pinned two other ways:
  * arithmetic: the forward follows, operation by operation, the PTX that nvcc 12.9 emits for the reference
    kernels (which products are fused into FMAs, in which order).  FMAs are emulated exactly in float64.
    Integer artefacts (radii, tile rects, sort keys, sorted lists, tile ranges, n_contrib) are therefore
    expected to agree bit-for-bit with the reference CUDA build, and that is what the GPU tests check
    against `oracle/_ref` (the reference's own sources compiled on the GPU box).
  * gradients: autograd through this forward, cross-checked against float64 finite differences
    (tests/test_oracle.py).  Gradient conventions that are not derivable from the forward are taken from the
    reference backward: straight-through 0.99 alpha clamp (R/cuda_rasterizer/backward.cu:560-565), NDC-scaled
    units for the 2D-mean gradient (backward.cu:520-521,640-641), zero gradient for culled Gaussians
    (backward.cu:158,386).

Reference lines followed:
  preprocess            R/cuda_rasterizer/forward.cu:166-269  (+ auxiliary.h:41-77,145-170)
  computeCov3D          forward.cu:129-163
  computeCov2D          forward.cu:75-124
  computeColorFromSH    forward.cu:20-71
  duplicateWithKeys     R/cuda_rasterizer/rasterizer_impl.cu:70-111
  sort / bit range      rasterizer_impl.cu:35-50,306-314
  identifyTileRanges    rasterizer_impl.cu:116-138
  renderCUDA (forward)  forward.cu:274-408
"""
from __future__ import annotations

import math
from typing import NamedTuple, Optional

import numpy as np
import torch

BLOCK_X = 16
BLOCK_Y = 16
BLOCK_SIZE = BLOCK_X * BLOCK_Y

SH_C0 = 0.28209479177387814
SH_C1 = 0.4886025119029199
SH_C2 = [1.0925484305920792, -1.0925484305920792, 0.31539156525252005, -1.0925484305920792, 0.5462742152960396]
SH_C3 = [-0.5900435899266435, 2.890611442640554, -0.4570457994644658, 0.3731763325901154,
         -0.4570457994644658, 1.445305721320277, -0.5900435899266435]


class Camera(NamedTuple):
    """Per-view camera, same conventions as GaussianRasterizationSettings (R/diff_gaussian_rasterization/__init__.py:176-191).

    viewmatrix / projmatrix are the *transposed* 4x4 matrices of REF/helpers.py:70-77, flattened row-major, so that
    element [i] is what the kernels call matrix[i] (auxiliary.h:58-77).
    """
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    c_x: float
    c_y: float
    bg: torch.Tensor            # [3] (or [C] when extra colour channels are rendered)
    scale_modifier: float
    viewmatrix: torch.Tensor    # [16] or [4,4] or [1,4,4]
    projmatrix: torch.Tensor
    campos: torch.Tensor        # [3]
    sh_degree: int = 0


def get_higher_msb(n: int) -> int:
    """rasterizer_impl.cu:35-50."""
    msb = 16
    step = msb
    while step > 1:
        step //= 2
        if n >> msb:
            msb += step
        else:
            msb -= step
    if n >> msb:
        msb += 1
    return msb


# ----------------------------------------------------------------------------------------------------------------
# arithmetic helpers: float32 with exact FMA emulation, or plain float64 (for finite-difference checks)
# ----------------------------------------------------------------------------------------------------------------
class _Arith:
    def __init__(self, dtype: torch.dtype):
        assert dtype in (torch.float32, torch.float64)
        self.dt = dtype
        self.exact32 = dtype == torch.float32

    def c(self, v):
        return torch.as_tensor(v, dtype=self.dt)

    def fma(self, a, b, c):
        a = torch.as_tensor(a, dtype=self.dt) if not torch.is_tensor(a) else a
        b = torch.as_tensor(b, dtype=self.dt) if not torch.is_tensor(b) else b
        c = torch.as_tensor(c, dtype=self.dt) if not torch.is_tensor(c) else c
        if self.exact32:
            # a*b is exact in float64 (24+24 bits); one float64 rounding then one float32 rounding.
            return (a.double() * b.double() + c.double()).float()
        return a * b + c

    def dot3(self, a0, b0, a1, b1, a2, b2):
        """nvcc's contraction of a0*b0 + a1*b1 + a2*b2:  fma(a2,b2, fma(a0,b0, a1*b1))."""
        return self.fma(a2, b2, self.fma(a0, b0, a1 * b1))


class _ClampMaxST(torch.autograd.Function):
    """min(x, hi) with a straight-through gradient, as the reference backward treats it (backward.cu:560-562)."""

    @staticmethod
    def forward(ctx, x, hi):
        return torch.clamp(x, max=hi)

    @staticmethod
    def backward(ctx, g):
        return g, None


def _flat16(m: torch.Tensor, dt) -> torch.Tensor:
    return m.detach().reshape(-1).to(dt).cpu()


# ----------------------------------------------------------------------------------------------------------------
# (a) preprocess
# ----------------------------------------------------------------------------------------------------------------
def compute_cov3d(A: _Arith, scales, rotations, mod):
    """forward.cu:129-163.  rotations are (r,x,y,z) and are NOT normalised here (line 138)."""
    sx = mod * scales[:, 0]
    sy = mod * scales[:, 1]
    sz = mod * scales[:, 2]
    r, x, y, z = rotations[:, 0], rotations[:, 1], rotations[:, 2], rotations[:, 3]
    # Besides the FMAs nvcc's front end forms, ptxas fuses a product into a following add/sub whenever all uses of
    # the product are such adds/subs (read off the reference SASS): x*y, r*y and y*z are never rounded on their own.
    yy = y * y
    zz = z * z
    rz = r * z
    xz = x * z
    rx = r * x
    xx_zz = A.fma(x, x, zz)
    xx_yy = A.fma(x, x, yy)
    yy_zz = yy + zz
    one = A.c(1.0)

    def twice(a):
        return a + a
    # R columns (GLM column-major constructor, forward.cu:145-149)
    R00 = one - twice(yy_zz)
    R01 = twice(A.fma(x, y, -rz))
    R02 = twice(A.fma(r, y, xz))
    R10 = twice(A.fma(x, y, rz))
    R11 = one - twice(xx_zz)
    R12 = twice(A.fma(y, z, -rx))
    R20 = twice(A.fma(-r, y, xz))
    R21 = twice(A.fma(y, z, rx))
    R22 = one - twice(xx_yy)
    # M = S * R  (column j of M = S * column j of R)
    M00, M01, M02 = sx * R00, sy * R01, sz * R02
    M10, M11, M12 = sx * R10, sy * R11, sz * R12
    M20, M21, M22 = sx * R20, sy * R21, sz * R22
    # Sigma = M^T M, upper triangle
    c0 = A.dot3(M00, M00, M01, M01, M02, M02)
    c1 = A.dot3(M10, M00, M11, M01, M12, M02)
    c2 = A.dot3(M20, M00, M21, M01, M22, M02)
    c3 = A.dot3(M10, M10, M11, M11, M12, M12)
    c4 = A.dot3(M20, M10, M21, M11, M22, M12)
    c5 = A.dot3(M20, M20, M21, M21, M22, M22)
    return torch.stack([c0, c1, c2, c3, c4, c5], dim=1)


def _transform4(A: _Arith, p, m, row, with_w=True):
    """matrix[row]*x + matrix[row+4]*y + matrix[row+8]*z + matrix[row+12]  (auxiliary.h:58-77), nvcc's contraction."""
    x, y, z = p[:, 0], p[:, 1], p[:, 2]
    acc = A.fma(z, m[row + 8], A.fma(x, m[row], y * m[row + 4]))
    return m[row + 12] + acc


def eval_sh(A: _Arith, deg, means3D, campos, shs):
    """forward.cu:20-71.  shs is [P, M, 3]."""
    d = means3D - campos[None, :]
    dx, dy, dz = d[:, 0], d[:, 1], d[:, 2]
    ln = torch.sqrt(A.fma(dz, dz, A.fma(dx, dx, dy * dy)))
    x, y, z = dx / ln, dy / ln, dz / ln
    sh = shs
    res = A.c(SH_C0) * sh[:, 0]
    if deg > 0:
        t = (y * A.c(SH_C1))[:, None]
        res = A.fma(-t, sh[:, 1], res)
        t = (z * A.c(SH_C1))[:, None]
        res = A.fma(t, sh[:, 2], res)
        t = (x * A.c(SH_C1))[:, None]
        res = A.fma(-t, sh[:, 3], res)
        if deg > 1:
            xx, yy, zz = x * x, y * y, z * z
            xy, yz, xz = x * y, y * z, x * z
            res = A.fma((xy * A.c(SH_C2[0]))[:, None], sh[:, 4], res)
            res = A.fma((yz * A.c(SH_C2[1]))[:, None], sh[:, 5], res)
            zz2 = zz + zz
            res = A.fma((((zz2 - xx) - yy) * A.c(SH_C2[2]))[:, None], sh[:, 6], res)
            res = A.fma((xz * A.c(SH_C2[3]))[:, None], sh[:, 7], res)
            xx_yy = xx - yy
            res = A.fma((xx_yy * A.c(SH_C2[4]))[:, None], sh[:, 8], res)
            if deg > 2:
                three, four = A.c(3.0), A.c(4.0)
                res = A.fma(((y * A.c(SH_C3[0])) * A.fma(xx, three, -yy))[:, None], sh[:, 9], res)
                res = A.fma((z * (xy * A.c(SH_C3[1])))[:, None], sh[:, 10], res)
                f401 = A.fma(zz, four, -xx) - yy
                res = A.fma(((y * A.c(SH_C3[2])) * f401)[:, None], sh[:, 11], res)
                res = A.fma(((z * A.c(SH_C3[3])) * A.fma(-yy, three, A.fma(-xx, three, zz2)))[:, None], sh[:, 12], res)
                res = A.fma(((x * A.c(SH_C3[4])) * f401)[:, None], sh[:, 13], res)
                res = A.fma(((z * A.c(SH_C3[5])) * xx_yy)[:, None], sh[:, 14], res)
                res = A.fma(((x * A.c(SH_C3[6])) * A.fma(-yy, three, xx))[:, None], sh[:, 15], res)
    res = res + 0.5
    clamped = res < 0
    return torch.where(clamped, torch.zeros_like(res), res), clamped


def preprocess(cam: Camera, means3D, opacities, scales=None, rotations=None, cov3D_precomp=None,
               shs=None, dtype=torch.float32):
    """forward.cu:166-269.  Returns a dict; float outputs carry autograd history, integer ones are plain."""
    A = _Arith(dtype)
    dt = dtype
    P = means3D.shape[0]
    W, H = int(cam.image_width), int(cam.image_height)
    V = _flat16(cam.viewmatrix, dt)
    Pm = _flat16(cam.projmatrix, dt)
    # host-side float arithmetic of rasterizer_impl.cu:227-228 (C++ float)
    tanx = A.c(cam.tanfovx)
    tany = A.c(cam.tanfovy)
    focal_x = A.c(float(W)) / (A.c(2.0) * tanx)
    focal_y = A.c(float(H)) / (A.c(2.0) * tany)
    c_x = A.c(cam.c_x)
    c_y = A.c(cam.c_y)
    grid_x = (W + BLOCK_X - 1) // BLOCK_X
    grid_y = (H + BLOCK_Y - 1) // BLOCK_Y

    p = means3D.to(dt)
    depth = _transform4(A, p, V, 2)                      # p_view.z
    in_front = depth.detach() > 0                        # auxiliary.h:158  (culled iff z <= 0)

    hx = _transform4(A, p, Pm, 0)
    hy = _transform4(A, p, Pm, 1)
    hw = _transform4(A, p, Pm, 3)
    pw = torch.reciprocal(hw + A.c(0.0000001))
    ndc_x = hx * pw
    ndc_y = hy * pw

    if cov3D_precomp is not None:
        cov3D = cov3D_precomp.to(dt)
    else:
        cov3D = compute_cov3d(A, scales.to(dt), rotations.to(dt), A.c(cam.scale_modifier))
    c0, c1, c2, c3, c4, c5 = [cov3D[:, i] for i in range(6)]

    # ---- computeCov2D (forward.cu:75-124) ----
    tx = _transform4(A, p, V, 0)
    ty = _transform4(A, p, V, 1)
    tz = depth
    # guard culled points so the discarded branch cannot put NaNs into autograd
    tz_s = torch.where(in_front, tz, torch.ones_like(tz))
    lim_x_pos = A.fma(tanx, A.c(0.3), (A.c(float(W)) - c_x) / focal_x)
    lim_x_neg = A.fma(tanx, A.c(0.3), c_x / focal_x)
    lim_y_pos = A.fma(tany, A.c(0.3), (A.c(float(H)) - c_y) / focal_y)
    lim_y_neg = A.fma(tany, A.c(0.3), c_y / focal_y)
    txtz = tx / tz_s
    tytz = ty / tz_s
    cxl = torch.minimum(lim_x_pos, torch.maximum(-lim_x_neg, txtz))
    cyl = torch.minimum(lim_y_pos, torch.maximum(-lim_y_neg, tytz))
    J00 = focal_x / tz_s
    tzz = tz_s * tz_s
    J02 = (focal_x * (cxl * (-tz_s))) / tzz
    J11 = focal_y / tz_s
    J12 = (focal_y * (cyl * (-tz_s))) / tzz
    # T = W * J ; only columns 0 and 1 are non-zero
    T00 = A.fma(V[2], J02, V[0] * J00)
    T01 = A.fma(V[6], J02, V[4] * J00)
    T02 = A.fma(V[10], J02, V[8] * J00)
    T10 = A.fma(V[2], J12, J11 * V[1])
    T11 = A.fma(V[6], J12, J11 * V[5])
    T12 = A.fma(V[10], J12, J11 * V[9])
    # Vrk * T columns
    u0 = A.dot3(T00, c0, T01, c1, T02, c2)
    u1 = A.dot3(T00, c1, T01, c3, T02, c4)
    u2 = A.dot3(T00, c2, T01, c4, T02, c5)
    w0 = A.dot3(T10, c0, T11, c1, T12, c2)
    w1 = A.dot3(T10, c1, T11, c3, T12, c4)
    w2 = A.dot3(T10, c2, T11, c4, T12, c5)
    cov_a = A.dot3(T00, u0, T01, u1, T02, u2) + A.c(0.3)
    cov_b = A.dot3(T00, w0, T01, w1, T02, w2)
    cov_c = A.dot3(T10, w0, T11, w1, T12, w2) + A.c(0.3)

    det = A.fma(cov_a, cov_c, -(cov_b * cov_b))
    det_ok = det.detach() != 0
    det_s = torch.where(det_ok, det, torch.ones_like(det))
    det_inv = torch.reciprocal(det_s)
    conic_x = cov_c * det_inv
    conic_y = det_inv * (-cov_b)
    conic_z = cov_a * det_inv

    mid = (cov_a + cov_c) * 0.5
    disc = torch.clamp(A.fma(mid, mid, -det), min=A.c(0.1).item())
    sq = torch.sqrt(disc)
    lam = torch.maximum(mid + sq, mid - sq)
    rad_f = torch.ceil(torch.sqrt(torch.clamp(lam, min=0.0)) * 3.0).detach()
    rad_f = torch.nan_to_num(rad_f, nan=0.0, posinf=2147483520.0)
    radius_i = rad_f.clamp(max=2147483520.0).to(torch.int64)       # cvt.rzi.s32.f32 (saturating)
    rf = radius_i.to(dt)

    # ndc2Pix in double, then rounded to float (auxiliary.h:41-44)
    pix_x = (((ndc_x.double() + 1.0) * float(W) - 1.0) * 0.5).to(dt)
    pix_y = (((ndc_y.double() + 1.0) * float(H) - 1.0) * 0.5).to(dt)

    def _to_int(v):
        v = torch.nan_to_num(v.detach(), nan=0.0)
        return torch.trunc(v).clamp(-2147483648.0, 2147483520.0).to(torch.int64)

    rmin_x = _to_int((pix_x - rf) * 0.0625).clamp(min=0).clamp(max=grid_x)
    rmin_y = _to_int((pix_y - rf) * 0.0625).clamp(min=0).clamp(max=grid_y)
    rmax_x = _to_int((((pix_x + rf) + 16.0) + (-1.0)) * 0.0625).clamp(min=0).clamp(max=grid_x)
    rmax_y = _to_int((((pix_y + rf) + 16.0) + (-1.0)) * 0.0625).clamp(min=0).clamp(max=grid_y)
    tiles = (rmax_x - rmin_x) * (rmax_y - rmin_y)

    visible = in_front & det_ok & (tiles != 0)
    radii = torch.where(visible, radius_i, torch.zeros_like(radius_i)).to(torch.int32)
    tiles_touched = torch.where(visible, tiles, torch.zeros_like(tiles)).to(torch.int64)

    out = dict(
        P=P, W=W, H=H, grid_x=grid_x, grid_y=grid_y,
        depths=depth, means2D=torch.stack([pix_x, pix_y], dim=1), cov3D=cov3D,
        conic=torch.stack([conic_x, conic_y, conic_z], dim=1), opacity=opacities.reshape(-1).to(dt),
        radii=radii, tiles_touched=tiles_touched, visible=visible,
        rect_min=torch.stack([rmin_x, rmin_y], dim=1), rect_max=torch.stack([rmax_x, rmax_y], dim=1),
        cov2D=torch.stack([cov_a, cov_b, cov_c], dim=1).detach(),
        # distance (in output units) of knife-edge quantities to their decision boundary, for test triage
        edge_radius=(torch.sqrt(torch.clamp(lam, min=0.0)) * 3.0).detach(),
    )
    if shs is not None:
        rgb, clamped = eval_sh(A, int(cam.sh_degree), p, cam.campos.to(dt).cpu(), shs.to(dt))
        out["rgb"] = rgb
        out["clamped"] = clamped
    return out


# ----------------------------------------------------------------------------------------------------------------
# (b) binning: key emission, stable sort, tile ranges   (integer work, numpy)
# ----------------------------------------------------------------------------------------------------------------
def build_sorted_lists(pre: dict):
    """rasterizer_impl.cu:70-138,283-323.  Returns keys_sorted (u64), point_list (i64), ranges (T,2) and offsets."""
    grid_x, grid_y = pre["grid_x"], pre["grid_y"]
    tiles = pre["tiles_touched"].numpy().astype(np.int64)
    offsets = np.cumsum(tiles)                               # InclusiveSum
    R = int(offsets[-1]) if len(offsets) else 0
    vis = np.nonzero(tiles > 0)[0]
    rmin = pre["rect_min"].numpy()
    rmax = pre["rect_max"].numpy()
    depth_bits = pre["depths"].detach().to(torch.float32).numpy().view(np.uint32).astype(np.uint64)
    rep = np.repeat(vis, tiles[vis])                          # gaussian id per instance, ascending
    start = np.repeat(offsets[vis] - tiles[vis], tiles[vis])
    k = np.arange(R, dtype=np.int64) - start                   # instance number inside its gaussian
    w = (rmax[rep, 0] - rmin[rep, 0]).astype(np.int64)
    ty = rmin[rep, 1] + k // np.maximum(w, 1)                  # y outer, x inner
    tx = rmin[rep, 0] + k % np.maximum(w, 1)
    tile_id = (ty * grid_x + tx).astype(np.uint64)
    keys = (tile_id << np.uint64(32)) | depth_bits[rep]
    bits = 32 + get_higher_msb(grid_x * grid_y)
    mask = np.uint64((1 << bits) - 1)
    order = np.argsort(keys & mask, kind="stable")            # CUB LSD radix sort on [0, bits) is stable
    keys_sorted = keys[order]
    point_list = rep[order].astype(np.int64)
    n_tiles = grid_x * grid_y
    ranges = np.zeros((n_tiles, 2), dtype=np.int64)
    if R > 0:
        t_sorted = (keys_sorted >> np.uint64(32)).astype(np.int64)
        first = np.searchsorted(t_sorted, np.arange(n_tiles), side="left")
        last = np.searchsorted(t_sorted, np.arange(n_tiles), side="right")
        nz = last > first
        ranges[nz, 0] = first[nz]
        ranges[nz, 1] = last[nz]
    return dict(keys_unsorted=keys, values_unsorted=rep.astype(np.int64), keys_sorted=keys_sorted,
                point_list=point_list, ranges=ranges, offsets=offsets, num_rendered=R, sort_bits=bits)


# ----------------------------------------------------------------------------------------------------------------
# (c) tile blend forward (differentiable)
# ----------------------------------------------------------------------------------------------------------------
def blend_forward(pre: dict, lists: dict, colors, bg, dtype=torch.float32, collect_fragile=False):
    """forward.cu:274-408.  `colors` is [P, C] (C = 3, or 6 when the fg/bg segmentation colours ride along).

    Returns dict(color [C,H,W], depth [1,H,W], alpha [1,H,W], n_contrib [H,W] int, final_T [H,W]).
    alpha is the accumulated weight sum(alpha_i * T_i) the reference computes but forgets to store.
    """
    A = _Arith(dtype)
    dt = dtype
    W, H, gx, gy = pre["W"], pre["H"], pre["grid_x"], pre["grid_y"]
    nT = gx * gy
    C = colors.shape[1]
    ranges = lists["ranges"]
    plist = torch.from_numpy(lists["point_list"])
    lens = torch.from_numpy(ranges[:, 1] - ranges[:, 0])
    Lmax = int(lens.max()) if nT > 0 else 0
    starts = torch.from_numpy(ranges[:, 0])

    tid = torch.arange(nT)
    tile_x = (tid % gx) * BLOCK_X
    tile_y = (tid // gx) * BLOCK_Y
    lx = torch.arange(BLOCK_SIZE) % BLOCK_X
    ly = torch.arange(BLOCK_SIZE) // BLOCK_X
    px = tile_x[:, None] + lx[None, :]                # [nT,256] int
    py = tile_y[:, None] + ly[None, :]
    inside = (px < W) & (py < H)
    pxf = px.to(dt)
    pyf = py.to(dt)

    means2D = pre["means2D"].to(dt)
    conic = pre["conic"].to(dt)
    opac = pre["opacity"].to(dt)
    depths = pre["depths"].to(dt)
    colors = colors.to(dt)

    Tr = torch.ones(nT, BLOCK_SIZE, dtype=dt)
    done = ~inside
    contributor = torch.zeros(nT, BLOCK_SIZE, dtype=torch.int64)
    last = torch.zeros(nT, BLOCK_SIZE, dtype=torch.int64)
    Cacc = torch.zeros(nT, BLOCK_SIZE, C, dtype=dt)
    Dacc = torch.zeros(nT, BLOCK_SIZE, dtype=dt)
    Wacc = torch.zeros(nT, BLOCK_SIZE, dtype=dt)
    fragile = torch.zeros(nT, BLOCK_SIZE, dtype=torch.bool)
    pairs = 0
    thr_alpha = A.c(1.0) / A.c(255.0)
    thr_T = A.c(0.0001)

    for j in range(Lmax):
        active = (j < lens)
        valid = active[:, None] & (~done)
        if not bool(valid.any()):
            break
        pos = torch.where(active, starts + j, torch.zeros_like(starts))
        g = plist[pos] if plist.numel() else pos
        xy = means2D[g]
        con = conic[g]
        o = opac[g]
        dx = xy[:, 0:1] - pxf
        dy = xy[:, 1:2] - pyf
        t3 = dy * (dy * con[:, 2:3])
        q = A.fma(dx, dx * con[:, 0:1], t3)
        power = A.fma(q, A.c(-0.5), -(dy * (dx * con[:, 1:2])))      # ptxas fuses the *(-0.5) into the subtraction
        skip1 = power.detach() > 0
        power_s = torch.where(skip1 | ~valid, torch.zeros_like(power), power)
        G = torch.exp(power_s)
        alpha = _ClampMaxST.apply(o[:, None] * G, 0.99)
        skip2 = alpha.detach() < thr_alpha
        test_T = Tr * (1.0 - alpha)
        term = test_T.detach() < thr_T
        contributor = contributor + valid.to(torch.int64)
        pairs += int(valid.sum())
        live = valid & ~skip1 & ~skip2
        contributes = live & ~term
        if collect_fragile:
            ad = alpha.detach()
            near = (power.detach().abs() < 1e-6) | ((ad - thr_alpha).abs() < 4e-7) | \
                   ((test_T.detach() - thr_T).abs() < 2e-9)
            fragile |= valid & near
        w_c = (alpha[:, :, None] * colors[g][:, None, :])
        Cacc = torch.where(contributes[:, :, None], A.fma(Tr[:, :, None], w_c, Cacc), Cacc)
        Dacc = torch.where(contributes, A.fma(Tr, alpha * depths[g][:, None], Dacc), Dacc)
        Wacc = torch.where(contributes, A.fma(alpha, Tr, Wacc), Wacc)
        Tr = torch.where(contributes, test_T, Tr)
        last = torch.where(contributes, contributor, last)
        done = done | (live & term)

    bgv = bg.to(dt).reshape(-1)
    if bgv.numel() < C:
        bgv = torch.cat([bgv, torch.zeros(C - bgv.numel(), dtype=dt)])
    col_t = A.fma(Tr[:, :, None], bgv[None, None, :C], Cacc)

    def to_img(t):   # [nT,256,...] -> [H,W,...]
        full = t.reshape(gy, gx, BLOCK_Y, BLOCK_X, *t.shape[2:])
        full = full.permute(0, 2, 1, 3, *range(4, full.dim())).reshape(gy * BLOCK_Y, gx * BLOCK_X, *t.shape[2:])
        return full[:H, :W]

    out = dict(
        color=to_img(col_t).permute(2, 0, 1),
        depth=to_img(Dacc)[None],
        alpha=to_img(Wacc)[None],
        n_contrib=to_img(last),
        final_T=to_img(Tr),
        pairs=pairs,
    )
    if collect_fragile:
        out["fragile"] = to_img(fragile)
    return out


# ----------------------------------------------------------------------------------------------------------------
# whole op
# ----------------------------------------------------------------------------------------------------------------
def rasterize(cam: Camera, means3D, opacities, colors_precomp=None, scales=None, rotations=None,
              cov3D_precomp=None, shs=None, extra_colors=None, dtype=torch.float32, collect_fragile=False):
    """One view, forward.  `extra_colors` ([P,3], e.g. the fg/bg segmentation colours of REF/train.py:247-249)
    are blended in the same pass with the same geometry; the reference renders them with a second call.
    """
    pre = preprocess(cam, means3D, opacities, scales, rotations, cov3D_precomp, shs, dtype=dtype)
    pre["means2D"].retain_grad() if pre["means2D"].requires_grad else None
    lists = build_sorted_lists(pre)
    cols = pre["rgb"] if shs is not None else colors_precomp.to(dtype)
    if extra_colors is not None:
        cols = torch.cat([cols, extra_colors.to(dtype)], dim=1)
    img = blend_forward(pre, lists, cols, cam.bg, dtype=dtype, collect_fragile=collect_fragile)
    return dict(pre=pre, lists=lists, **img)


def means2D_grad_ndc(pre: dict) -> Optional[torch.Tensor]:
    """dL/dmeans2D in the reference's units (backward.cu:520-521,640-641): pixel gradient times (W/2, H/2); z = 0."""
    g = pre["means2D"].grad
    if g is None:
        return None
    s = torch.tensor([0.5 * pre["W"], 0.5 * pre["H"]], dtype=g.dtype)
    out = torch.zeros(g.shape[0], 3, dtype=g.dtype)
    out[:, :2] = g * s
    return out


def camera_from_spec(w, h, k, w2c, near=0.01, far=100.0, bg=(0.0, 0.0, 0.0)) -> Camera:
    """REF/helpers.py:68-95 (setup_camera) restated on the CPU."""
    fx, fy, cx, cy = k[0][0], k[1][1], k[0][2], k[1][2]
    w2c_t = torch.tensor(np.asarray(w2c), dtype=torch.float32)
    cam_center = torch.inverse(w2c_t)[:3, 3]
    w2c_T = w2c_t.unsqueeze(0).transpose(1, 2)
    opengl_proj = torch.tensor([[2 * fx / w, 0.0, -(w - 2 * cx) / w, 0.0],
                                [0.0, 2 * fy / h, -(h - 2 * cy) / h, 0.0],
                                [0.0, 0.0, far / (far - near), -(far * near) / (far - near)],
                                [0.0, 0.0, 1.0, 0.0]], dtype=torch.float32).unsqueeze(0).transpose(1, 2)
    full_proj = w2c_T.bmm(opengl_proj)
    return Camera(image_height=h, image_width=w, tanfovx=w / (2 * fx), tanfovy=h / (2 * fy), c_x=cx, c_y=cy,
                  bg=torch.tensor(bg, dtype=torch.float32), scale_modifier=1.0,
                  viewmatrix=w2c_T.reshape(4, 4).contiguous(), projmatrix=full_proj.reshape(4, 4).contiguous(),
                  campos=cam_center, sh_degree=0)
