"""Host side of the B200 rasterizer: the reference's operator API re-hosted on the C ABI.

Mirrors R/diff_gaussian_rasterization/__init__.py (R/ = submodules_fsgs/diff-gaussian-rasterization-confidence/):
  GaussianRasterizationSettings   :176-191   (same field names and order; c_x, c_y, confidence may be omitted)
  GaussianRasterizer              :194-245   (same kwargs, same two "exactly one of" errors)
  _RasterizeGaussians             :48-174    (same argument order, same gradient order)
plus a batched entry point, `rasterize_views`, that renders all V cameras of a timestep in one set of launches and
blends the fg/bg segmentation colours in the same pass as the colours (the reference needs two calls per view,
REF/train.py:142,249).

PyTorch is used for device memory, streams and autograd plumbing only; every kernel lives in libgsr_b200.so.
CPU tensors are rejected: there is no CPU path.
"""
from __future__ import annotations

import ctypes as C
from typing import List, NamedTuple, Optional, Sequence

import torch
import torch.nn as nn

from . import _lib

NUM_SEMANTIC_CHANNELS = 32          # R/cuda_rasterizer/config.h:17


class GaussianRasterizationSettings(NamedTuple):
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    c_x: Optional[float] = None          # principal point; None -> image centre (Inria-style callers omit it)
    c_y: Optional[float] = None
    bg: Optional[torch.Tensor] = None
    scale_modifier: float = 1.0
    viewmatrix: Optional[torch.Tensor] = None
    projmatrix: Optional[torch.Tensor] = None
    sh_degree: int = 0
    campos: Optional[torch.Tensor] = None
    prefiltered: bool = False
    debug: bool = False
    confidence: Optional[torch.Tensor] = None   # dead field in the reference; accepted and ignored


def _host_floats(t, n: int, what: str) -> List[float]:
    if t is None:
        return [0.0] * n
    if torch.is_tensor(t):
        v = t.detach().reshape(-1).to(torch.float32).cpu().tolist()
    else:
        v = [float(x) for x in t]
    if len(v) < n:
        raise ValueError(f"{what} must have at least {n} elements, got {len(v)}")
    return v[:n]


_CAM_CACHE: dict = {}


def _versions(s: GaussianRasterizationSettings, bg_extra):
    """(data_ptr, in-place version) of every tensor a GsrCamera is built from: an in-place update of a pose / background
    tensor (camera optimisation, random backgrounds) must be seen, as the reference reads the tensors on every call."""
    return tuple((t.data_ptr(), t._version) if torch.is_tensor(t) else t
                 for t in (s.viewmatrix, s.projmatrix, s.campos, s.bg, bg_extra))


def camera_struct(s: GaussianRasterizationSettings, bg_extra=None, band=None) -> _lib.GsrCamera:
    """GsrCamera for one settings tuple.  The matrices live on the device; reading them back costs a sync, so the struct
    is cached per settings object and rebuilt when any of its tensors (or `bg_extra`) was replaced or modified in place.
    `band` = (first tile row, one past the last) restricts the render to a band of 16-pixel rows (None = all)."""
    band = None if band is None else (int(band[0]), int(band[1]))
    key = (id(s), id(bg_extra), band)
    hit = _CAM_CACHE.get(key)
    if hit is not None and hit[0] is s and hit[2] is bg_extra and hit[3] == _versions(s, bg_extra):
        return hit[1]
    cam = _lib.GsrCamera()
    if band is not None:
        if not (0 <= band[0] < band[1]):
            raise ValueError(f"band must be a non-empty range of tile rows, got {band}")
        cam.tile_row_begin, cam.tile_row_end = band
    cam.image_width = int(s.image_width)
    cam.image_height = int(s.image_height)
    cam.tanfovx = float(s.tanfovx)
    cam.tanfovy = float(s.tanfovy)
    cam.c_x = float(s.c_x) if s.c_x is not None else 0.5 * int(s.image_width)
    cam.c_y = float(s.c_y) if s.c_y is not None else 0.5 * int(s.image_height)
    cam.scale_modifier = float(s.scale_modifier)
    cam.sh_degree = int(s.sh_degree)
    cam.prefiltered = int(bool(s.prefiltered))
    cam.debug = int(bool(s.debug))
    bg = _host_floats(s.bg, 3, "bg") + _host_floats(bg_extra, 3, "bg_extra") + [0.0, 0.0]
    for i in range(8):
        cam.bg[i] = bg[i]
    vm = _host_floats(s.viewmatrix, 16, "viewmatrix")
    pm = _host_floats(s.projmatrix, 16, "projmatrix")
    cp = _host_floats(s.campos, 3, "campos")
    for i in range(16):
        cam.viewmatrix[i] = vm[i]
        cam.projmatrix[i] = pm[i]
    for i in range(3):
        cam.campos[i] = cp[i]
    if len(_CAM_CACHE) > 4096:
        _CAM_CACHE.clear()
    _CAM_CACHE[key] = (s, cam, bg_extra, _versions(s, bg_extra))      # holds s and bg_extra alive: ids stay unique
    return cam


def _cuda_f32(t: Optional[torch.Tensor], name: str, shape_tail=None) -> Optional[torch.Tensor]:
    if t is None or (torch.is_tensor(t) and t.numel() == 0):
        return None
    if not torch.is_tensor(t):
        raise ValueError(f"{name} must be a tensor")
    if not t.is_cuda:
        raise ValueError(f"{name} must be a CUDA tensor (this rasterizer has no CPU path)")
    if t.dtype != torch.float32:
        raise ValueError(f"{name} must be float32, got {t.dtype}")
    if shape_tail is not None and tuple(t.shape[1:]) != tuple(shape_tail):
        raise ValueError(f"{name} must have shape [P, {', '.join(map(str, shape_tail))}], got {tuple(t.shape)}")
    return t.contiguous()


def _stream_ptr(device) -> int:
    return torch.cuda.current_stream(device).cuda_stream


class RasterState:
    """Everything the backward needs from the forward (the reference keeps geomBuffer/binningBuffer/imgBuffer on ctx)."""
    __slots__ = ("V", "P", "W", "H", "channels", "sh_coeffs", "capacity", "num_rendered", "geom", "binning", "image",
                 "radii", "layout", "feature_map", "flags")


def workspace_layout(P, V, W, H, channels, sh_coeffs, capacity) -> _lib.GsrLayout:
    L = _lib.GsrLayout()
    _lib.check(_lib.load().gsr_workspace_layout(P, V, W, H, channels, sh_coeffs, capacity, C.byref(L)), "workspace_layout")
    return L


def forward_views(settings: Sequence[GaussianRasterizationSettings], means3D, opacities, colors_precomp=None,
                  scales=None, rotations=None, cov3D_precomp=None, shs=None, extra_colors=None, bg_extra=None,
                  capacity: Optional[int] = None, semantic_feature=None, bands=None):
    """Runs the forward for V views.  Returns (color [V,3,H,W], extra [V,3,H,W] | None, depth [V,1,H,W],
    alpha [V,1,H,W], radii [V,P] int32, RasterState).  With `semantic_feature` ([P,32] or the reference's [P,1,32],
    R/diff_gaussian_rasterization/__init__.py:215) the 32-channel feature map [V,32,H,W] is left in
    RasterState.feature_map.  `bands[v]` = (first, last + 1) tile rows to render for view v, or None for the whole
    view: the rest of that image is background and takes no part in the backward (dist.partition_units shards single
    views over GPUs this way).  With capacity=None the instance count is read back from the
    device once (the reference's one host sync, rasterizer_impl.cu:287); otherwise no sync happens and the caller
    checks `read_counters` later.
    """
    lib = _lib.load()
    V = len(settings)
    if V == 0:
        raise ValueError("need at least one view")
    if not torch.is_tensor(means3D) or means3D.dim() != 2 or means3D.shape[1] != 3:
        raise ValueError("means3D must have dimensions (num_points, 3)")       # R/rasterize_points.cu:60-62
    if not means3D.is_cuda or means3D.dtype != torch.float32:
        raise ValueError("means3D must be a float32 CUDA tensor (this rasterizer has no CPU path)")
    means3D = means3D.contiguous()
    dev = means3D.device
    P = means3D.shape[0]
    W, H = int(settings[0].image_width), int(settings[0].image_height)
    opacities = _cuda_f32(opacities, "opacities")
    colors_precomp = _cuda_f32(colors_precomp, "colors_precomp", (3,))
    extra_colors = _cuda_f32(extra_colors, "extra_colors", (3,))
    scales = _cuda_f32(scales, "scales", (3,))
    rotations = _cuda_f32(rotations, "rotations", (4,))
    cov3D_precomp = _cuda_f32(cov3D_precomp, "cov3D_precomp", (6,))
    shs = _cuda_f32(shs, "shs")
    sh_coeffs = 0
    if shs is not None:
        if shs.dim() != 3 or shs.shape[2] != 3:
            raise ValueError("shs must have shape [P, M, 3]")
        sh_coeffs = int(shs.shape[1])
    if P > 0 and opacities is None:
        raise ValueError("opacities are required")
    channels = 6 if extra_colors is not None else 3
    semantic_feature = _cuda_f32(semantic_feature, "semantic_feature")
    if semantic_feature is not None:
        if semantic_feature.numel() != P * NUM_SEMANTIC_CHANNELS:
            raise ValueError(f"semantic_feature must hold {NUM_SEMANTIC_CHANNELS} channels per Gaussian "
                             f"(R/cuda_rasterizer/config.h:17), got {tuple(semantic_feature.shape)}")
        semantic_feature = semantic_feature.reshape(P, NUM_SEMANTIC_CHANNELS)

    if bands is not None and len(bands) != V:
        raise ValueError("bands must have one entry per view")
    cams = (_lib.GsrCamera * V)(*[camera_struct(s, bg_extra, None if bands is None else bands[i])
                                  for i, s in enumerate(settings)])
    sizes = _lib.GsrWorkspaceSizes()
    _lib.check(lib.gsr_workspace_sizes(P, V, W, H, channels, sh_coeffs, 0, C.byref(sizes)), "workspace_sizes")
    geom = torch.empty(sizes.geom_bytes, dtype=torch.uint8, device=dev)
    image = torch.empty(sizes.image_bytes, dtype=torch.uint8, device=dev)
    radii = torch.empty((V, P), dtype=torch.int32, device=dev)
    stream = _stream_ptr(dev)
    flags = _lib.FLAG_DEBUG if any(bool(s.debug) for s in settings) else 0
    if flags:
        cams[0].debug = 1
    with torch.cuda.device(dev):
        _lib.check(lib.gsr_forward_geometry(cams, V, P, channels, sh_coeffs, _lib.ptr(means3D), _lib.ptr(opacities),
                                            _lib.ptr(scales), _lib.ptr(rotations), _lib.ptr(cov3D_precomp),
                                            _lib.ptr(shs), geom.data_ptr(), sizes.geom_bytes, radii.data_ptr(), None,
                                            stream), "forward_geometry")
        num_rendered = None
        if capacity is None:
            num_rendered = 0
            if P > 0:
                c = geom[:12].view(torch.int32).tolist()                              # the one host sync
                num_rendered = c[0]
                if c[2]:
                    # R/cuda_rasterizer/auxiliary.h:160-167: with prefiltered = True a culled point is an error
                    raise RuntimeError("Point is filtered although prefiltered is set. This shouldn't happen!")
            capacity = num_rendered
        _lib.check(lib.gsr_workspace_sizes(P, V, W, H, channels, sh_coeffs, capacity, C.byref(sizes)), "workspace_sizes")
        binning = torch.empty(max(sizes.binning_bytes, 16), dtype=torch.uint8, device=dev)
        color = torch.empty((V, 3, H, W), dtype=torch.float32, device=dev)
        extra = torch.empty((V, 3, H, W), dtype=torch.float32, device=dev) if channels == 6 else None
        depth = torch.empty((V, 1, H, W), dtype=torch.float32, device=dev)
        alpha = torch.empty((V, 1, H, W), dtype=torch.float32, device=dev)
        feature_map = None
        if semantic_feature is not None:
            feature_map = torch.empty((V, NUM_SEMANTIC_CHANNELS, H, W), dtype=torch.float32, device=dev)
        _lib.check(lib.gsr_forward_render(V, P, W, H, channels, sh_coeffs, capacity, _lib.ptr(colors_precomp),
                                          _lib.ptr(extra_colors), _lib.ptr(semantic_feature), NUM_SEMANTIC_CHANNELS,
                                          geom.data_ptr(), binning.data_ptr(), image.data_ptr(), color.data_ptr(),
                                          _lib.ptr(extra), depth.data_ptr(), alpha.data_ptr(), _lib.ptr(feature_map),
                                          flags, stream), "forward_render")
    st = RasterState()
    st.V, st.P, st.W, st.H, st.channels, st.sh_coeffs = V, P, W, H, channels, sh_coeffs
    st.capacity, st.num_rendered = int(capacity), num_rendered
    st.geom, st.binning, st.image, st.radii = geom, binning, image, radii
    st.layout = None
    st.feature_map = feature_map
    st.flags = flags
    return color, extra, depth, alpha, radii, st


def read_counters(st: RasterState):
    """(num_rendered, overflowed) -- synchronises the current stream."""
    c = st.geom[:8].view(torch.int32).cpu()
    return int(c[0]), bool(c[1])


def read_flags(st: RasterState) -> dict:
    """All device-side counters / flags of a forward (gsr_read_counters) -- synchronises the current stream.  With an
    explicit `capacity` the forward never looks at them itself: check `overflowed` (instances were dropped) and
    `prefilter_violation` (a camera said prefiltered=True and a Gaussian was culled anyway; the reference traps there,
    R/cuda_rasterizer/auxiliary.h:160-167) afterwards."""
    c = st.geom[:16].view(torch.int32).tolist()
    return dict(num_rendered=c[0], overflowed=bool(c[1]), prefilter_violation=bool(c[2]))


def tile_row_costs(st: RasterState) -> torch.Tensor:
    """[V, tile rows] float64: summed tile-list lengths of every row of 16-pixel tiles (the blend work of that row);
    what dist.partition_units() balances.  Reads the forward's tile ranges from the image workspace."""
    L = workspace_layout(st.P, st.V, st.W, st.H, st.channels, st.sh_coeffs, st.capacity)
    tx, ty = (st.W + 15) // 16, (st.H + 15) // 16
    r = st.image[L.ranges: L.ranges + st.V * tx * ty * 8].view(torch.int32).view(st.V, ty, tx, 2)
    return (r[..., 1] - r[..., 0]).double().sum(dim=2)


def backward_views(st: RasterState, means3D, opacities, colors_precomp, scales, rotations, cov3D_precomp, shs,
                   extra_colors, dL_dcolor, dL_dextra, dL_ddepth, dL_dalpha, label=None, want_means2D=True, out=None,
                   semantic_feature=None, dL_dfeature=None, want_means2D_color=False):
    """Returns dict of gradients summed over the views of the batch (means2D: [V,P,3]; features: [P,32] when
    `semantic_feature` is given, with `dL_dfeature` [V,32,H,W] the cotangent of the feature map; means2D_color [V,P,3]
    with want_means2D_color: the colour render's share of means2D -- everything but the fused extra colours' cotangent --
    which is what the reference's densification statistic is made of, REF/train.py:245).  `out` may carry preallocated
    contiguous float32 tensors for means3D / colors / opacity / scales / rotations (e.g. views into one all-reduce
    slab, dist.GradSlab); every gradient is written, never accumulated."""
    lib = _lib.load()
    dev = st.geom.device
    V, P = st.V, st.P
    f32 = dict(dtype=torch.float32, device=dev)

    def cot(t, shape):
        if t is None:
            return None
        t = t.to(torch.float32).contiguous()
        if tuple(t.shape) != shape:
            t = t.reshape(shape)
        return t

    H, W = st.H, st.W
    dL_dcolor = cot(dL_dcolor, (V, 3, H, W))
    dL_dextra = cot(dL_dextra, (V, 3, H, W)) if st.channels == 6 else None
    dL_ddepth = cot(dL_ddepth, (V, 1, H, W))
    dL_dalpha = cot(dL_dalpha, (V, 1, H, W))
    if semantic_feature is not None:
        semantic_feature = semantic_feature.reshape(P, NUM_SEMANTIC_CHANNELS).contiguous()
        dL_dfeature = cot(dL_dfeature, (V, NUM_SEMANTIC_CHANNELS, H, W))
    out = out or {}

    def buf(key, shape):
        t = out.get(key)
        if t is None:
            return torch.empty(shape, **f32)
        if tuple(t.shape) != tuple(shape) or t.dtype != torch.float32 or not t.is_contiguous() or t.device != dev:
            raise ValueError(f"out[{key!r}] must be a contiguous float32 {tuple(shape)} tensor on {dev}")
        return t

    g = dict(
        means3D=buf("means3D", (P, 3)),
        means2D=torch.empty((V, P, 3), **f32) if want_means2D else None,
        means2D_color=torch.empty((V, P, 3), **f32) if want_means2D_color else None,
        colors=buf("colors", (P, 3)) if colors_precomp is not None else None,
        extra=torch.empty((P, 3), **f32) if st.channels == 6 else None,
        opacity=buf("opacity", (P, 1)),
        scales=buf("scales", (P, 3)) if scales is not None else None,
        rotations=buf("rotations", (P, 4)) if rotations is not None else None,
        cov3D=torch.empty((P, 6), **f32) if cov3D_precomp is not None else None,
        sh=torch.empty((P, st.sh_coeffs, 3), **f32) if shs is not None else None,
        features=torch.empty((P, NUM_SEMANTIC_CHANNELS), **f32) if semantic_feature is not None else None,
    )
    if label is not None:
        label = label.to(torch.float32).reshape(-1).contiguous()
        if label.numel() != P:
            raise ValueError("label must have one entry per Gaussian")
    p = _lib.ptr
    with torch.cuda.device(dev):
        _lib.check(lib.gsr_backward(V, P, W, H, st.channels, st.sh_coeffs, st.capacity, p(means3D), p(opacities),
                                    p(scales), p(rotations), p(cov3D_precomp), p(shs), p(colors_precomp),
                                    p(extra_colors), p(semantic_feature), NUM_SEMANTIC_CHANNELS, p(st.radii),
                                    st.geom.data_ptr(), st.binning.data_ptr(), st.image.data_ptr(), p(dL_dcolor),
                                    p(dL_dextra), p(dL_ddepth), p(dL_dalpha), p(dL_dfeature), p(label),
                                    p(g["means3D"]), p(g["means2D"]), p(g["means2D_color"]), p(g["colors"]),
                                    p(g["extra"]), p(g["opacity"]), p(g["scales"]), p(g["rotations"]), p(g["cov3D"]),
                                    p(g["sh"]), p(g["features"]), getattr(st, "flags", 0), _stream_ptr(dev)), "backward")
    return g


# ----------------------------------------------------------------------------------------------------------------
# autograd
# ----------------------------------------------------------------------------------------------------------------
def _opt(t):
    return None if (t is None or (torch.is_tensor(t) and t.numel() == 0)) else t


class _RasterizeViews(torch.autograd.Function):
    """Batched op: V views, colour (+ optional extra colours), depth and alpha in one pass."""

    @staticmethod
    def forward(ctx, means3D, means2D, sh, colors_precomp, extra_colors, opacities, scales, rotations, cov3Ds_precomp,
                settings, label, bg_extra, capacity, state_out, bands=None, means2D_color=None):
        color, extra, depth, alpha, radii, st = forward_views(
            settings, means3D, opacities, _opt(colors_precomp), _opt(scales), _opt(rotations), _opt(cov3Ds_precomp),
            _opt(sh), _opt(extra_colors), bg_extra, capacity=capacity, bands=bands)
        ctx.state = st
        if state_out is not None:
            state_out.append(st)
        ctx.has = [_opt(t) is not None for t in (sh, colors_precomp, extra_colors, scales, rotations, cov3Ds_precomp)]
        ctx.means2D_shape = None if means2D is None else tuple(means2D.shape)
        ctx.means2D_color_shape = None if means2D_color is None else tuple(means2D_color.shape)
        ctx.label = label if torch.is_tensor(label) else None
        ctx.save_for_backward(means3D, sh, colors_precomp, extra_colors, opacities, scales, rotations, cov3Ds_precomp)
        ctx.mark_non_differentiable(radii)
        if extra is None:
            extra = color.new_empty(0)
        return color, extra, depth, alpha, radii

    @staticmethod
    def backward(ctx, g_color, g_extra, g_depth, g_alpha, _g_radii):
        means3D, sh, colors_precomp, extra_colors, opacities, scales, rotations, cov3Ds_precomp = ctx.saved_tensors
        st = ctx.state
        g = backward_views(st, means3D.contiguous(), opacities.contiguous(),
                           _cuda_f32(_opt(colors_precomp), "colors_precomp"), _cuda_f32(_opt(scales), "scales"),
                           _cuda_f32(_opt(rotations), "rotations"), _cuda_f32(_opt(cov3Ds_precomp), "cov3D_precomp"),
                           _cuda_f32(_opt(sh), "shs"), _cuda_f32(_opt(extra_colors), "extra_colors"),
                           g_color, g_extra if st.channels == 6 else None, g_depth, g_alpha, label=ctx.label,
                           want_means2D=ctx.means2D_shape is not None,
                           want_means2D_color=ctx.means2D_color_shape is not None)

        def handle_grad(t, shape):
            if shape is None:
                return None
            if len(shape) == 2:                    # [P,3] dummy shared by all views: sum over views
                t = t.sum(dim=0) if st.V > 1 else t[0]
            return t.reshape(shape)
        g_means2D = handle_grad(g["means2D"], ctx.means2D_shape)
        g_means2D_color = handle_grad(g["means2D_color"], ctx.means2D_color_shape)
        g_op = g["opacity"].reshape(opacities.shape)
        return (g["means3D"], g_means2D, g["sh"], g["colors"], g["extra"], g_op, g["scales"], g["rotations"],
                g["cov3D"], None, None, None, None, None, None, g_means2D_color)


def rasterize_views(settings: Sequence[GaussianRasterizationSettings], means3D, opacities, means2D=None, shs=None,
                    colors_precomp=None, scales=None, rotations=None, cov3D_precomp=None, extra_colors=None,
                    label=None, bg_extra=None, capacity: Optional[int] = None, state_out: Optional[list] = None,
                    bands=None, means2D_color=None):
    """Renders all `settings` (V cameras of equal resolution) in one batch.

    Returns (color [V,3,H,W], extra [V,3,H,W] or None, depth [V,1,H,W], alpha [V,1,H,W], radii [V,P]).
    `extra_colors` ([P,3], e.g. the fg/bg segmentation colours of REF/train.py:100,247-249) are blended in the same
    pass.  Gradients of the per-Gaussian inputs are summed over the views.

    `capacity` (instances the binning workspace is sized for) removes the reference's one host sync per call
    (rasterizer_impl.cu:287): pass e.g. 1.1x the previous step's count; append-only `state_out` receives the
    RasterState, whose `read_counters` tells afterwards whether the capacity was exceeded (surplus instances dropped).
    `bands[v]` = (first, last + 1) tile rows of view v to render (see forward_views); pixels outside the band come back
    as background and their cotangents are ignored.
    `means2D_color` ([P,3] or [V,P,3], requires_grad) is a second gradient handle like `means2D`: its .grad receives the
    screen-space gradient of the COLOUR render alone (colour, depth, alpha cotangents; not `extra_colors`'), which is what
    the reference accumulates for densification -- it renders colour and segmentation separately and keeps the colour
    call's means2D (REF/train.py:245: "Gradient only accum from colour render for densification").
    """
    if (shs is None) == (colors_precomp is None):
        raise Exception('Please provide excatly one of either SHs or precomputed colors!')
    if ((scales is None or rotations is None) and cov3D_precomp is None) or \
            ((scales is not None or rotations is not None) and cov3D_precomp is not None):
        raise Exception('Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!')
    e = means3D.new_empty(0)
    color, extra, depth, alpha, radii = _RasterizeViews.apply(
        means3D, means2D, e if shs is None else shs, e if colors_precomp is None else colors_precomp,
        e if extra_colors is None else extra_colors, opacities, e if scales is None else scales,
        e if rotations is None else rotations, e if cov3D_precomp is None else cov3D_precomp, list(settings), label,
        bg_extra, capacity, state_out, bands, means2D_color)
    return color, (extra if extra.numel() else None), depth, alpha, radii


class _RasterizeGaussians(torch.autograd.Function):
    """Single-view op with the reference's argument and gradient order (R/diff_gaussian_rasterization/__init__.py:48-174).
    Returns (color [3,H,W], radii [P], feature_map, depth [1,H,W], alpha [1,H,W]) like the vendored variant (:108)."""

    @staticmethod
    def forward(ctx, means3D, means2D, sh, colors_precomp, semantic_feature, opacities, scales, rotations,
                cov3Ds_precomp, raster_settings, label):
        feat = _opt(semantic_feature) if torch.is_tensor(semantic_feature) else None
        if raster_settings.debug:
            cpu_args = tuple(a.detach().cpu().clone() if torch.is_tensor(a) else a for a in
                             (means3D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp))
            try:
                out = forward_views([raster_settings], means3D, opacities, _opt(colors_precomp), _opt(scales),
                                    _opt(rotations), _opt(cov3Ds_precomp), _opt(sh), semantic_feature=feat)
                torch.cuda.synchronize(means3D.device)
            except Exception:
                torch.save(cpu_args, "snapshot_fw.dump")
                print("\nAn error occured in forward. Please forward snapshot_fw.dump for debugging.")
                raise
        else:
            out = forward_views([raster_settings], means3D, opacities, _opt(colors_precomp), _opt(scales),
                                _opt(rotations), _opt(cov3Ds_precomp), _opt(sh), semantic_feature=feat)
        color, _extra, depth, alpha, radii, st = out
        ctx.state = st
        ctx.raster_settings = raster_settings
        ctx.label = label if torch.is_tensor(label) else None
        ctx.has_feat = feat is not None
        ctx.save_for_backward(means3D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                              feat if feat is not None else means3D.new_empty(0))
        radii = radii[0]
        if feat is not None:
            feature_map = st.feature_map[0]
            st.feature_map = None
            ctx.mark_non_differentiable(radii)
        else:
            feature_map = color.new_empty(0)
            ctx.mark_non_differentiable(radii, feature_map)
        return color[0], radii, feature_map, depth[0], alpha[0]

    @staticmethod
    def backward(ctx, grad_color, _grad_radii, grad_feature, grad_depth, grad_alpha):
        means3D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp, feat = ctx.saved_tensors
        st = ctx.state
        feat = feat if ctx.has_feat else None

        def run():
            return backward_views(st, means3D.contiguous(), opacities.contiguous(),
                                  _cuda_f32(_opt(colors_precomp), "colors_precomp"),
                                  _cuda_f32(_opt(scales), "scales"), _cuda_f32(_opt(rotations), "rotations"),
                                  _cuda_f32(_opt(cov3Ds_precomp), "cov3D_precomp"), _cuda_f32(_opt(sh), "shs"), None,
                                  grad_color, None, grad_depth, grad_alpha, label=ctx.label,
                                  semantic_feature=feat, dL_dfeature=grad_feature if feat is not None else None)
        if ctx.raster_settings.debug:
            try:
                g = run()
                torch.cuda.synchronize(means3D.device)
            except Exception:
                torch.save(tuple(t.detach().cpu() for t in ctx.saved_tensors), "snapshot_bw.dump")
                print("\nAn error occured in backward. Writing snapshot_bw.dump for debugging.\n")
                raise
        else:
            g = run()
        g_op = g["opacity"].reshape(opacities.shape)
        # order: means3D, means2D, sh, colors_precomp, semantic_feature, opacities, scales, rotations, cov3Ds, settings, label
        g_feat = None
        if feat is not None:
            g_feat = g["features"].reshape(ctx.saved_tensors[7].shape)
        return (g["means3D"], g["means2D"][0], g["sh"], g["colors"], g_feat, g_op, g["scales"], g["rotations"],
                g["cov3D"], None, None)


def rasterize_gaussians(means3D, means2D, sh, colors_precomp, semantic_feature, opacities, scales, rotations,
                        cov3Ds_precomp, raster_settings, label):
    return _RasterizeGaussians.apply(means3D, means2D, sh, colors_precomp, semantic_feature, opacities, scales,
                                     rotations, cov3Ds_precomp, raster_settings, label)


DEFAULT_RETURNS = 3


class GaussianRasterizer(nn.Module):
    """Drop-in for R/diff_gaussian_rasterization/__init__.py:194-245.

    `returns` selects the tuple the caller family expects:
      3 -> (color, radii, depth)                       REF/train.py:142, REF/visualize_org.py:112
      4 -> (color, radii, depth, alpha)                REF/cvpr_dyn.py:284
      5 -> (color, radii, feature_map, depth, alpha)   the vendored variant as written (:108), REF/dyn_train.py:244
    """

    def __init__(self, raster_settings, returns: Optional[int] = None):
        super().__init__()
        self.raster_settings = raster_settings
        self.returns = DEFAULT_RETURNS if returns is None else returns
        if self.returns not in (3, 4, 5):
            raise ValueError("returns must be 3, 4 or 5")

    def markVisible(self, positions):
        with torch.no_grad():
            positions = _cuda_f32(positions, "positions", (3,))
            P = positions.shape[0]
            present = torch.zeros(P, dtype=torch.uint8, device=positions.device)
            if P:
                cam = camera_struct(self.raster_settings)
                with torch.cuda.device(positions.device):
                    _lib.check(_lib.load().gsr_mark_visible(C.byref(cam), P, positions.data_ptr(), present.data_ptr(),
                                                            _stream_ptr(positions.device)), "mark_visible")
        return present.bool()

    def forward(self, means3D, means2D, opacities=None, shs=None, semantic_feature=None, colors_precomp=None,
                scales=None, rotations=None, cov3D_precomp=None, label=1):
        raster_settings = self.raster_settings
        if (shs is None and colors_precomp is None) or (shs is not None and colors_precomp is not None):
            raise Exception('Please provide excatly one of either SHs or precomputed colors!')
        if ((scales is None or rotations is None) and cov3D_precomp is None) or \
                ((scales is not None or rotations is not None) and cov3D_precomp is not None):
            raise Exception('Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!')
        e = torch.Tensor([])
        shs = e if shs is None else shs
        colors_precomp = e if colors_precomp is None else colors_precomp
        scales = e if scales is None else scales
        rotations = e if rotations is None else rotations
        cov3D_precomp = e if cov3D_precomp is None else cov3D_precomp
        color, radii, feature_map, depth, alpha = rasterize_gaussians(
            means3D, means2D, shs, colors_precomp, semantic_feature, opacities, scales, rotations, cov3D_precomp,
            raster_settings, label)
        if self.returns == 3:
            return color, radii, depth
        if self.returns == 4:
            return color, radii, depth, alpha
        return color, radii, feature_map, depth, alpha
