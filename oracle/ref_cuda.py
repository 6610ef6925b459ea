"""TEST INFRASTRUCTURE ONLY.  ctypes wrapper of oracle/_ref/libref_rasterizer.so / libref_knn.so: the reference's own
CUDA code (unmodified, compiled by oracle/build_ref.py) as a second witness for parity and as the timing comparator.
Needs a GPU.  Never imported by the product package."""
from __future__ import annotations

import ctypes as C
import os

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
RAS = os.path.join(HERE, "_ref", "libref_rasterizer.so")
RAS_PATCHED = os.path.join(HERE, "_ref", "libref_rasterizer_patched.so")   # SURVEY.md 2.4 repairs, see build_ref.PATCHES
KNN = os.path.join(HERE, "_ref", "libref_knn.so")

_ras = {}
_knn = None


def available() -> bool:
    return os.path.exists(RAS)


def patched_available() -> bool:
    return os.path.exists(RAS_PATCHED)


def _load(patched: bool = False):
    lib = _ras.get(patched)
    if lib is None:
        lib = C.CDLL(RAS_PATCHED if patched else RAS)
        vp, i, f = C.c_void_p, C.c_int, C.c_float
        lib.ref_forward.argtypes = [i, i, i, vp, i, i, vp, vp, vp, vp, vp, vp, f, vp, vp, vp, vp, vp, f, f, f, f,
                                    vp, vp, vp, vp, vp]
        lib.ref_backward.argtypes = [i, i, i, vp, i, i, vp, vp, vp, vp, vp, vp, f, vp, vp, vp, vp, vp, f, f, f, f, vp,
                                     vp, vp, vp, vp] + [vp] * 11
        lib.ref_export_state.argtypes = [vp] * 8
        _ras[patched] = lib
    return lib


def _p(t):
    return None if t is None else t.data_ptr()


class RefRasterizer:
    """One view at a time, like the reference.  `cam` is any object with the GaussianRasterizationSettings fields."""

    def __init__(self, cam, patched: bool = False):
        self.cam = cam
        self.patched = patched
        dev = "cuda"
        self.view = cam.viewmatrix.reshape(-1).float().to(dev).contiguous()
        self.proj = cam.projmatrix.reshape(-1).float().to(dev).contiguous()
        self.campos = cam.campos.reshape(-1).float().to(dev).contiguous()
        bg = torch.zeros(32, device=dev)            # the reference reads bg[ch] for ch < 32
        bg[:3] = cam.bg.reshape(-1).float().to(dev)[:3]
        self.bg = bg
        self.W, self.H = int(cam.image_width), int(cam.image_height)
        self.c_x = float(cam.c_x) if cam.c_x is not None else 0.5 * self.W
        self.c_y = float(cam.c_y) if cam.c_y is not None else 0.5 * self.H

    def forward(self, means3D, opacities, colors, scales, rotations, semantic=None, shs=None, cov3D_precomp=None):
        """Exactly one of (colors | shs [P,M,3]) and one of (scales + rotations | cov3D_precomp [P,6]), as
        R/diff_gaussian_rasterization/__init__.py:214-218; D = cam.sh_degree, M = shs.shape[1] (rasterize_points.cu:85-89)."""
        lib = _load(self.patched)
        P = means3D.shape[0]
        dev = means3D.device
        if semantic is None:
            semantic = torch.zeros(P, 32, device=dev)
        self.semantic = semantic
        H, W = self.H, self.W
        color = torch.zeros(3, H, W, device=dev)
        feat = torch.zeros(32, H, W, device=dev)
        depth = torch.zeros(1, H, W, device=dev)
        alpha = torch.zeros(1, H, W, device=dev)
        radii = torch.zeros(P, dtype=torch.int32, device=dev)
        cam = self.cam
        D = int(cam.sh_degree) if shs is not None else 0
        M = int(shs.shape[1]) if shs is not None else 0
        self.DM = (D, M)
        R = lib.ref_forward(P, D, M, _p(self.bg), W, H, _p(means3D), _p(shs), _p(colors), _p(semantic), _p(opacities),
                            _p(scales), float(cam.scale_modifier), _p(rotations), _p(cov3D_precomp), _p(self.view), _p(self.proj),
                            _p(self.campos), self.c_x, self.c_y, float(cam.tanfovx), float(cam.tanfovy), _p(color),
                            _p(feat), _p(depth), _p(alpha), _p(radii))
        if R < 0:
            raise RuntimeError("reference forward failed")
        self.saved = (means3D, opacities, colors, scales, rotations, radii, alpha)
        self.saved_opt = (shs, cov3D_precomp)
        return dict(color=color, feature=feat, depth=depth, alpha=alpha, radii=radii, num_rendered=R)

    def backward(self, dL_dcolor, dL_ddepth, dL_dalpha, alphas=None):
        """Calls Rasterizer::backward with the argument order the C++ expects (the Python swap bug Q2 is not
        reproduced here).  `alphas` defaults to what the forward returned (all zeros)."""
        lib = _load(self.patched)
        means3D, opacities, colors, scales, rotations, radii, alpha = self.saved
        shs, cov3D_precomp = self.saved_opt
        D, M = self.DM
        P = means3D.shape[0]
        dev = means3D.device
        H, W = self.H, self.W
        z = lambda *s: torch.zeros(*s, device=dev)
        g = dict(means2D=z(P, 3), conic=z(P, 4), opacity=z(P, 1), colors=z(P, 3), semantic=z(P, 32), depth=z(P, 1),
                 means3D=z(P, 3), cov3D=z(P, 6), sh=z(P, max(M, 1), 3), scales=z(P, 3), rotations=z(P, 4))
        dfeat = z(32, H, W)
        cam = self.cam
        rc = lib.ref_backward(P, D, M, _p(self.bg), W, H, _p(means3D), _p(shs), _p(colors), _p(self.semantic),
                              _p(alpha if alphas is None else alphas), _p(scales), float(cam.scale_modifier),
                              _p(rotations), _p(cov3D_precomp), _p(self.view), _p(self.proj), _p(self.campos), self.c_x, self.c_y,
                              float(cam.tanfovx), float(cam.tanfovy), _p(radii), _p(dL_dcolor.contiguous()), _p(dfeat),
                              _p(dL_ddepth.contiguous()), _p(dL_dalpha.contiguous()), _p(g["means2D"]), _p(g["conic"]),
                              _p(g["opacity"]), _p(g["colors"]), _p(g["semantic"]), _p(g["depth"]), _p(g["means3D"]),
                              _p(g["cov3D"]), _p(g["sh"]), _p(g["scales"]), _p(g["rotations"]))
        if rc != 0:
            raise RuntimeError("reference backward failed")
        return g

    def export_state(self, R):
        lib = _load(self.patched)
        dev = "cuda"
        P = self.saved[0].shape[0]
        tiles = ((self.W + 15) // 16) * ((self.H + 15) // 16)
        out = dict(
            keys=torch.zeros(max(R, 1), dtype=torch.int64, device=dev),
            point_list=torch.zeros(max(R, 1), dtype=torch.int32, device=dev),
            ranges=torch.zeros(tiles, 2, dtype=torch.int32, device=dev),
            n_contrib=torch.zeros(self.H, self.W, dtype=torch.int32, device=dev),
            tiles_touched=torch.zeros(P, dtype=torch.int32, device=dev),
            means2D=torch.zeros(P, 2, device=dev), depths=torch.zeros(P, device=dev),
            conic_opacity=torch.zeros(P, 4, device=dev))
        rc = lib.ref_export_state(*[_p(out[k]) for k in ("keys", "point_list", "ranges", "n_contrib", "tiles_touched",
                                                          "means2D", "depths", "conic_opacity")])
        if rc != 0:
            raise RuntimeError("reference export failed")
        out["keys"] = out["keys"][:R]
        out["point_list"] = out["point_list"][:R]
        return out


def knn(points: torch.Tensor):
    global _knn
    if _knn is None:
        _knn = C.CDLL(KNN)
        _knn.ref_knn.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
    P = points.shape[0]
    pts = points.float().contiguous()
    d = torch.zeros(P, device=pts.device)
    idx = torch.zeros(P, 3, dtype=torch.int32, device=pts.device)
    if _knn.ref_knn(P, pts.data_ptr(), d.data_ptr(), idx.data_ptr()) != 0:
        raise RuntimeError("reference knn failed")
    return d, idx
