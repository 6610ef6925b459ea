"""Synthetic CMU-Panoptic-shaped scenes and cameras.  Host-side numpy/torch only.

The real capture data is not in the reference tree (REF/.gitignore:3), so every test and benchmark runs on seeded
synthetic scenes with the statistics the reference's own initialisation produces:
log-scales from the mean 3-NN distance (REF/train.py:95-103), sigmoid opacities, unit quaternions, RGB in [0,1],
fg/bg segmentation colours (seg, 0, 1-seg) (REF/train.py:100), cameras built like REF/helpers.py:68-95.
"""
from __future__ import annotations

import math
from typing import List, NamedTuple, Optional

import numpy as np
import torch


class Scene(NamedTuple):
    means3D: torch.Tensor        # [P,3]
    log_scales: torch.Tensor     # [P,3]
    unnorm_rotations: torch.Tensor  # [P,4]
    logit_opacities: torch.Tensor   # [P,1]
    rgb_colors: torch.Tensor     # [P,3]
    seg_colors: torch.Tensor     # [P,3]

    def rendervar(self):
        """REF/helpers.py:98-107 activations."""
        return dict(
            means3D=self.means3D,
            colors_precomp=self.rgb_colors,
            rotations=torch.nn.functional.normalize(self.unnorm_rotations),
            opacities=torch.sigmoid(self.logit_opacities),
            scales=torch.exp(self.log_scales),
        )

    def to(self, device):
        return Scene(*[t.to(device) for t in self])


def _mean_knn_dist2(pts: np.ndarray, k: int = 3) -> np.ndarray:
    from scipy.spatial import cKDTree
    tree = cKDTree(pts)
    d, _ = tree.query(pts, k=k + 1, workers=-1)
    return (d[:, 1:] ** 2).mean(axis=1)


def _finish_scene(pts: np.ndarray, seg: np.ndarray, rng: np.random.Generator, scale_jitter=0.3,
                  max_scale: Optional[float] = None) -> Scene:
    P = pts.shape[0]
    d2 = np.clip(_mean_knn_dist2(pts.astype(np.float64)), 1e-7, None)
    if max_scale is not None:
        d2 = np.minimum(d2, max_scale ** 2)
    log_s = np.log(np.sqrt(d2))[:, None] + rng.normal(0.0, scale_jitter, size=(P, 3))
    q = rng.normal(size=(P, 4))
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    logit_o = rng.normal(0.0, 1.5, size=(P, 1))
    rgb = rng.uniform(0.0, 1.0, size=(P, 3))
    segc = np.stack([seg, np.zeros_like(seg), 1.0 - seg], axis=1)
    f = lambda a: torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32))
    return Scene(f(pts), f(log_s), f(q), f(logit_o), f(rgb), f(segc))


def make_scene_c1(P: int = 10_000, seed: int = 0) -> Scene:
    """Config C1: P Gaussians uniform in [-1,1]^3."""
    rng = np.random.default_rng(seed)
    pts = rng.uniform(-1.0, 1.0, size=(P, 3))
    seg = (rng.uniform(size=P) < 0.5).astype(np.float64)
    return _finish_scene(pts, seg, rng)


def make_scene_c2(P: int = 300_000, seed: int = 0, scale: float = 1.0) -> Scene:
    """Config C2: 60 % foreground in a cylinder (r=1, y in [-1.8,0], floor at y=0, up = -y: REF/train.py:275),
    40 % background on the floor disc (r=4) and a dome (r=5).  `scale` stretches the layout (C5 uses 2.0 / 1.6)."""
    rng = np.random.default_rng(seed)
    n_fg = int(P * 0.6)
    n_bg = P - n_fg
    r = np.sqrt(rng.uniform(size=n_fg)) * 1.0 * scale
    th = rng.uniform(0, 2 * math.pi, size=n_fg)
    fg = np.stack([r * np.cos(th), rng.uniform(-1.8, 0.0, size=n_fg) * scale, r * np.sin(th)], axis=1)
    n_floor = n_bg // 2
    n_dome = n_bg - n_floor
    r = np.sqrt(rng.uniform(size=n_floor)) * 4.0 * scale
    th = rng.uniform(0, 2 * math.pi, size=n_floor)
    floor = np.stack([r * np.cos(th), np.zeros(n_floor), r * np.sin(th)], axis=1)
    v = rng.normal(size=(n_dome, 3))
    v /= np.linalg.norm(v, axis=1, keepdims=True)
    v[:, 1] = -np.abs(v[:, 1])
    dome = v * 5.0 * scale
    pts = np.concatenate([fg, floor, dome], axis=0)
    seg = np.concatenate([np.ones(n_fg), np.zeros(n_bg)])
    return _finish_scene(pts, seg, rng, max_scale=0.25 * scale)


class CameraSpec(NamedTuple):
    w: int
    h: int
    k: list          # 3x3 intrinsics
    w2c: np.ndarray  # 4x4


def look_at_w2c(eye, target, up=(0.0, -1.0, 0.0)) -> np.ndarray:
    """World-to-camera with +z forward, +x right, +y down (the convention of REF/helpers.py:68-95 data)."""
    eye = np.asarray(eye, dtype=np.float64)
    target = np.asarray(target, dtype=np.float64)
    up = np.asarray(up, dtype=np.float64)
    f = target - eye
    f /= np.linalg.norm(f)
    r = np.cross(-up, f)         # down x forward = right
    r /= np.linalg.norm(r)
    d = np.cross(f, r)
    Rm = np.stack([r, d, f], axis=0)
    w2c = np.eye(4)
    w2c[:3, :3] = Rm
    w2c[:3, 3] = -Rm @ eye
    return w2c


def camera_c1() -> CameraSpec:
    """C1: one 256x256 camera 3 m from the origin, fx=fy=280, off-centre principal point."""
    w = h = 256
    k = [[280.0, 0.0, 128.0 + 3.5], [0.0, 280.0, 128.0 - 2.25], [0.0, 0.0, 1.0]]
    return CameraSpec(w, h, k, look_at_w2c((0.6, -0.5, -2.9), (0.0, 0.0, 0.0)))


def cameras_c2(n: int = 27, w: int = 640, h: int = 360, f: float = 467.0, radius: float = 3.0,
               target=(0.0, -0.9, 0.0), seed: int = 1) -> List[CameraSpec]:
    """C2: 27 cameras on a dome, elevations {10,35,60} deg x 9 azimuths, principal point = centre +- U(0,10) px."""
    rng = np.random.default_rng(seed)
    cams = []
    elevs = [10.0, 35.0, 60.0]
    per = int(math.ceil(n / len(elevs)))
    i = 0
    for e in elevs:
        for a in range(per):
            if i >= n:
                break
            az = 2 * math.pi * (a + 0.3 * elevs.index(e)) / per
            el = math.radians(e)
            eye = np.array([radius * math.cos(el) * math.cos(az), -radius * math.sin(el) + target[1],
                            radius * math.cos(el) * math.sin(az)])
            cx = w / 2 + rng.uniform(-10, 10)
            cy = h / 2 + rng.uniform(-10, 10)
            k = [[f, 0.0, cx], [0.0, f, cy], [0.0, 0.0, 1.0]]
            cams.append(CameraSpec(w, h, k, look_at_w2c(eye, target)))
            i += 1
    return cams
