"""Camera construction for the rasterizer: host-side mirror of REF/helpers.py:68-95 (`setup_camera`)."""
from __future__ import annotations

import numpy as np
import torch

from .rasterizer import GaussianRasterizationSettings


def setup_camera(w, h, k, w2c, near=0.01, far=100.0, bg=(0.0, 0.0, 0.0), device="cuda") -> GaussianRasterizationSettings:
    """Same arithmetic and conventions as the reference's setup_camera: viewmatrix = w2c^T, projmatrix = w2c^T @ P^T,
    tanfov = size / (2 f), principal point passed through as c_x, c_y."""
    fx, fy, cx, cy = k[0][0], k[1][1], k[0][2], k[1][2]
    # the 4x4 algebra runs on the host (the reference does it on the GPU: ~15 tiny launches per camera) and the three
    # small tensors are moved to `device` once
    w2c = torch.tensor(np.asarray(w2c), dtype=torch.float32)
    cam_center = torch.inverse(w2c)[:3, 3].to(device)
    w2c = w2c.unsqueeze(0).transpose(1, 2)
    opengl_proj = torch.tensor([[2 * fx / w, 0.0, -(w - 2 * cx) / w, 0.0],
                                [0.0, 2 * fy / h, -(h - 2 * cy) / h, 0.0],
                                [0.0, 0.0, far / (far - near), -(far * near) / (far - near)],
                                [0.0, 0.0, 1.0, 0.0]], dtype=torch.float32).unsqueeze(0).transpose(1, 2)
    full_proj = w2c.bmm(opengl_proj).contiguous().to(device)
    w2c = w2c.contiguous().to(device)
    return GaussianRasterizationSettings(
        image_height=h, image_width=w, tanfovx=w / (2 * fx), tanfovy=h / (2 * fy), c_x=cx, c_y=cy,
        bg=torch.tensor(bg, dtype=torch.float32, device=device), scale_modifier=1.0, viewmatrix=w2c,
        projmatrix=full_proj, sh_degree=0, campos=cam_center, prefiltered=False, debug=False, confidence=None)


def settings_from_spec(cs, bg=(0.0, 0.0, 0.0), device="cuda") -> GaussianRasterizationSettings:
    return setup_camera(cs.w, cs.h, cs.k, cs.w2c, bg=bg, device=device)
