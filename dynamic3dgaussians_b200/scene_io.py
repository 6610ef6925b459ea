"""On-disk formats either side of the hot path.

  * `params.npz` -- the reference's training output: per-timestep stacks of means3D / rgb_colors / unnorm_rotations and
    the static rest (REF/helpers.py:149-168 `params2cpu` / `save_params`; read back by REF/vis_utils.py:83-99);
  * the 3DGS-style binary PLY the fork dumps next to it (REF/helpers.py:40-66 `save_ply_splat`: x y z, normals, DC SH
    colour, opacity logit, log scales, rotation) -- written and read here without the `plyfile` dependency;
  * `init_pt_cld.npz` / `train_meta.json` readers (REF/train.py:89-96,396-397);
  * a headless image dump (binary PPM) in place of the Open3D viewer, which needs a graphics stack (unavailable on
    this box)."""
from __future__ import annotations

import json
import os
from typing import Dict, List, Sequence

import numpy as np
import torch

SH_C0 = 0.28209479177387814
PER_TIMESTEP_KEYS = ('means3D', 'rgb_colors', 'unnorm_rotations')          # REF/helpers.py:153
PLY_ATTRS = ['x', 'y', 'z', 'nx', 'ny', 'nz', 'f_dc_0', 'f_dc_1', 'f_dc_2', 'opacity',
             'scale_0', 'scale_1', 'scale_2', 'rot_0', 'rot_1', 'rot_2', 'rot_3']      # REF/helpers.py:51-56


def params2cpu(params: Dict[str, torch.Tensor], is_initial_timestep: bool) -> Dict[str, np.ndarray]:
    """REF/helpers.py:149-155: everything at t = 0, only the moving quantities afterwards."""
    keys = params.keys() if is_initial_timestep else [k for k in params if k in PER_TIMESTEP_KEYS]
    return {k: params[k].detach().cpu().contiguous().numpy() for k in keys}


def save_params(output_params: Sequence[Dict[str, np.ndarray]], path: str) -> str:
    """REF/helpers.py:158-168: keys present at later timesteps are stacked [T, ...], the rest saved once."""
    to_save = {}
    for k in output_params[0].keys():
        if len(output_params) > 1 and k in output_params[1].keys():
            to_save[k] = np.stack([p[k] for p in output_params])
        else:
            to_save[k] = output_params[0][k]
    os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
    np.savez(path, **to_save)
    return path if path.endswith('.npz') else path + '.npz'


def load_scene_data(path: str, seg_as_col: bool = False, device="cuda"):
    """REF/vis_utils.py:83-118: one rendervar dict per timestep (+ the foreground mask).  Accepts both the stacked
    multi-timestep file and a single-timestep dump."""
    params = {k: torch.tensor(v).to(device).float() for k, v in dict(np.load(path)).items()}
    is_fg = params['seg_colors'][:, 0] > 0.5
    stacked = params['means3D'].dim() == 3
    T = params['means3D'].shape[0] if stacked else 1
    scene = []
    for t in range(T):
        pick = (lambda k: params[k][t]) if stacked else (lambda k: params[k])
        scene.append({
            'means3D': pick('means3D'),
            'colors_precomp': pick('rgb_colors') if not seg_as_col else params['seg_colors'],
            'rotations': torch.nn.functional.normalize(pick('unnorm_rotations')),
            'opacities': torch.sigmoid(params['logit_opacities']),
            'scales': torch.exp(params['log_scales']),
            'means2D': torch.zeros_like(pick('means3D')),
        })
    return scene, is_fg


def rgb_to_spherical_harmonic(rgb):
    return (rgb - 0.5) / SH_C0                                              # REF/helpers.py:9-10


def spherical_harmonic_to_rgb(sh):
    return sh * SH_C0 + 0.5                                                 # REF/helpers.py:13-14


def save_ply_splat(path: str, means, scales, rotations, rgbs, opacities, normals=None) -> None:
    """REF/helpers.py:40-66.  Inputs are numpy arrays in the PARAMETER domain the fork dumps (log scales, opacity
    logits, unnormalised rotations, RGB in [0, 1]); [P,1] scales are tiled to 3."""
    means = np.asarray(means, np.float32)
    P = means.shape[0]
    normals = np.zeros_like(means) if normals is None else np.asarray(normals, np.float32)
    colors = rgb_to_spherical_harmonic(np.asarray(rgbs, np.float32))
    scales = np.asarray(scales, np.float32)
    if scales.shape[1] == 1:
        scales = np.tile(scales, (1, 3))
    table = np.concatenate((means, normals, colors, np.asarray(opacities, np.float32).reshape(P, 1), scales,
                            np.asarray(rotations, np.float32)), axis=1).astype('<f4')
    assert table.shape[1] == len(PLY_ATTRS)
    header = "ply\nformat binary_little_endian 1.0\n" + f"element vertex {P}\n" + \
             "".join(f"property float {a}\n" for a in PLY_ATTRS) + "end_header\n"
    os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
    with open(path, "wb") as f:
        f.write(header.encode("ascii"))
        f.write(table.tobytes())


def load_ply_splat(path: str) -> Dict[str, np.ndarray]:
    """Reads a PLY written by save_ply_splat (or any binary little-endian float-only vertex PLY)."""
    with open(path, "rb") as f:
        names, n = [], 0
        line = f.readline().strip()
        if line != b"ply":
            raise ValueError("not a PLY file")
        while True:
            raw = f.readline()
            if not raw:
                raise ValueError("truncated PLY header: reached the end of the file before 'end_header'")
            line = raw.decode("ascii", errors="replace").strip()
            if line.startswith("format") and "binary_little_endian" not in line:
                raise ValueError("only binary little-endian PLY is supported")
            if line.startswith("element vertex"):
                n = int(line.split()[-1])
            elif line.startswith("property"):
                _, ty, name = line.split()
                if ty not in ("float", "float32"):
                    raise ValueError(f"unsupported property type {ty}")
                names.append(name)
            elif line == "end_header":
                break
        buf = f.read(n * len(names) * 4)
        if len(buf) != n * len(names) * 4:
            raise ValueError(f"truncated PLY body: header promises {n} vertices x {len(names)} floats "
                             f"({n * len(names) * 4} bytes), file holds {len(buf)}")
        table = np.frombuffer(buf, dtype='<f4').reshape(n, len(names))
    col = {a: table[:, i] for i, a in enumerate(names)}
    g = lambda *ks: np.stack([col[k] for k in ks], axis=1)
    return {'means3D': g('x', 'y', 'z'), 'normals': g('nx', 'ny', 'nz'),
            'rgb_colors': spherical_harmonic_to_rgb(g('f_dc_0', 'f_dc_1', 'f_dc_2')),
            'logit_opacities': g('opacity'), 'log_scales': g('scale_0', 'scale_1', 'scale_2'),
            'unnorm_rotations': g('rot_0', 'rot_1', 'rot_2', 'rot_3')}


def load_init_point_cloud(path: str):
    """`init_pt_cld.npz` (REF/train.py:90-94): data [N, 7] = xyz, rgb, seg.  Returns (points, rgb, seg) float32."""
    data = np.load(path)["data"].astype(np.float32)
    return data[:, :3], data[:, 3:6], data[:, 6]


def load_train_meta(path: str) -> dict:
    """`train_meta.json` (REF/train.py:396-397): 'w', 'h', per-timestep per-camera 'k', 'w2c', 'fn', 'cam_id'."""
    with open(path, "r") as f:
        return json.load(f)


def save_image_ppm(path: str, img: torch.Tensor) -> None:
    """[3,H,W] float image in [0,1] -> binary PPM (P6).  The headless stand-in for the reference's viewers."""
    a = (img.detach().clamp(0, 1) * 255.0 + 0.5).to(torch.uint8).permute(1, 2, 0).contiguous().cpu().numpy()
    os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
    with open(path, "wb") as f:
        f.write(f"P6\n{a.shape[1]} {a.shape[0]}\n255\n".encode("ascii"))
        f.write(a.tobytes())
