"""params.npz / PLY / PPM round trips (CPU)."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import helpers  # noqa: E402,F401
from dynamic3dgaussians_b200 import scene_io as IO  # noqa: E402


def _params(P, seed):
    g = torch.Generator().manual_seed(seed)
    seg = (torch.rand(P, generator=g) > 0.4).float()
    return {'means3D': torch.randn(P, 3, generator=g), 'rgb_colors': torch.rand(P, 3, generator=g),
            'seg_colors': torch.stack((seg, torch.zeros(P), 1 - seg), -1), 'unnorm_rotations': torch.randn(P, 4, generator=g),
            'logit_opacities': torch.randn(P, 1, generator=g), 'log_scales': torch.randn(P, 3, generator=g) - 3,
            'cam_m': torch.zeros(5, 3), 'cam_c': torch.zeros(5, 3)}


def test_params_npz_layout_and_scene_data(tmp_path):
    p0, p1, p2 = _params(50, 0), _params(50, 1), _params(50, 2)
    out = [IO.params2cpu(p0, True), IO.params2cpu(p1, False), IO.params2cpu(p2, False)]
    assert set(out[1]) == set(IO.PER_TIMESTEP_KEYS)                       # REF/helpers.py:153
    path = IO.save_params(out, str(tmp_path / "exp" / "seq" / "params"))
    z = dict(np.load(path))
    assert z['means3D'].shape == (3, 50, 3) and z['unnorm_rotations'].shape == (3, 50, 4)
    assert z['log_scales'].shape == (50, 3) and z['seg_colors'].shape == (50, 3) and z['cam_m'].shape == (5, 3)
    scene, is_fg = IO.load_scene_data(path, device="cpu")
    assert len(scene) == 3 and torch.equal(is_fg, p0['seg_colors'][:, 0] > 0.5)
    assert torch.allclose(scene[2]['means3D'], p2['means3D'])
    assert torch.allclose(scene[1]['rotations'], torch.nn.functional.normalize(p1['unnorm_rotations']))
    assert torch.allclose(scene[0]['opacities'], torch.sigmoid(p0['logit_opacities']))
    assert torch.allclose(scene[0]['scales'], torch.exp(p0['log_scales']))
    seg_scene, _ = IO.load_scene_data(path, seg_as_col=True, device="cpu")
    assert torch.equal(seg_scene[1]['colors_precomp'], p0['seg_colors'])
    single = IO.save_params(out[:1], str(tmp_path / "one"))
    s1, _ = IO.load_scene_data(single, device="cpu")
    assert len(s1) == 1 and torch.allclose(s1[0]['means3D'], p0['means3D'])


def test_ply_header_and_round_trip(tmp_path):
    p = {k: v.numpy() for k, v in _params(123, 3).items()}
    path = str(tmp_path / "points.ply")
    IO.save_ply_splat(path, p['means3D'], p['log_scales'][:, :1], p['unnorm_rotations'], p['rgb_colors'], p['logit_opacities'])
    head = open(path, "rb").read(600).decode("ascii", errors="ignore")
    assert head.startswith("ply\nformat binary_little_endian 1.0\nelement vertex 123\n")
    order = [ln.split()[-1] for ln in head.split("end_header")[0].splitlines() if ln.startswith("property")]
    assert order == IO.PLY_ATTRS
    back = IO.load_ply_splat(path)
    assert np.allclose(back['means3D'], p['means3D']) and np.allclose(back['unnorm_rotations'], p['unnorm_rotations'])
    assert np.allclose(back['rgb_colors'], p['rgb_colors'], atol=1e-6)     # through the DC spherical-harmonic coefficient
    assert np.allclose(back['log_scales'], np.tile(p['log_scales'][:, :1], (1, 3)))
    assert np.allclose(back['logit_opacities'], p['logit_opacities']) and float(np.abs(back['normals']).max()) == 0.0


def test_init_point_cloud_meta_and_ppm(tmp_path):
    data = np.random.default_rng(0).random((20, 7))
    np.savez(tmp_path / "init_pt_cld.npz", data=data)
    pts, rgb, seg = IO.load_init_point_cloud(str(tmp_path / "init_pt_cld.npz"))
    assert pts.shape == (20, 3) and rgb.shape == (20, 3) and seg.shape == (20,) and pts.dtype == np.float32
    with open(tmp_path / "train_meta.json", "w") as f:
        f.write('{"w": 640, "h": 360, "k": [[[[1,0,0],[0,1,0],[0,0,1]]]], "fn": [["a.jpg"]]}')
    md = IO.load_train_meta(str(tmp_path / "train_meta.json"))
    assert md["w"] == 640 and len(md["fn"]) == 1
    img = torch.zeros(3, 4, 6)
    img[0] = 1.0
    IO.save_image_ppm(str(tmp_path / "im.ppm"), img)
    raw = open(tmp_path / "im.ppm", "rb").read()
    assert raw.startswith(b"P6\n6 4\n255\n") and raw[-3:] == bytes([255, 0, 0]) and len(raw) == 11 + 72
