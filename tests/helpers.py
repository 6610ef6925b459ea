"""Shared test helpers (test infrastructure; may import oracle/)."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from dynamic3dgaussians_b200 import rasterizer as RZ  # noqa: E402
from oracle import eager_rasterizer as O  # noqa: E402

GOLDEN = os.path.join(ROOT, "tests", "golden", "small_scene.npz")


def settings_from_spec(cs, bg=(0.0, 0.0, 0.0), device="cuda"):
    """GaussianRasterizationSettings (product type) + the oracle camera for one CameraSpec."""
    oc = O.camera_from_spec(cs.w, cs.h, cs.k, cs.w2c, bg=bg)
    st = RZ.GaussianRasterizationSettings(
        image_height=cs.h, image_width=cs.w, tanfovx=oc.tanfovx, tanfovy=oc.tanfovy, c_x=oc.c_x, c_y=oc.c_y,
        bg=oc.bg.to(device), scale_modifier=1.0, viewmatrix=oc.viewmatrix.to(device),
        projmatrix=oc.projmatrix.to(device), sh_degree=0, campos=oc.campos.to(device), prefiltered=False, debug=False)
    return st, oc


def rel_err(a, b):
    """max |a-b| / max |b|: the gradient metric of the parity tests (tolerance 1e-4, atomics reorder sums)."""
    a = a.detach().double().cpu()
    b = b.detach().double().cpu()
    return float((a - b).abs().max() / (b.abs().max() + 1e-30))


def elem_violations(a, b, rtol=1e-4, floor=1e-6):
    """Per-element gradient criterion |a - b| <= rtol |b| + floor max|b|  (BASELINE.json: "gradients within 1e-4
    relative"; the floor keeps elements that are sums of cancelling terms from being judged against their own rounding
    noise).  Returns (number of violating elements, number of elements)."""
    a = a.detach().double().cpu().reshape(-1)
    b = b.detach().double().cpu().reshape(-1)
    bad = (a - b).abs() > rtol * b.abs() + floor * b.abs().max()
    return int(bad.sum()), a.numel()


def final_T(st):
    """[V,H,W] final transmittance per pixel, as the forward left it in the image workspace."""
    L = RZ.workspace_layout(st.P, st.V, st.W, st.H, st.channels, st.sh_coeffs, st.capacity)
    n = st.V * st.H * st.W
    return st.image[L.final_T: L.final_T + 4 * n].view(torch.float32).view(st.V, st.H, st.W).clone()


def export_view(st, view=0):
    """(keys u64-as-int64, point_list, ranges [T,2], n_contrib [H,W]) of one view in the reference's format."""
    from dynamic3dgaussians_b200 import _lib
    lib = _lib.load()
    dev = st.geom.device
    T = ((st.W + 15) // 16) * ((st.H + 15) // 16)
    ranges = torch.zeros(T, 2, dtype=torch.int32, device=dev)
    ncon = torch.zeros(st.H, st.W, dtype=torch.int32, device=dev)
    s = torch.cuda.current_stream().cuda_stream
    args = (view, st.V, st.P, st.W, st.H, st.channels, st.sh_coeffs, st.capacity, st.geom.data_ptr(),
            st.binning.data_ptr(), st.image.data_ptr())
    _lib.check(lib.gsr_export_view_binning(*args, None, None, ranges.data_ptr(), ncon.data_ptr(), s), "export")
    n = int(ranges[:, 1].max().item()) if T else 0
    keys = torch.zeros(max(n, 1), dtype=torch.int64, device=dev)
    plist = torch.zeros(max(n, 1), dtype=torch.int32, device=dev)
    _lib.check(lib.gsr_export_view_binning(*args, keys.data_ptr(), plist.data_ptr(), None, None, s), "export")
    torch.cuda.synchronize()
    return keys[:n], plist[:n], ranges, ncon
