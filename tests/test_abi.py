"""CPU tests of the drop-in boundary: the C-ABI library loads, exports every symbol include/gsr_b200.h declares,
answers the host-only queries, and the Python host layer mirrors the reference API's validation (no GPU needed)."""
import ctypes as C
import os
import re
import sys

import pytest
import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from helpers import ROOT  # noqa: E402
from dynamic3dgaussians_b200 import _lib, rasterizer as RZ  # noqa: E402


def _declared_symbols():
    src = open(os.path.join(ROOT, "include", "gsr_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(gsr_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    lib = _lib.load()
    names = _declared_symbols()
    assert len(names) >= 15
    for n in names:
        assert hasattr(lib, n), n
    assert sorted(_lib.EXPORTED_SYMBOLS) == names
    assert lib.gsr_abi_version() == 4
    assert lib.gsr_strerror(0) == b"ok" and lib.gsr_strerror(-2) == b"workspace too small"
    # host-only switches (the library reads no environment variables): unknown keys / values are error codes
    v = C.c_int(-1)
    assert lib.gsr_get_option(b"bwd_cost_order", C.byref(v)) == 0 and v.value == 1
    assert lib.gsr_set_option(b"nope", 1) == -1 and lib.gsr_set_option(b"blend_ilp", 3) == -1


def test_workspace_protocol_host_side():
    lib = _lib.load()
    s = _lib.GsrWorkspaceSizes()
    assert lib.gsr_workspace_sizes(300_000, 27, 640, 360, 6, 0, 40_000_000, C.byref(s)) == 0
    assert s.geom_bytes > 27 * 300_000 * 100 and s.binning_bytes > 40_000_000 * 64 and s.image_bytes > 27 * 640 * 360 * 8
    L = RZ.workspace_layout(300_000, 27, 640, 360, 3, 0, 1000)
    assert (L.tiles_x, L.tiles_y) == (40, 23)
    assert (L.stiles_x, L.stiles_y) == (10, 6)                 # supertiles of 4x4 tiles
    assert L.tile_bits == 11 and L.tile_passes == 1 and L.depth_passes == 4      # key = 6 supertile bits << 5 | view:
    # one sort pass on the supertile bits; 32 depth bits per view
    assert L.counters == 0
    offs = [getattr(L, n) for n in _lib._LAYOUT_SIZE_T[:18]]
    assert all(o % 256 == 0 for o in offs)
    # what a forward needs zeroed sits at the head of its workspace (one memset per workspace)
    assert max(L.counters, L.dhist, L.dlookback, L.dcounters) < min(L.cams, L.radii, L.rects, L.geomrec, L.dkeys0, L.grad_rec)
    assert max(L.ihist, L.ilookback, L.st_ranges) < min(L.ikeys0, L.ivals0, L.imasks, L.chunk_base, L.recA)
    # error codes, not exceptions, across the ABI
    assert lib.gsr_workspace_sizes(10, 1, 0, 10, 3, 0, 0, C.byref(s)) == -1
    assert lib.gsr_workspace_sizes(10, 1, 16, 16, 4, 0, 0, C.byref(s)) == -4
    assert lib.gsr_sort_temp_bytes(1000, 8) > 1000 * 24
    with pytest.raises(RuntimeError):
        _lib.check(-2, "x")


def test_settings_tuple_matches_reference_fields():
    ref_fields = ("image_height", "image_width", "tanfovx", "tanfovy", "c_x", "c_y", "bg", "scale_modifier",
                  "viewmatrix", "projmatrix", "sh_degree", "campos", "prefiltered", "debug", "confidence")
    assert RZ.GaussianRasterizationSettings._fields == ref_fields      # R/.../__init__.py:176-191
    s = RZ.GaussianRasterizationSettings(image_height=4, image_width=8, tanfovx=1.0, tanfovy=0.5,
                                         bg=torch.zeros(3), viewmatrix=torch.eye(4), projmatrix=torch.eye(4),
                                         campos=torch.zeros(3))
    cam = RZ.camera_struct(s)
    assert cam.c_x == 4.0 and cam.c_y == 2.0 and cam.image_width == 8       # default principal point = centre
    assert cam.debug == 0 and cam.prefiltered == 0
    assert RZ.camera_struct(s) is cam                                       # cached per settings object ...
    s.bg[0] = 0.5                                                           # ... until a tensor changes in place
    cam2 = RZ.camera_struct(s)
    assert cam2 is not cam and cam2.bg[0] == 0.5
    assert RZ.camera_struct(s._replace(debug=True, prefiltered=True)).debug == 1


def test_reference_validation_errors_and_no_cpu_path():
    s = RZ.GaussianRasterizationSettings(image_height=16, image_width=16, tanfovx=1.0, tanfovy=1.0, bg=torch.zeros(3),
                                         viewmatrix=torch.eye(4), projmatrix=torch.eye(4), campos=torch.zeros(3))
    r = RZ.GaussianRasterizer(s)
    m = torch.zeros(4, 3)
    with pytest.raises(Exception, match="SHs or precomputed colors"):
        r(m, m, opacities=torch.ones(4, 1), scales=torch.ones(4, 3), rotations=torch.ones(4, 4))
    with pytest.raises(Exception, match="scale/rotation pair or precomputed 3D covariance"):
        r(m, m, opacities=torch.ones(4, 1), colors_precomp=torch.ones(4, 3), scales=torch.ones(4, 3))
    with pytest.raises(ValueError, match="CUDA tensor"):
        r(m, m, opacities=torch.ones(4, 1), colors_precomp=torch.ones(4, 3), scales=torch.ones(4, 3),
          rotations=torch.ones(4, 4))
    with pytest.raises(ValueError):
        RZ.GaussianRasterizer(s, returns=2)


def test_drop_in_packages_reexport_the_reference_names():
    import diff_gaussian_rasterization as D
    import simple_knn._C as K
    for n in ("GaussianRasterizationSettings", "GaussianRasterizer", "_RasterizeGaussians", "rasterize_gaussians"):
        assert hasattr(D, n)
    assert hasattr(K, "distCUDA2")
