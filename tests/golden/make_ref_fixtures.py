"""TEST INFRASTRUCTURE.  Generates tests/golden/ref_python_fixtures.npz by RUNNING THE REFERENCE'S OWN PYTHON, unmodified,
on seeded inputs, so that the oracles for SURVEY.md rows a22 and 8f-1..4 carry a reference pin:

    REF/external.py   calc_ssim :102, build_rotation :61, accumulate_mean2d_gradient :136, densify :244 (with
                      cat_params_to_optimizer :159, remove_points :186, update_params_and_optimizer :143)
    REF/helpers.py    setup_camera :68, params2rendervar :98, l1_loss_v1/_v2 :110-115, weighted_l2_loss_v1/_v2 :118-123,
                      quat_mult :126, params2cpu :149, save_params :158
    REF/train.py      initialize_optimizer :119, initialize_per_timestep :294, initialize_post_first_timestep :316 and the
                      rigidity / rotation / isometry block of get_loss :253-273.  train.py does not parse (:284-285), so
                      the three functions are exec'ed from their own source lines and the block is re-assembled, statement
                      by statement, from the IMPORTED helpers.

How the reference is made to run here (this container has /root/reference but no GPU, no open3d, no plyfile):
  (i)   an EMPTY module is registered as `open3d` (an import stub only; REF/helpers.py:3, REF/external.py:212 import it at
        module level) -- the one function that needs it, o3d_knn (REF/helpers.py:136-146: exact kNN, self excluded,
        squared distances), is replaced in the exec'ed train.py functions by the same query on scipy's cKDTree;
  (ii)  /root/reference and the repository root are put on sys.path, so REF/helpers.py:5's
        `from diff_gaussian_rasterization import GaussianRasterizationSettings as Camera` resolves to THIS repository's
        drop-in package -- setup_camera() below therefore builds the product's settings type with the reference's code;
  (iii) the reference hard-codes device="cuda" / .cuda() (REF/helpers.py:70-93,105; REF/external.py:64,258-280): a small
        shim redirects those requests to the CPU for the duration of the calls.  No arithmetic is changed.
REF/helpers.py save_ply_splat (:40-66) needs `plyfile`, which is absent: the PLY writer stays pinned by its format
description only (tests/test_scene_io.py).

Run from the repository root:  python tests/golden/make_ref_fixtures.py   (needs /root/reference; the GPU box has none,
which is why the outputs are committed).  tests/test_ref_fixtures.py compares oracle/ and the product's host code with
the file, and re-runs this script when /root/reference exists to prove the file is reproducible.
"""
from __future__ import annotations

import contextlib
import importlib.util
import io
import os
import sys
import tempfile
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = os.environ.get("GSR_REFERENCE_ROOT", "/root/reference")
OUT = os.path.join(HERE, "ref_python_fixtures.npz")


def available() -> bool:
    return os.path.isfile(os.path.join(REF, "helpers.py")) and os.path.isfile(os.path.join(REF, "external.py"))


# ---------------------------------------------------------------------------------------------------------------------
# (iii) device shim
# ---------------------------------------------------------------------------------------------------------------------
@contextlib.contextmanager
def cuda_requests_go_to_cpu():
    names = ["zeros", "ones", "tensor", "zeros_like", "ones_like", "normal", "full", "empty", "arange"]
    saved = {n: getattr(torch, n) for n in names}
    saved_cuda = torch.Tensor.cuda
    saved_empty_cache = torch.cuda.empty_cache

    def strip(fn):
        def wrapped(*a, **k):
            if str(k.get("device", "")).startswith("cuda"):
                k = dict(k, device="cpu")
            return fn(*a, **k)
        return wrapped

    try:
        for n in names:
            setattr(torch, n, strip(saved[n]))
        torch.Tensor.cuda = lambda self, *a, **k: self
        torch.cuda.empty_cache = lambda: None
        yield
    finally:
        for n in names:
            setattr(torch, n, saved[n])
        torch.Tensor.cuda = saved_cuda
        torch.cuda.empty_cache = saved_empty_cache


def import_reference():
    """(i) + (ii): returns the reference's `helpers` and `external` modules, imported from /root/reference unmodified."""
    sys.modules.setdefault("open3d", types.ModuleType("open3d"))
    for p in (ROOT, REF):
        if p not in sys.path:
            sys.path.append(p)          # appended: tests/helpers.py keeps the bare name `helpers` for the test-suite
    mods = {}
    for name in ("helpers", "external"):
        spec = importlib.util.spec_from_file_location(f"ref_{name}", os.path.join(REF, f"{name}.py"))
        m = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(m)
        mods[name] = m
    return mods["helpers"], mods["external"]


def train_py_functions(ref_helpers, ref_external, names):
    """exec()s the named top-level functions of REF/train.py from their own source lines (the file as a whole has a
    syntax error at :284-285) in a namespace holding what REF/train.py:1-12 imports."""
    with open(os.path.join(REF, "train.py"), "r") as f:
        lines = f.read().split("\n")
    ns = {"torch": torch, "np": np, "o3d_knn": ckdtree_knn}
    for n in ("setup_camera", "l1_loss_v1", "l1_loss_v2", "weighted_l2_loss_v1", "weighted_l2_loss_v2", "quat_mult",
              "params2rendervar", "params2cpu", "save_params"):
        ns[n] = getattr(ref_helpers, n)
    for n in ("calc_ssim", "calc_psnr", "build_rotation", "densify", "update_params_and_optimizer"):
        ns[n] = getattr(ref_external, n)
    for name in names:
        start = next(i for i, l in enumerate(lines) if l.startswith(f"def {name}("))
        end = next(i for i in range(start + 1, len(lines)) if lines[i] and not lines[i][0].isspace())
        exec(compile("\n".join(lines[start:end]), f"REF/train.py:{start + 1}-{end}", "exec"), ns)
    return ns


def ckdtree_knn(pts, num_knn):
    """Stand-in for o3d_knn (REF/helpers.py:136-146): (squared distances [N,k], indices [N,k]) of the k nearest OTHER
    points, ascending."""
    from scipy.spatial import cKDTree
    pts = np.ascontiguousarray(pts, np.float64)
    d, i = cKDTree(pts).query(pts, k=num_knn + 1)
    return d[:, 1:] ** 2, i[:, 1:]


# ---------------------------------------------------------------------------------------------------------------------
# seeded inputs (shared with tests/test_ref_fixtures.py through the file itself: every input is stored next to its output)
# ---------------------------------------------------------------------------------------------------------------------
def generate() -> dict:
    H_, E_ = import_reference()
    out = {}
    g = torch.Generator().manual_seed(20260)
    f32 = lambda t: t.detach().cpu().numpy().astype(np.float32)

    with cuda_requests_go_to_cpu():
        # ---- image loss: REF/train.py:183 = 0.8 * l1_loss_v1 + 0.2 * (1 - calc_ssim), on [3,H,W] tensors ----
        im = torch.rand(3, 45, 61, generator=g).requires_grad_(True)
        tgt = (im.detach() + 0.15 * torch.randn(3, 45, 61, generator=g)).clamp(0, 1)
        ssim = E_.calc_ssim(im, tgt)
        l1 = H_.l1_loss_v1(im, tgt)
        loss = 0.8 * l1 + 0.2 * (1.0 - ssim)
        loss.backward()
        out.update(img_x=f32(im), img_y=f32(tgt), img_ssim=f32(ssim), img_l1=f32(l1), img_loss=f32(loss),
                   img_grad=f32(im.grad), img_ssim_same=f32(E_.calc_ssim(tgt, tgt)),
                   img_psnr=f32(E_.calc_psnr(im.detach(), tgt)))

        # ---- small helpers ----
        q1 = torch.randn(64, 4, generator=g)
        q2 = torch.randn(64, 4, generator=g)
        x = torch.randn(50, 20, 3, generator=g)
        y = torch.randn(50, 20, 3, generator=g)
        w = torch.rand(50, 20, generator=g)
        out.update(q1=f32(q1), q2=f32(q2), quat_mult=f32(H_.quat_mult(q1, q2)), build_rotation=f32(E_.build_rotation(q1)),
                   wl_x=f32(x), wl_y=f32(y), wl_w=f32(w),
                   weighted_l2_v1=f32(H_.weighted_l2_loss_v1(x[..., 0], y[..., 0], w)),
                   weighted_l2_v2=f32(H_.weighted_l2_loss_v2(x, y, w)),
                   l1_v1=f32(H_.l1_loss_v1(x, y)), l1_v2=f32(H_.l1_loss_v2(x, y)))

        # ---- setup_camera + params2rendervar: the reference's call contract (REF/train.py:142) ----
        wv, hv = 96, 64
        k = [[110.0, 0.0, 48.0 + 2.5], [0.0, 112.0, 32.0 - 1.75], [0.0, 0.0, 1.0]]
        c, s = np.cos(0.3), np.sin(0.3)
        w2c = np.array([[c, 0.0, s, 0.1], [0.0, 1.0, 0.0, -0.2], [-s, 0.0, c, 3.0], [0.0, 0.0, 0.0, 1.0]])
        cam = H_.setup_camera(wv, hv, k, w2c)
        assert type(cam).__name__ == "GaussianRasterizationSettings" and type(cam).__module__.startswith(
            "dynamic3dgaussians_b200"), "REF/helpers.py:5 must resolve to this repository's drop-in"
        out.update(cam_w=np.int32(wv), cam_h=np.int32(hv), cam_k=np.asarray(k, np.float64), cam_w2c=w2c,
                   cam_tanfovx=np.float64(cam.tanfovx), cam_tanfovy=np.float64(cam.tanfovy),
                   cam_c_x=np.float64(cam.c_x), cam_c_y=np.float64(cam.c_y), cam_bg=f32(cam.bg),
                   cam_scale_modifier=np.float64(cam.scale_modifier), cam_viewmatrix=f32(cam.viewmatrix),
                   cam_projmatrix=f32(cam.projmatrix), cam_campos=f32(cam.campos), cam_sh_degree=np.int32(cam.sh_degree),
                   cam_prefiltered=np.bool_(cam.prefiltered), cam_debug=np.bool_(cam.debug),
                   cam_confidence_shape=np.asarray(cam.confidence.shape, np.int64),
                   cam_fields=np.asarray(type(cam)._fields))
        P = 400
        pts = torch.rand(P, 3, generator=g) * 2 - 1
        params = {
            'means3D': pts, 'rgb_colors': torch.rand(P, 3, generator=g),
            'seg_colors': torch.stack(((torch.arange(P) % 3 != 0).float(), torch.zeros(P), (torch.arange(P) % 3 == 0).float()), -1),
            'unnorm_rotations': torch.randn(P, 4, generator=g), 'logit_opacities': torch.randn(P, 1, generator=g),
            'log_scales': torch.log(torch.rand(P, 3, generator=g) * 0.05 + 0.01),
            'cam_m': torch.zeros(4, 3), 'cam_c': torch.zeros(4, 3),
        }
        rv = H_.params2rendervar({kk: v.clone().requires_grad_(True) for kk, v in params.items()})
        out.update({f"scene_{kk}": f32(v) for kk, v in params.items()})
        out.update({f"rendervar_{kk}": f32(v) for kk, v in rv.items()})
        out["rendervar_keys"] = np.asarray(list(rv.keys()))
        assert rv['means2D'].requires_grad and not rv['means2D'].is_leaf

        # ---- REF/train.py functions, exec'ed from their own lines ----
        T_ = train_py_functions(H_, E_, ["initialize_optimizer", "initialize_per_timestep",
                                         "initialize_post_first_timestep"])
        tp = {kk: torch.nn.Parameter(v.clone().float().contiguous().requires_grad_(True)) for kk, v in params.items()}
        variables = {'max_2D_radius': torch.zeros(P), 'scene_radius': 2.75, 'means2D_gradient_accum': torch.zeros(P),
                     'denom': torch.zeros(P)}
        opt = T_["initialize_optimizer"](tp, variables)
        out["opt_names"] = np.asarray([gr['name'] for gr in opt.param_groups])
        out["opt_lrs"] = np.asarray([gr['lr'] for gr in opt.param_groups], np.float64)
        out["opt_eps"] = np.float64(opt.defaults['eps'])
        out["opt_scene_radius"] = np.float64(variables['scene_radius'])
        # one Adam step so that moments exist, then the t = 0 -> 1 hand-over
        for v in tp.values():
            v.grad = torch.randn(v.shape, generator=g) * 0.01
        opt.step()
        opt.zero_grad(set_to_none=True)
        variables = T_["initialize_post_first_timestep"](tp, variables, opt)
        out["post_lrs"] = np.asarray([gr['lr'] for gr in opt.param_groups], np.float64)
        for kk in ("neighbor_indices", "neighbor_weight", "neighbor_dist", "init_bg_pts", "init_bg_rot", "prev_pts", "prev_rot"):
            v = variables[kk]
            out[f"post_{kk}"] = v.numpy().astype(np.int64) if v.dtype == torch.int64 else f32(v)
        out.update({f"t0_{kk}": f32(tp[kk]) for kk in ("means3D", "unnorm_rotations", "rgb_colors", "seg_colors")})
        # t = 0 -> 1: the reference calls initialize_per_timestep right after initialize_post_first_timestep
        # (REF/train.py:416-417,445-446), and `prev_pts = params['means3D'].detach()` (:331) aliases the parameter anyway:
        # the first hand-over always sees zero velocity
        tp, variables = T_["initialize_per_timestep"](tp, variables, opt)
        out["step1_means3D"] = f32(tp['means3D'])
        out["step1_unnorm_rotations"] = f32(tp['unnorm_rotations'])
        out["step1_prev_offset"] = f32(variables["prev_offset"])
        # timestep 1 is optimised: the (new) parameters move in place, prev_pts keeps the old tensor
        with torch.no_grad():
            tp['means3D'].add_(0.01 * torch.randn(P, 3, generator=g))
            tp['unnorm_rotations'].add_(0.02 * torch.randn(P, 4, generator=g))
        out.update({f"t1in_{kk}": f32(tp[kk]) for kk in ("means3D", "unnorm_rotations", "rgb_colors")})
        tp, variables = T_["initialize_per_timestep"](tp, variables, opt)
        for kk in ("prev_inv_rot_fg", "prev_offset", "prev_pts", "prev_rot", "prev_col"):
            out[f"step_{kk}"] = f32(variables[kk])
        out["step_means3D"] = f32(tp['means3D'])
        out["step_unnorm_rotations"] = f32(tp['unnorm_rotations'])
        st = opt.state[tp['means3D']]
        out["step_means3D_exp_avg_absmax"] = f32(st["exp_avg"].abs().max())        # moments reset (external.py:148-149)

        # ---- the rigidity block of get_loss (REF/train.py:253-273), statement by statement from imported helpers ----
        with torch.no_grad():
            tp['means3D'].add_(0.004 * torch.randn(P, 3, generator=g))
            tp['unnorm_rotations'].add_(0.03 * torch.randn(P, 4, generator=g))
        rendervar = H_.params2rendervar(tp)
        is_fg = (tp['seg_colors'][:, 0] > 0.5).detach()
        fg_pts = rendervar['means3D'][is_fg]
        fg_rot = rendervar['rotations'][is_fg]
        rel_rot = H_.quat_mult(fg_rot, variables["prev_inv_rot_fg"])
        rot = E_.build_rotation(rel_rot)
        neighbor_pts = fg_pts[variables["neighbor_indices"]]
        curr_offset = neighbor_pts - fg_pts[:, None]
        curr_offset_in_prev_coord = (rot.transpose(2, 1)[:, None] @ curr_offset[:, :, :, None]).squeeze(-1)
        losses = {}
        losses['rigid'] = H_.weighted_l2_loss_v2(curr_offset_in_prev_coord, variables["prev_offset"], variables["neighbor_weight"])
        losses['rot'] = H_.weighted_l2_loss_v2(rel_rot[variables["neighbor_indices"]], rel_rot[:, None], variables["neighbor_weight"])
        curr_offset_mag = torch.sqrt((curr_offset ** 2).sum(-1) + 1e-20)
        losses['iso'] = H_.weighted_l2_loss_v1(curr_offset_mag, variables["neighbor_dist"], variables["neighbor_weight"])
        losses['floor'] = torch.clamp(fg_pts[:, 1], min=0).mean()
        bg_pts = rendervar['means3D'][~is_fg]
        bg_rot = rendervar['rotations'][~is_fg]
        losses['bg'] = H_.l1_loss_v2(bg_pts, variables["init_bg_pts"]) + H_.l1_loss_v2(bg_rot, variables["init_bg_rot"])
        losses['soft_col_cons'] = H_.l1_loss_v2(tp['rgb_colors'], variables["prev_col"])
        up = {'rigid': 4.0, 'rot': 4.0, 'iso': 2.0, 'floor': 2.0, 'bg': 20.0, 'soft_col_cons': 0.01}
        sum(up[kk] * v for kk, v in losses.items()).backward()
        out.update({f"dyn_loss_{kk}": f32(v) for kk, v in losses.items()})
        out["dyn_weights"] = np.asarray([up[kk] for kk in ('rigid', 'rot', 'iso', 'floor', 'bg', 'soft_col_cons')], np.float64)
        out["dyn_in_means3D"] = f32(tp['means3D'])
        out["dyn_in_unnorm_rotations"] = f32(tp['unnorm_rotations'])
        out["dyn_grad_means3D"] = f32(tp['means3D'].grad)
        out["dyn_grad_unnorm_rotations"] = f32(tp['unnorm_rotations'].grad)
        out["dyn_is_fg"] = is_fg.numpy()

        # ---- densify (REF/external.py:244-292) at the three schedule regimes ----
        for it in (600, 3000, 5000):
            out.update(_densify_case(E_, it, g))

        # ---- params2cpu / save_params (REF/helpers.py:149-168): the on-disk format ----
        t0 = {kk: torch.nn.Parameter(v.clone()) for kk, v in params.items()}
        seq = [H_.params2cpu(t0, True)]
        for t in (1, 2):
            t0['means3D'].data.add_(0.01 * t)
            seq.append(H_.params2cpu(t0, False))
        cwd = os.getcwd()
        with tempfile.TemporaryDirectory() as tmp:
            os.chdir(tmp)
            try:
                H_.save_params(seq, "seqname", "expname")
                with open(os.path.join(tmp, "output", "expname", "seqname", "params.npz"), "rb") as fh:
                    blob = fh.read()
            finally:
                os.chdir(cwd)
        z = np.load(io.BytesIO(blob))
        out["npz_keys"] = np.asarray(sorted(z.files))
        for kk in z.files:
            out[f"npz_{kk}"] = z[kk]
    return out


def _densify_case(E_, it: int, g: torch.Generator) -> dict:
    """One call of the reference's densify() at iteration `it` on a seeded state; the Gaussian noise torch.normal draws
    for the split (external.py:266) is recorded so that a restatement can be fed the same samples."""
    P, radius = 800, 3.0
    t = {
        'means3D': torch.randn(P, 3, generator=g),
        'rgb_colors': torch.rand(P, 3, generator=g),
        'seg_colors': torch.rand(P, 3, generator=g).round(),
        'unnorm_rotations': torch.randn(P, 4, generator=g),
        'logit_opacities': torch.randn(P, 1, generator=g) * 3.0,
        'log_scales': torch.log(torch.rand(P, 3, generator=g) * 0.08 + 1e-3),
    }
    params = {k: torch.nn.Parameter(v.clone()) for k, v in t.items()}
    params['cam_m'] = torch.nn.Parameter(torch.zeros(4, 3))
    params['cam_c'] = torch.nn.Parameter(torch.zeros(4, 3))
    opt = torch.optim.Adam([{'params': [v], 'name': k, 'lr': 1e-3} for k, v in params.items()], lr=0.0, eps=1e-15)
    for v in params.values():
        v.grad = torch.randn(v.shape, generator=g)
    opt.step()
    keys = list(t.keys())
    pre = {k: params[k].detach().clone() for k in keys}
    pre_m = {k: (opt.state[params[k]]['exp_avg'].clone(), opt.state[params[k]]['exp_avg_sq'].clone()) for k in keys}
    denom = torch.randint(0, 5, (P,), generator=g).float()
    accum = torch.rand(P, generator=g) * 0.0005 * denom
    seen = torch.rand(P, generator=g) < 0.7
    m2d = torch.zeros(P, 3, requires_grad=True)
    m2d.grad = torch.randn(P, 3, generator=g) * 1e-4
    variables = {'means2D_gradient_accum': accum.clone(), 'denom': denom.clone(), 'scene_radius': radius,
                 'max_2D_radius': torch.rand(P, generator=g) * 30, 'seen': seen, 'means2D': m2d}
    recorded = {}
    real_normal = torch.normal

    def recording_normal(*a, **k):
        torch.manual_seed(it)
        r = real_normal(*a, **k)
        recorded['std'] = k['std'].clone()
        recorded['z'] = r.clone()
        return r

    torch.normal = recording_normal
    try:
        with contextlib.redirect_stdout(io.StringIO()):
            params2, variables2 = E_.densify(dict(params), variables, opt, it)
    finally:
        torch.normal = real_normal
    f32 = lambda v: v.detach().numpy().astype(np.float32)
    o = {f"dens{it}_in_{k}": f32(pre[k]) for k in keys}
    o.update({f"dens{it}_in_m1_{k}": f32(pre_m[k][0]) for k in keys})
    o.update({f"dens{it}_in_m2_{k}": f32(pre_m[k][1]) for k in keys})
    o[f"dens{it}_in_accum"] = f32(accum)
    o[f"dens{it}_in_denom"] = f32(denom)
    o[f"dens{it}_in_seen"] = seen.numpy()
    o[f"dens{it}_in_means2D_grad"] = f32(m2d.grad)
    o[f"dens{it}_noise"] = f32(recorded['z'])
    o[f"dens{it}_noise_std"] = f32(recorded['std'])
    for k in keys:
        o[f"dens{it}_out_{k}"] = f32(params2[k])
        stt = opt.state[params2[k]]
        o[f"dens{it}_out_m1_{k}"] = f32(stt['exp_avg'])
        o[f"dens{it}_out_m2_{k}"] = f32(stt['exp_avg_sq'])
    for k in ('means2D_gradient_accum', 'denom', 'max_2D_radius'):
        o[f"dens{it}_out_{k}"] = f32(variables2[k])
    o[f"dens{it}_radius"] = np.float64(radius)
    return o


def main():
    if not available():
        print(f"reference not found under {REF}: nothing generated")
        return 1
    out = generate()
    np.savez_compressed(OUT, **out)
    print(f"wrote {OUT}: {len(out)} arrays, {os.path.getsize(OUT) / 1024:.0f} KiB")
    return 0


if __name__ == "__main__":
    sys.exit(main())
