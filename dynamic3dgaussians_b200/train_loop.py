"""A `train.py`-shaped optimisation loop driving the B200 rasterizer end to end.

Same structure as the reference's flat scripts -- params / variables dicts, `params2rendervar`, `get_loss`, per-timestep
initialisation, one Adam group per parameter (REF/train.py:89-136,138-291,294-336,392-443; REF/helpers.py:98-155) -- with
the hot path replaced:

  * colour and fg/bg segmentation are rendered in ONE fused pass (`rasterize_views(extra_colors=seg_colors)`) instead of
    two rasterizer calls (REF/train.py:142,249); the densification statistic still comes from the COLOUR render's
    screen-space gradient alone, as in the reference (REF/train.py:245, REF/README.md:89): the fused backward returns that
    share separately (`means2D_color`, gsr_backward's dL_dmeans2D_color_only);
  * the image losses are the fused L1 + SSIM kernels (`image_loss.l1_ssim_loss`, REF/train.py:183);
  * the rigidity / rotation / isometry terms are the fused kNN-loss kernels (`dyn_losses.rigidity_losses`,
    REF/train.py:253-273) on a 20-NN graph built on the GPU (`knn.knn`, REF/train.py:316-326 uses Open3D on the CPU);
  * densification (REF/external.py:244-292) builds ONE gather list for clone + split + prune and applies it once to every
    parameter and Adam moment (`densify.py`) instead of four rounds of torch.cat / boolean indexing + empty_cache();
  * nothing in a step reads a device value back to the host (no `.item()` logging inside the loop).

The fork's own scripts cannot run as written (hard-coded data paths REF/train.py:26,39,168; a syntax error in the loss
weights REF/train.py:284-285; open3d / wandb dependencies), so the loss weights and learning rates default to the
upstream Dynamic3DGaussians values the fork was derived from and can be overridden."""
from __future__ import annotations

import os
from typing import Dict, List, Optional, Sequence

import torch

from . import rasterizer as RZ
from . import train_ops
from .densify import accumulate_mean2d_gradient, densify, replace_param
from .image_loss import l1_ssim_loss

# The small ops around the rasterizer (activations, camera colour correction, densification statistics) run as one CUDA
# kernel each (train_ops.py) instead of ~40 eager kernels per iteration; False keeps the eager torch ops (the comparison
# arm of the tests).
FUSED_GLUE = True
# The segmentation image loss runs on a side stream next to the colour correction and the colour image loss; autograd
# replays the same fork in the backward, and a captured iteration keeps it as parallel graph branches (+2 % it/s).
# Forking the render-independent terms of a dynamic timestep (rigidity / background / colour consistency) the same
# way measured 2-3 % SLOWER than leaving them in line before the render, so they stay on the main stream.
OVERLAP_LOSSES = os.environ.get("GSR_OVERLAP_LOSSES", "1") != "0"
_SIDE_STREAMS: Dict[torch.device, torch.cuda.Stream] = {}


def _side_stream(dev: torch.device) -> torch.cuda.Stream:
    s = _SIDE_STREAMS.get(dev)
    if s is None:
        s = _SIDE_STREAMS[dev] = torch.cuda.Stream(device=dev)
    return s

# REF/train.py:120-129 as written has means3D 0.0000016 * scene_radius, rgb 0.000025, rotations 0.0 (the fork froze
# most of them for its ego-camera experiments); the upstream values are the defaults here
LEARNING_RATES = {
    'means3D': 0.00016, 'rgb_colors': 0.0025, 'seg_colors': 0.0, 'unnorm_rotations': 0.001,
    'logit_opacities': 0.05, 'log_scales': 0.001, 'cam_m': 1e-4, 'cam_c': 1e-4,
}
# REF/train.py:281-283 (the fork zeroes most of them and does not parse); upstream weights
LOSS_WEIGHTS = {'im': 1.0, 'seg': 3.0, 'rigid': 4.0, 'rot': 4.0, 'iso': 2.0, 'floor': 2.0, 'bg': 20.0,
                'soft_col_cons': 0.01}


def initialize_params(points: torch.Tensor, rgb: torch.Tensor, seg: torch.Tensor, num_cams: int, scene_radius: float):
    """REF/train.py:89-116.  points [P,3], rgb [P,3], seg [P] (1 = foreground) on the GPU."""
    from .knn import distCUDA2_dist
    dev = points.device
    P = points.shape[0]
    mean3_sq_dist = distCUDA2_dist(points.contiguous()).clamp(min=1e-7)           # o3d_knn(.., 3).mean(-1).clip(...)
    rot = torch.zeros(P, 4, device=dev)
    rot[:, 0] = 1.0
    raw = {
        'means3D': points, 'rgb_colors': rgb,
        'seg_colors': torch.stack((seg, torch.zeros_like(seg), 1 - seg), -1),
        'unnorm_rotations': rot,
        'logit_opacities': torch.zeros(P, 1, device=dev),
        'log_scales': torch.log(torch.sqrt(mean3_sq_dist))[:, None].repeat(1, 3),
        'cam_m': torch.zeros(num_cams, 3, device=dev), 'cam_c': torch.zeros(num_cams, 3, device=dev),
    }
    params = {k: torch.nn.Parameter(v.float().contiguous().clone().requires_grad_(True)) for k, v in raw.items()}
    variables = {'max_2D_radius': torch.zeros(P, device=dev), 'scene_radius': float(scene_radius),
                 'means2D_gradient_accum': torch.zeros(P, device=dev), 'denom': torch.zeros(P, device=dev)}
    return params, variables


def initialize_optimizer(params: Dict[str, torch.nn.Parameter], variables: dict, lrs: Optional[dict] = None,
                         fused: bool = True):
    """REF/train.py:119-135: one Adam group per tensor, eps 1e-15, means3D scaled by the scene radius.  fused=True
    (default) is optim.FusedAdam -- one kernel for all groups, CUDA-graph capturable; fused=False is torch.optim.Adam
    in capturable mode (the comparison arm of the tests)."""
    lrs = dict(LEARNING_RATES if lrs is None else lrs)
    lrs['means3D'] = lrs['means3D'] * variables['scene_radius']
    groups = [{'params': [v], 'name': k, 'lr': lrs[k]} for k, v in params.items()]
    if fused:
        from .optim import FusedAdam
        return FusedAdam(groups, lr=0.0, eps=1e-15)
    return torch.optim.Adam(groups, lr=0.0, eps=1e-15, capturable=True)


def params2rendervar(params, fused: Optional[bool] = None):
    """REF/helpers.py:98-107."""
    fused = FUSED_GLUE if fused is None else fused
    if fused:
        rotations, opacities, scales = train_ops.activations(params['unnorm_rotations'], params['logit_opacities'],
                                                             params['log_scales'])
        # the rasterizer never reads means2D (it is the handle the screen-space gradient comes back on): a fresh leaf
        means2D = torch.zeros_like(params['means3D']).requires_grad_(True)
    else:
        rotations = torch.nn.functional.normalize(params['unnorm_rotations'])
        opacities = torch.sigmoid(params['logit_opacities'])
        scales = torch.exp(params['log_scales'])
        means2D = torch.zeros_like(params['means3D'], requires_grad=True) + 0
    return {
        'means3D': params['means3D'],
        'colors_precomp': params['rgb_colors'],
        'rotations': rotations,
        'opacities': opacities,
        'scales': scales,
        'means2D': means2D,
    }


def get_loss(params, curr_data, variables, is_initial_timestep, weights: Optional[dict] = None,
             capacity: Optional[int] = None, fused: Optional[bool] = None, overlap: Optional[bool] = None,
             state_out: Optional[list] = None):
    """REF/train.py:138-291.  curr_data: {'cam': GaussianRasterizationSettings, 'im': [3,H,W], 'seg': [3,H,W], 'id'}.
    `state_out` (append-only list) receives the render's RasterState: with a fixed `capacity` nothing in the step looks
    at the overflow flag, so whoever fixes the capacity checks it afterwards (rasterizer.read_flags)."""
    weights = LOSS_WEIGHTS if weights is None else weights
    fused = FUSED_GLUE if fused is None else fused
    losses = {}
    # eager iterations are bound by the host (~1 ms of Python per step): the fork's events only cost there
    overlap = False if overlap is None else overlap       # GraphedTrainer asks for it (OVERLAP_LOSSES)
    dev = params['means3D'].device
    cur = torch.cuda.current_stream(dev)
    side = _side_stream(dev) if overlap else cur
    rv = params2rendervar(params, fused)
    # handle for the colour render's screen-space gradient (the fused pass also blends the segmentation colours, whose
    # cotangent must not leak into the densification statistic: REF/train.py:245)
    # -- only the initial timestep densifies, the later ones skip the extra reduction
    means2D_color = torch.zeros_like(params['means3D']).requires_grad_(True) if is_initial_timestep else None

    if not is_initial_timestep:
        from .dyn_losses import rigidity_losses
        with torch.cuda.stream(cur):
            # index lists instead of the reference's boolean masks (REF/train.py:254-257): static shapes, no host sync,
            # so the iteration can be captured into a CUDA graph
            fg_idx, bg_idx = variables['fg_idx'], variables['bg_idx']
            fg_pts = rv['means3D'].index_select(0, fg_idx)
            fg_rot = rv['rotations'].index_select(0, fg_idx)
            dyn = {}
            dyn['rigid'], dyn['rot'], dyn['iso'] = rigidity_losses(fg_pts, fg_rot, variables['graph'])
            dyn['floor'] = torch.clamp(fg_pts[:, 1], min=0).mean()
            bg_pts = rv['means3D'].index_select(0, bg_idx)
            bg_rot = rv['rotations'].index_select(0, bg_idx)
            dyn['bg'] = (torch.abs(bg_pts - variables['init_bg_pts']).sum(-1)).mean() + \
                        (torch.abs(bg_rot - variables['init_bg_rot']).sum(-1)).mean()
            dyn['soft_col_cons'] = (torch.abs(params['rgb_colors'] - variables['prev_col']).sum(-1)).mean()

    color, seg, depth, alpha, radii = RZ.rasterize_views(
        [curr_data['cam']], rv['means3D'], rv['opacities'], means2D=None, colors_precomp=rv['colors_precomp'],
        scales=rv['scales'], rotations=rv['rotations'], extra_colors=params['seg_colors'], capacity=capacity,
        state_out=state_out, means2D_color=means2D_color)
    im, seg, radius = color[0], seg[0], radii[0]
    side.wait_stream(cur)
    with torch.cuda.stream(side):
        loss_seg = l1_ssim_loss(seg, curr_data['seg'])
    cid = curr_data['id']
    if fused:
        im = train_ops.cam_affine(im, params['cam_m'], params['cam_c'], cid)
    else:
        im = torch.exp(params['cam_m'][cid])[:, None, None] * im + params['cam_c'][cid][:, None, None]
    losses['im'] = l1_ssim_loss(im, curr_data['im'])
    if means2D_color is not None:
        variables['means2D'] = means2D_color             # "Gradient only accum from colour render for densification"
    losses['seg'] = loss_seg
    if not is_initial_timestep:
        losses.update(dyn)
    cur.wait_stream(side)

    loss = None
    for k, v in losses.items():
        term = v if weights[k] == 1.0 else weights[k] * v
        loss = term if loss is None else loss + term
    # in place (REF/train.py:288-289), so that the statistics also accumulate when the step is replayed from a CUDA graph
    if fused:
        seen = train_ops.update_max_radius(radius, variables['max_2D_radius'])
    else:
        seen = radius > 0
        variables['max_2D_radius'].copy_(torch.where(seen, torch.maximum(radius.float(), variables['max_2D_radius']),
                                                     variables['max_2D_radius']))
    variables['seen'] = seen
    return loss, variables, losses


def initialize_per_timestep(params, variables, optimizer):
    """REF/train.py:294-313: constant-velocity initialisation of the next timestep + the rigidity references."""
    from .dyn_losses import refresh_graph
    pts = params['means3D']
    rot = torch.nn.functional.normalize(params['unnorm_rotations'])
    new_pts = pts + (pts - variables['prev_pts'])
    new_rot = torch.nn.functional.normalize(rot + (rot - variables['prev_rot']))
    is_fg = variables['is_fg']
    variables['graph'] = refresh_graph(variables['graph'], pts[is_fg].detach(), rot[is_fg].detach())
    variables['prev_pts'] = pts.detach()
    variables['prev_rot'] = rot.detach()
    variables['prev_col'] = params['rgb_colors'].detach().clone()
    for k, v in {'means3D': new_pts, 'unnorm_rotations': new_rot}.items():
        replace_param(params, optimizer, k, v.detach().clone(), reset_state=True)
    return params, variables


def initialize_post_first_timestep(params, variables, optimizer, num_knn: int = 20):
    """REF/train.py:316-336: 20-NN graph of the foreground, background anchors, freeze opacity / scale / camera terms."""
    from .dyn_losses import build_neighbor_graph
    is_fg = (params['seg_colors'][:, 0] > 0.5).detach()
    variables['is_fg'] = is_fg
    variables['fg_idx'] = torch.nonzero(is_fg).squeeze(1)
    variables['bg_idx'] = torch.nonzero(~is_fg).squeeze(1)
    rot = torch.nn.functional.normalize(params['unnorm_rotations']).detach()
    variables['graph'] = build_neighbor_graph(params['means3D'][is_fg].detach(), rot[is_fg], num_knn)
    variables['init_bg_pts'] = params['means3D'][~is_fg].detach().clone()
    variables['init_bg_rot'] = rot[~is_fg].clone()
    variables['prev_pts'] = params['means3D'].detach().clone()
    variables['prev_rot'] = rot.clone()
    for group in optimizer.param_groups:
        if group['name'] in ('logit_opacities', 'log_scales', 'cam_m', 'cam_c'):
            group['lr'] = 0.0
    return variables


def train_timestep(params, variables, optimizer, dataset: Sequence[dict], num_iters: int, is_initial_timestep: bool,
                   generator: Optional[torch.Generator] = None, weights: Optional[dict] = None,
                   do_densify: bool = True, capacity: Optional[int] = None, start_iter: int = 0) -> List[torch.Tensor]:
    """The inner loop of REF/train.py:422-438: one random view per step (get_batch), loss, backward, densify at t = 0,
    Adam step.  Returns the per-step loss tensors (still on the device: nothing synchronises inside the loop)."""
    order: List[int] = []
    history = []
    for i in range(start_iter, start_iter + num_iters):
        if not order:                                           # get_batch: a shuffled pass over the cameras
            order = torch.randperm(len(dataset), generator=generator).tolist()
        curr = dataset[order.pop()]
        loss, variables, _ = get_loss(params, curr, variables, is_initial_timestep, weights, capacity)
        loss.backward()
        with torch.no_grad():
            if is_initial_timestep and do_densify:
                params, variables = densify(params, variables, optimizer, i, generator=generator)
            optimizer.step()
            optimizer.zero_grad(set_to_none=True)
        history.append(loss.detach())
    return history


class GraphedTrainer:
    """One CUDA graph per training camera, each holding a WHOLE iteration -- render, losses, backward, statistics, Adam
    step -- so that an iteration is a single launch (the eager loop spends ~2 ms of Python per iteration on ~1 ms of GPU
    work).  The graphs share one memory pool (they never run concurrently).  Valid while the parameter tensors keep
    their identity and size: build it after the densification phase (or rebuild it after every densify round) with an
    optimizer created by initialize_optimizer() and a fixed per-view instance `capacity`.

    `capacity` is baked into every captured iteration: when the Gaussians move or grow so that a view produces more
    (Gaussian, tile) instances than that, the binning drops the surplus (the render is then wrong, silently).  Size it
    with headroom (bench.py uses 1.25x the largest view's count at build time) and call `overflowed()` every few hundred
    steps and at the end of a timestep -- one 16-byte read per camera; on True rebuild the trainer with more capacity.
    Building the trainer does not train: the warm-up iterations run on the live tensors, but the parameters, the Adam
    state and the densification statistics are restored afterwards.  Learning rates are read when a graph is captured;
    `step()` refuses to replay after a `param_group['lr']` changed (rebuild instead)."""

    def __init__(self, params, variables, optimizer, dataset: Sequence[dict], is_initial_timestep: bool, capacity: int,
                 weights: Optional[dict] = None, accumulate_stats: bool = True, warmup: int = 3):
        dev = params['means3D'].device
        self.params, self.variables, self.optimizer, self.dataset = params, variables, optimizer, list(dataset)
        self.graphs: List[torch.cuda.CUDAGraph] = []
        self.losses: List[torch.Tensor] = []
        # [camera][num_rendered, overflow, prefilter violation, -]: copied out of the render's workspace at the end of every
        # captured iteration.  The workspaces themselves live in the graphs' SHARED memory pool, where another camera's
        # replay reuses them, so they cannot be read back later.
        self.counters = torch.zeros(len(self.dataset), 4, dtype=torch.int32, device=dev)
        self.order: List[int] = []
        self.capacity = int(capacity)

        def iteration(curr, state_out=None):
            loss, _, _ = get_loss(params, curr, variables, is_initial_timestep, weights, capacity,
                                  overlap=OVERLAP_LOSSES, state_out=state_out)
            loss.backward()
            with torch.no_grad():
                if accumulate_stats and is_initial_timestep:
                    _accumulate_inplace(variables)
                optimizer.step()
            return loss.detach()

        # the warm-up must leave no trace: snapshot what an iteration updates in place
        stat_keys = [k for k in ('means2D_gradient_accum', 'denom', 'max_2D_radius') if torch.is_tensor(variables.get(k))]
        snap_params = {k: v.detach().clone() for k, v in params.items()}
        snap_stats = {k: variables[k].clone() for k in stat_keys}
        snap_state = {k: {n: t.clone() for n, t in optimizer.state[v].items() if torch.is_tensor(t)}
                      for k, v in params.items() if v in optimizer.state}
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):
            for i in range(max(1, warmup)):
                optimizer.zero_grad(set_to_none=True)
                iteration(self.dataset[i % len(self.dataset)])
        torch.cuda.current_stream(dev).wait_stream(side)
        with torch.no_grad():
            for k, v in params.items():
                v.copy_(snap_params[k])
                st = optimizer.state.get(v)
                if st is None:
                    continue
                for n, t in st.items():
                    if torch.is_tensor(t):
                        if k in snap_state and n in snap_state[k]:
                            t.copy_(snap_state[k][n])
                        else:
                            t.zero_()                 # state created by the warm-up itself (first ever step)
            for k in stat_keys:
                variables[k].copy_(snap_stats[k])
        torch.cuda.synchronize(dev)
        self._lrs = [float(g['lr']) for g in optimizer.param_groups]
        pool = None
        if hasattr(torch.autograd.graph, "set_warn_on_accumulate_grad_stream_mismatch"):
            # the side-stream branches hand gradients to AccumulateGrad nodes of the capture stream on purpose
            torch.autograd.graph.set_warn_on_accumulate_grad_stream_mismatch(False)
        for curr in self.dataset:
            optimizer.zero_grad(set_to_none=True)
            g = torch.cuda.CUDAGraph()
            st_out: list = []
            with torch.cuda.graph(g, pool=pool):
                loss = iteration(curr, st_out)
                self.counters[len(self.graphs)].copy_(st_out[0].geom[:16].view(torch.int32))
            pool = g.pool() if pool is None else pool
            self.graphs.append(g)
            self.losses.append(loss)

    def step(self, generator: Optional[torch.Generator] = None) -> torch.Tensor:
        """One optimizer iteration on the next camera of a shuffled pass (get_batch); returns that graph's loss tensor
        (overwritten by its next replay)."""
        if not self.order:
            self.order = torch.randperm(len(self.dataset), generator=generator).tolist()
            if [float(g['lr']) for g in self.optimizer.param_groups] != self._lrs:      # once per pass over the cameras
                raise RuntimeError("a learning rate changed after the iterations were captured (they are baked into the "
                                   "graphs): rebuild the GraphedTrainer")
        i = self.order.pop()
        self.graphs[i].replay()
        return self.losses[i]

    def overflowed(self) -> bool:
        """True when any camera's most recent replay needed more instances than `capacity` (its render dropped the
        surplus).  Synchronises; reads 16 bytes per camera."""
        return bool(self.counters[:, 1].any().item())

    def max_num_rendered(self) -> int:
        """Largest per-view instance count of the most recent replays (how much of `capacity` is in use)."""
        return int(self.counters[:, 0].max().item())


def _accumulate_inplace(variables: dict, fused: Optional[bool] = None):
    """accumulate_mean2d_gradient (REF/external.py:136-140) with in-place updates, for graph replay."""
    seen = variables['seen']
    if FUSED_GLUE if fused is None else fused:
        train_ops.accumulate_mean2d_gradient_(seen, variables['means2D'].grad, variables['means2D_gradient_accum'],
                                              variables['denom'])
        return
    norm = torch.norm(variables['means2D'].grad[:, :2], dim=-1)
    variables['means2D_gradient_accum'].add_(torch.where(seen, norm, torch.zeros_like(norm)))
    variables['denom'].add_(seen.to(variables['denom'].dtype))
