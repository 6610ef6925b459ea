"""TEST INFRASTRUCTURE (oracle): the reference's densification restated pass by pass on plain tensors.

Follows REF/external.py:244-292 literally -- clone (cat), split (cat, shrink), remove the split parents, prune by
opacity / size -- each as its own torch.cat / boolean-mask round like `cat_params_to_optimizer` (:159-183) and
`remove_points` (:186-208) do, with the Adam moments carried along (zeros for new rows).  The Gaussian noise of the split
(:264-266) is injected by the caller so that results are comparable.  Pinned against the reference's own densify() run
on a seeded scene with its Adam optimizer (tests/golden/ref_python_fixtures.npz, tests/test_ref_fixtures.py: same rows,
order, values and moments).  Only tests/ may import this file."""
import torch

KEYS = ('means3D', 'rgb_colors', 'seg_colors', 'unnorm_rotations', 'logit_opacities', 'log_scales')


def densify_step(tensors: dict, moments: dict, grads: torch.Tensor, scene_radius: float, i: int, noise_fn):
    """tensors / moments: {name: [P, d]} (moments: (exp_avg, exp_avg_sq) per name).  Returns (tensors, moments, src,
    kind) where src / kind give every surviving row's original row and 0 / 1 / 2 = original / clone / child."""
    t = {k: v.clone() for k, v in tensors.items()}
    m = {k: (a.clone(), b.clone()) for k, (a, b) in moments.items()}
    P = grads.shape[0]
    src = torch.arange(P)
    kind = torch.zeros(P, dtype=torch.int8)
    grad_thresh = 0.0002

    def cat(new, new_src, new_kind):
        nonlocal src, kind
        for k in KEYS:
            t[k] = torch.cat((t[k], new[k]), dim=0)
            m[k] = (torch.cat((m[k][0], torch.zeros_like(new[k])), dim=0),
                    torch.cat((m[k][1], torch.zeros_like(new[k])), dim=0))
        src = torch.cat((src, new_src))
        kind = torch.cat((kind, new_kind))

    def remove(mask):
        nonlocal src, kind
        keep = ~mask
        for k in KEYS:
            t[k] = t[k][keep]
            m[k] = (m[k][0][keep], m[k][1][keep])
        src, kind = src[keep], kind[keep]

    to_clone = torch.logical_and(grads >= grad_thresh, torch.max(torch.exp(t['log_scales']), dim=1).values <= 0.01 * scene_radius)
    cat({k: t[k][to_clone] for k in KEYS}, src[to_clone], torch.ones(int(to_clone.sum()), dtype=torch.int8))
    num_pts = t['means3D'].shape[0]
    padded = torch.zeros(num_pts)
    padded[:P] = grads
    to_split = torch.logical_and(padded >= grad_thresh, torch.max(torch.exp(t['log_scales']), dim=1).values > 0.01 * scene_radius)
    n = 2
    new = {k: t[k][to_split].repeat(n, 1) for k in KEYS}
    stds = torch.exp(t['log_scales'])[to_split].repeat(n, 1)
    new_src = src[to_split].repeat(n)
    samples = noise_fn(new_src, stds)
    q = torch.nn.functional.normalize(t['unnorm_rotations'][to_split])
    r, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    rot = torch.zeros(q.shape[0], 3, 3)
    rot[:, 0, 0] = 1 - 2 * (y * y + z * z); rot[:, 0, 1] = 2 * (x * y - r * z); rot[:, 0, 2] = 2 * (x * z + r * y)
    rot[:, 1, 0] = 2 * (x * y + r * z); rot[:, 1, 1] = 1 - 2 * (x * x + z * z); rot[:, 1, 2] = 2 * (y * z - r * x)
    rot[:, 2, 0] = 2 * (x * z - r * y); rot[:, 2, 1] = 2 * (y * z + r * x); rot[:, 2, 2] = 1 - 2 * (x * x + y * y)
    rots = rot.repeat(n, 1, 1)
    new['means3D'] = new['means3D'] + torch.bmm(rots, samples.unsqueeze(-1)).squeeze(-1)
    new['log_scales'] = torch.log(torch.exp(new['log_scales']) / (0.8 * n))
    cat(new, new_src, torch.full((new_src.numel(),), 2, dtype=torch.int8))
    remove(torch.cat((to_split, torch.zeros(n * int(to_split.sum()), dtype=torch.bool))))
    remove_threshold = 0.25 if i == 5000 else 0.005
    to_remove = (torch.sigmoid(t['logit_opacities']) < remove_threshold).squeeze(-1)
    if i >= 3000:
        to_remove = torch.logical_or(to_remove, torch.exp(t['log_scales']).max(dim=1).values > 0.1 * scene_radius)
    remove(to_remove)
    return t, m, src, kind
