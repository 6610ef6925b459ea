"""TEST INFRASTRUCTURE ONLY.  PyTorch-eager CPU restatement of the rigidity / rotation / isometry losses, line by line
from REF/train.py:253-273 with the helpers of REF/helpers.py:118-133 (weighted_l2_loss_v1/_v2, quat_mult) and
REF/external.py:61-78 (build_rotation); gradients by autograd.  The reference ships no test or fixture for this code, so
it is pinned against the reference's OWN Python run on seeded inputs: tests/golden/ref_python_fixtures.npz (written by
tests/golden/make_ref_fixtures.py, which imports REF/helpers.py and REF/external.py unmodified and exec's the functions of
REF/train.py) -- values and gradients <= 1e-6 / 1e-5 (tests/test_ref_fixtures.py) -- and with float64 finite differences."""
import torch


def weighted_l2_loss_v1(x, y, w):                       # REF/helpers.py:118-119
    return torch.sqrt(((x - y) ** 2) * w + 1e-20).mean()


def weighted_l2_loss_v2(x, y, w):                       # REF/helpers.py:122-123
    return torch.sqrt(((x - y) ** 2).sum(-1) * w + 1e-20).mean()


def quat_mult(q1, q2):                                  # REF/helpers.py:126-133
    w1, x1, y1, z1 = q1.T
    w2, x2, y2, z2 = q2.T
    w = w1 * w2 - x1 * x2 - y1 * y2 - z1 * z2
    x = w1 * x2 + x1 * w2 + y1 * z2 - z1 * y2
    y = w1 * y2 - x1 * z2 + y1 * w2 + z1 * x2
    z = w1 * z2 + x1 * y2 - y1 * x2 + z1 * w2
    return torch.stack([w, x, y, z]).T


def build_rotation(q):                                  # REF/external.py:61-78
    norm = torch.sqrt(q[:, 0] * q[:, 0] + q[:, 1] * q[:, 1] + q[:, 2] * q[:, 2] + q[:, 3] * q[:, 3])
    q = q / norm[:, None]
    rot = torch.zeros((q.size(0), 3, 3), dtype=q.dtype)
    r, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    rot[:, 0, 0] = 1 - 2 * (y * y + z * z)
    rot[:, 0, 1] = 2 * (x * y - r * z)
    rot[:, 0, 2] = 2 * (x * z + r * y)
    rot[:, 1, 0] = 2 * (x * y + r * z)
    rot[:, 1, 1] = 1 - 2 * (x * x + z * z)
    rot[:, 1, 2] = 2 * (y * z - r * x)
    rot[:, 2, 0] = 2 * (x * z - r * y)
    rot[:, 2, 1] = 2 * (y * z + r * x)
    rot[:, 2, 2] = 1 - 2 * (x * x + y * y)
    return rot


def rigidity_losses(fg_pts, fg_rot, prev_inv_rot_fg, neighbor_indices, neighbor_weight, prev_offset, neighbor_dist):
    """(rigid, rot, iso) exactly as REF/train.py:259-273."""
    rel_rot = quat_mult(fg_rot, prev_inv_rot_fg)
    rot = build_rotation(rel_rot)
    neighbor_pts = fg_pts[neighbor_indices]
    curr_offset = neighbor_pts - fg_pts[:, None]
    curr_offset_in_prev_coord = (rot.transpose(2, 1)[:, None] @ curr_offset[:, :, :, None]).squeeze(-1)
    rigid = weighted_l2_loss_v2(curr_offset_in_prev_coord, prev_offset, neighbor_weight)
    rot_l = weighted_l2_loss_v2(rel_rot[neighbor_indices], rel_rot[:, None], neighbor_weight)
    curr_offset_mag = torch.sqrt((curr_offset ** 2).sum(-1) + 1e-20)
    iso = weighted_l2_loss_v1(curr_offset_mag, neighbor_dist, neighbor_weight)
    return torch.stack([rigid, rot_l, iso])
