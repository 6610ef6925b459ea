"""Fused L1 + SSIM loss: the oracle against a float64 finite difference (CPU), the CUDA kernels against the oracle."""
import os
import sys

import pytest
import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from helpers import rel_err  # noqa: E402
from oracle import eager_image_loss as OL  # noqa: E402


def _pair(shape, seed):
    g = torch.Generator().manual_seed(seed)
    tgt = torch.rand(*shape, generator=g)
    img = (tgt + 0.2 * torch.randn(*shape, generator=g)).clamp(0, 1)
    return img, tgt


def test_oracle_known_answers_and_finite_difference():
    img, tgt = _pair((3, 24, 31), 0)
    assert abs(float(OL.calc_ssim(tgt, tgt)) - 1.0) < 1e-6           # identical images: SSIM 1, loss 0
    assert abs(float(OL.image_loss(tgt, tgt))) < 1e-6
    w = OL.gaussian(11, 1.5)
    assert abs(float(w.sum()) - 1.0) < 1e-6 and float(w[5]) == float(w.max()) and torch.equal(w, w.flip(0))
    x = img.double().requires_grad_(True)
    OL.image_loss(x, tgt.double()).backward()
    eps = 1e-6
    for idx in [(0, 0, 0), (1, 12, 17), (2, 23, 30), (0, 5, 5)]:
        xp, xm = img.double().clone(), img.double().clone()
        xp[idx] += eps
        xm[idx] -= eps
        fd = (OL.image_loss(xp, tgt.double()) - OL.image_loss(xm, tgt.double())) / (2 * eps)
        assert abs(float(fd) - float(x.grad[idx])) < 1e-7


@pytest.mark.gpu
@pytest.mark.parametrize("shape", [(3, 360, 640), (2, 3, 37, 53), (1, 5, 16)])
def test_cuda_loss_and_gradient_match_oracle(shape):
    from dynamic3dgaussians_b200.image_loss import l1_ssim_loss
    img, tgt = _pair(shape, 1)
    x = img.clone().requires_grad_(True)
    ref = OL.image_loss(x, tgt)
    ref.backward()
    xc = img.cuda().requires_grad_(True)
    loss, parts = l1_ssim_loss(xc, tgt.cuda(), return_parts=True)
    (3.0 * loss).backward()
    assert abs(float(loss) - float(ref)) <= 1e-5
    assert abs(float(parts[1]) - float(OL.l1_loss_v1(img, tgt))) <= 1e-5
    assert abs(float(parts[2]) - float(OL.calc_ssim(img, tgt))) <= 1e-5
    assert rel_err(xc.grad, 3.0 * x.grad) <= 1e-4
    with pytest.raises(ValueError):
        l1_ssim_loss(img, tgt)                                       # CPU tensors: no fallback
