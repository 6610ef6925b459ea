"""TEST INFRASTRUCTURE (oracle): CPU restatement of the reference's image loss in plain PyTorch.

Follows REF/helpers.py l1_loss_v1 (`torch.abs(x - y).mean()`), REF/external.py:90-97 (gaussian / create_window: the
1-D window is built with math.exp in double precision, stored as float32, normalised, and the 2-D window is its outer
product) and REF/external.py:112-133 (_ssim: five grouped conv2d with zero padding window//2, c1 = 0.01^2, c2 = 0.03^2,
mean over everything), and REF/train.py:183 for the 0.8 / 0.2 mix.  Pinned against the reference's own calc_ssim /
l1_loss_v1 run on seeded images (tests/golden/ref_python_fixtures.npz, tests/test_ref_fixtures.py, <= 1e-6 incl. the
gradient).  Only tests/ and bench.py may import this file."""
from math import exp

import torch
import torch.nn.functional as func


def gaussian(window_size: int, sigma: float) -> torch.Tensor:
    g = torch.Tensor([exp(-(x - window_size // 2) ** 2 / float(2 * sigma ** 2)) for x in range(window_size)])
    return g / g.sum()


def create_window(window_size: int, channel: int) -> torch.Tensor:
    w1 = gaussian(window_size, 1.5).unsqueeze(1)
    w2 = w1.mm(w1.t()).float().unsqueeze(0).unsqueeze(0)
    return w2.expand(channel, 1, window_size, window_size).contiguous()


def calc_ssim(img1: torch.Tensor, img2: torch.Tensor, window_size: int = 11) -> torch.Tensor:
    channel = img1.size(-3)
    window = create_window(window_size, channel).type_as(img1)
    pad = window_size // 2
    mu1 = func.conv2d(img1, window, padding=pad, groups=channel)
    mu2 = func.conv2d(img2, window, padding=pad, groups=channel)
    mu1_sq, mu2_sq, mu1_mu2 = mu1.pow(2), mu2.pow(2), mu1 * mu2
    sigma1_sq = func.conv2d(img1 * img1, window, padding=pad, groups=channel) - mu1_sq
    sigma2_sq = func.conv2d(img2 * img2, window, padding=pad, groups=channel) - mu2_sq
    sigma12 = func.conv2d(img1 * img2, window, padding=pad, groups=channel) - mu1_mu2
    c1, c2 = 0.01 ** 2, 0.03 ** 2
    ssim_map = ((2 * mu1_mu2 + c1) * (2 * sigma12 + c2)) / ((mu1_sq + mu2_sq + c1) * (sigma1_sq + sigma2_sq + c2))
    return ssim_map.mean()


def l1_loss_v1(x: torch.Tensor, y: torch.Tensor) -> torch.Tensor:
    return torch.abs(x - y).mean()


def image_loss(im: torch.Tensor, target: torch.Tensor, w_l1: float = 0.8, w_ssim: float = 0.2) -> torch.Tensor:
    return w_l1 * l1_loss_v1(im, target) + w_ssim * (1.0 - calc_ssim(im, target))
