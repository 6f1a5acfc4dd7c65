"""CIFAR10 DAWNBench data augmentation on the device, batched.

The reference prepares the data with per-sample NumPy transforms chosen once per epoch on the host
(``examples/dist/CIFAR10-dawndist/core.py:50-62`` pad / normalise / transpose, ``:69-110`` ``Crop(32, 32)``,
``FlipLR()``, ``Cutout(8, 8)`` driven by ``Transform.set_random_choices``).  Here the whole (reflect-padded) training
set lives on the GPU and one call draws the choices and produces an augmented batch with three indexing kernels --
nothing is staged through the host inside the training loop.
"""
from __future__ import annotations

import torch

CIFAR10_MEAN = (0.4914, 0.4822, 0.4465)
CIFAR10_STD = (0.2471, 0.2435, 0.2616)


def normalise(x_uint8_nhwc, mean=CIFAR10_MEAN, std=CIFAR10_STD):
    """uint8 NHWC -> float32 NCHW, (x - 255 mean) / (255 std)  (reference ``core.py:50-54,61-62``)."""
    x = x_uint8_nhwc.float()
    m = torch.tensor(mean, device=x.device).view(1, 1, 1, 3) * 255
    s = torch.tensor(std, device=x.device).view(1, 1, 1, 3) * 255
    return ((x - m) / s).permute(0, 3, 1, 2).contiguous()


def pad_reflect(x_nchw, border=4):
    """Reflect padding of the training images (reference ``core.py:57-58``)."""
    return torch.nn.functional.pad(x_nchw, (border,) * 4, mode="reflect")


class DawnAugment:
    """``Crop(h, w)`` at a random offset, ``FlipLR`` with probability 1/2, ``Cutout(ch, cw)`` at a random offset --
    the reference's ``train_transforms`` (``dawn.py:94``) -- for a whole batch at once."""

    def __init__(self, crop=(32, 32), cutout=(8, 8), flip=True, generator=None):
        self.crop, self.cutout, self.flip, self.generator = crop, cutout, flip, generator

    def __call__(self, padded):
        n, c, H, W = padded.shape
        h, w = self.crop
        dev, gen = padded.device, self.generator
        y0 = torch.randint(0, H - h + 1, (n,), device=dev, generator=gen)
        x0 = torch.randint(0, W - w + 1, (n,), device=dev, generator=gen)
        rows = (y0.view(n, 1) + torch.arange(h, device=dev).view(1, h))                  # [n, h]
        cols = (x0.view(n, 1) + torch.arange(w, device=dev).view(1, w))                  # [n, w]
        if self.flip:
            flip = torch.rand(n, device=dev, generator=gen) < 0.5
            cols = torch.where(flip.view(n, 1), cols.flip(1), cols)
        idx_n = torch.arange(n, device=dev).view(n, 1, 1, 1)
        idx_c = torch.arange(c, device=dev).view(1, c, 1, 1)
        out = padded[idx_n, idx_c, rows.view(n, 1, h, 1), cols.view(n, 1, 1, w)]
        if self.cutout is not None:
            ch, cw = self.cutout
            cy = torch.randint(0, h - ch + 1, (n,), device=dev, generator=gen).view(n, 1, 1)
            cx = torch.randint(0, w - cw + 1, (n,), device=dev, generator=gen).view(n, 1, 1)
            yy = torch.arange(h, device=dev).view(1, h, 1)
            xx = torch.arange(w, device=dev).view(1, 1, w)
            hole = (yy >= cy) & (yy < cy + ch) & (xx >= cx) & (xx < cx + cw)             # [n, h, w]
            out = out.masked_fill(hole.view(n, 1, h, w), 0.0)
        return out
