"""MS-SSIM as the reference's CURLLoss uses it (reference metric.py:75-214, `MSSSIMMetric`).

This is the one part of the loss that stays on the library path (SURVEY.md section 8, row f3:
"MS-SSIM stays on cuDNN"): Gaussian-window statistics via `conv2d` and `avg_pool2d`.  Same class
name, constructor and result as the reference; differences in form only: the five window
statistics of one scale come from ONE grouped convolution over a 5-plane stack instead of five
calls, and nothing is forced onto `.cuda()` (the reference hard-codes it, metric.py:116,150-157) --
tensors stay on the device they arrive on.
"""
import math

import torch
import torch.nn as nn
import torch.nn.functional as F


class MSSSIMMetric(nn.Module):
    def __init__(self, window_size=11, num_channel=3):
        super().__init__()
        self.msssim_weights = nn.Parameter(torch.tensor([0.0448, 0.2856, 0.3001, 0.2363, 0.1333]),
                                           requires_grad=False)                       # metric.py:78
        self.levels = self.msssim_weights.numel()
        self.window_size = window_size
        self.num_channel = num_channel
        self.gaussian_window = MSSSIMMetric.create_window(window_size, num_channel)
        # the reference keeps the window as a plain CPU attribute and copies it to the device on every call
        # (metric.py:131); here the device copy is cached per (device, dtype): no per-call H2D copy, and the
        # criterion can be captured in a CUDA graph.  Not a buffer, so the state dict is the reference's.
        self._window_cache = {}

    @staticmethod
    def gaussian(window_size, sigma):
        """Normalised 1-D Gaussian taps, fp32 (metric.py:106-117)."""
        taps = torch.tensor([math.exp(-(i - window_size // 2) ** 2 / float(2 * sigma ** 2))
                             for i in range(window_size)], dtype=torch.float32)
        return taps / taps.sum()

    @staticmethod
    def create_window(window_size, num_channel):
        """[C, 1, k, k] window: outer product of the taps, sigma 1.5 (metric.py:88-103)."""
        g = MSSSIMMetric.gaussian(window_size, 1.5).unsqueeze(1)
        return g.mm(g.t()).float().expand(num_channel, 1, window_size, window_size).contiguous()

    def _stats(self, img1, img2):
        """mu1, mu2, E[x^2], E[y^2], E[xy] under the window, one grouped conv."""
        C = self.num_channel
        key = (img1.device, img1.dtype)
        win = self._window_cache.get(key)
        if win is None:
            win = self.gaussian_window.to(device=img1.device, dtype=img1.dtype).repeat(5, 1, 1, 1)
            self._window_cache[key] = win
        planes = torch.cat((img1, img2, img1 * img1, img2 * img2, img1 * img2), dim=1)
        s = F.conv2d(planes, win, padding=self.window_size // 2, groups=5 * C)
        return s[:, 0:C], s[:, C:2 * C], s[:, 2 * C:3 * C], s[:, 3 * C:4 * C], s[:, 4 * C:5 * C]

    def compute_ssim(self, img1, img2):
        """(mean SSIM, mean contrast-structure) per image (metric.py:119-160)."""
        mu1, mu2, e11, e22, e12 = self._stats(img1, img2)
        mu1_sq, mu2_sq, mu1_mu2 = mu1.pow(2), mu2.pow(2), mu1 * mu2
        sigma1_sq, sigma2_sq, sigma12 = e11 - mu1_sq, e22 - mu2_sq, e12 - mu1_mu2
        C1, C2 = 0.01 ** 2, 0.03 ** 2
        v1 = 2.0 * sigma12 + C2
        v2 = sigma1_sq + sigma2_sq + C2
        ssim_map = ((2 * mu1_mu2 + C1) * (2 * sigma12 + C2)) / ((mu1_sq + mu2_sq + C1) * v2)
        return ssim_map.mean(dim=(1, 2, 3)), torch.mean(v1 / v2, dim=(1, 2, 3))

    def compute_msssim(self, img1, img2):
        """Per-image MS-SSIM over 5 dyadic scales (metric.py:162-211), including the reference's
        (x + 1) / 2 normalisation and its product form."""
        if img1.shape[2] != img2.shape[2]:
            img1 = img1.transpose(2, 3)
        if img1.shape != img2.shape:
            raise RuntimeError("Input images must have the same shape (%s vs. %s)." % (img1.shape, img2.shape))
        if img1.ndim != 4:
            raise RuntimeError("Input images must have four dimensions, not %d" % img1.ndim)
        ssims, mcs = [], []
        for _ in range(self.levels):
            s, cs = self.compute_ssim(img1, img2)
            ssims.append(s)
            mcs.append(cs)
            img1 = F.avg_pool2d(img1, (2, 2))
            img2 = F.avg_pool2d(img2, (2, 2))
        ssims = (torch.stack(ssims, dim=1) + 1) / 2
        mcs = (torch.stack(mcs, dim=1) + 1) / 2
        w = self.msssim_weights.reshape(1, -1)
        terms = (mcs ** w)[:, :-1] * (ssims ** w)[:, -1].reshape(-1, 1)
        # the reference's torch.prod(terms, dim=1) (metric.py:210), written out: prod's backward inspects its
        # input for zeros on the host (a device synchronisation, and not capturable in a CUDA graph)
        out = terms[:, 0]
        for k in range(1, terms.shape[1]):
            out = out * terms[:, k]
        return out

    def forward(self, img1, img2):
        return self.compute_msssim(img1, img2)
