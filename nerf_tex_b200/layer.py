"""Input encodings (network/layer.py).  In this package the Fourier features are computed inside
the fused MLP kernel; the class below only carries the configuration, plus a standalone
callable for users who want the features themselves."""

from __future__ import annotations

import torch


class FourierFeatures:
    """network/layer.py:8-23: concat([x] + [sin(2^k x), cos(2^k x) for k < n], -1)."""

    def __init__(self, n_freq_bands: int) -> None:
        self.n_freq_bands = int(n_freq_bands)

    def out_dim(self, d: int) -> int:
        return d * (1 + 2 * self.n_freq_bands)

    def __call__(self, inputs: torch.Tensor) -> torch.Tensor:
        feats = [inputs]
        for k in range(self.n_freq_bands):
            feats.append(torch.sin((2.0 ** k) * inputs))
            feats.append(torch.cos((2.0 ** k) * inputs))
        return torch.cat(feats, -1)


class IntegratedPositionalEncoding:
    """network/layer.py:25-41 — configuration holder; the mip renderers are not built yet."""

    def __init__(self, n_freq_bands: int) -> None:
        self.n_freq_bands = int(n_freq_bands)
