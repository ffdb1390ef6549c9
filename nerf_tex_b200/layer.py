"""Input encodings (network/layer.py).  In this package both encodings are computed inside the fused MLP
kernel; the classes below only carry the configuration, plus a standalone callable for users who want the
features themselves."""

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
    """network/layer.py:25-41: inputs [mean(3), diagonal covariance(3)] ->
    concat([sin(y), sin(y + pi/2)]) * exp(-y_var / 2) with y = mean * 2^k, y_var = cov * 4^k, frequency-major."""

    def __init__(self, n_freq_bands: int) -> None:
        self.n_freq_bands = int(n_freq_bands)

    def out_dim(self, d: int = 6) -> int:
        return 6 * self.n_freq_bands

    def __call__(self, inputs: torch.Tensor) -> torch.Tensor:
        import math
        freq = 2.0 ** torch.arange(self.n_freq_bands, dtype=torch.float32, device=inputs.device)
        y = (inputs[..., None, :3] * freq[:, None]).reshape(-1, 3 * self.n_freq_bands)
        y_var = (inputs[..., None, 3:] * freq[:, None] ** 2).reshape(-1, 3 * self.n_freq_bands)
        x = torch.cat([y, y + 0.5 * math.pi], -1)
        return torch.sin(x) * torch.exp(-0.5 * torch.cat([y_var, y_var], -1))
