"""Loss functions (network/loss.py) in torch; tiny elementwise work around the hot path."""

from __future__ import annotations

import torch

from . import util


def mse(y_true: torch.Tensor, y_pred: torch.Tensor) -> torch.Tensor:
    """network/loss.py:51-54."""
    return torch.mean((y_true - y_pred) ** 2)


def smape(y_true: torch.Tensor, y_pred: torch.Tensor, eps: float = 1e-2) -> torch.Tensor:
    """network/loss.py:56-58."""
    return torch.mean(torch.abs(y_true - y_pred) / (y_true + y_pred + eps))


class NerfLoss:
    """network/loss.py:6-18."""

    def __init__(self, loss_fn: str = "network.loss.mse") -> None:
        self.loss = util.get_attr_from_path(loss_fn)

    def __call__(self, color_true, color_pred, color_pred_coarse=None, **kwargs):
        loss = self.loss(color_true, color_pred)
        if color_pred_coarse is not None:
            loss = loss + self.loss(color_true, color_pred_coarse)
        return loss


class AlphaLoss:
    """network/loss.py:20-49."""

    def __init__(self, loss_fn: str = "network.loss.mse", alpha_loss_fn: str = None, gamma: float = 1,
                 filter_color_loss: bool = True, use_hard_mask: bool = True) -> None:
        self.loss = util.get_attr_from_path(loss_fn)
        self.alpha_loss = self.loss if alpha_loss_fn is None else util.get_attr_from_path(alpha_loss_fn)
        self.gamma = gamma
        self.filter_color_loss = filter_color_loss
        self.use_hard_mask = use_hard_mask

    def __call__(self, color_true, alpha_true, color_pred, alpha_pred, color_pred_coarse=None,
                 alpha_pred_coarse=None, **kwargs):
        alpha_mask = None
        if self.filter_color_loss:
            alpha_mask = (alpha_true[..., None] > 0).float() if self.use_hard_mask else alpha_true[..., None]
            color_true = color_true * alpha_mask
            color_pred = color_pred * alpha_mask
        loss = self.loss(color_true, color_pred)
        loss = loss + self.gamma * self.alpha_loss(alpha_true, alpha_pred)
        if color_pred_coarse is not None:
            if self.filter_color_loss:
                color_pred_coarse = color_pred_coarse * alpha_mask
            loss = loss + self.loss(color_true, color_pred_coarse)
            loss = loss + self.gamma * self.alpha_loss(alpha_true, alpha_pred_coarse)
        return loss
