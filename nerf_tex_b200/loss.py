"""Loss functions (network/loss.py).  The generic forms are torch expressions (any callable can be plugged in through
`loss_fn`); AlphaLoss over the shipped smape / mse pair on CUDA tensors takes one kernel (ntx_alpha_loss) for the
value and the gradient with respect to the renderer's outputs."""

from __future__ import annotations

import torch

from . import util


def mse(y_true: torch.Tensor, y_pred: torch.Tensor) -> torch.Tensor:
    """network/loss.py:51-54."""
    return torch.mean((y_true - y_pred) ** 2)


def smape(y_true: torch.Tensor, y_pred: torch.Tensor, eps: float = 1e-2) -> torch.Tensor:
    """network/loss.py:56-58."""
    return torch.mean(torch.abs(y_true - y_pred) / (y_true + y_pred + eps))


class _FusedAlphaLoss(torch.autograd.Function):
    """loss and d loss / d (color_pred, alpha_pred[, coarse pair]) from one launch of ntx_alpha_loss."""

    @staticmethod
    def forward(ctx, color_true, alpha_true, color_pred, alpha_pred, color_coarse, alpha_coarse, kinds, gamma, filt, hard):
        from . import _lib
        from ._lib import check, ptr, stream_ptr
        f = lambda x: None if x is None else x.detach().reshape(-1).contiguous().float()
        ct, at, cp, ap, cc, ac = (f(x) for x in (color_true, alpha_true, color_pred, alpha_pred, color_coarse, alpha_coarse))
        n = at.shape[0]
        loss = torch.empty(1, dtype=torch.float32, device=cp.device)
        d_cp, d_ap = torch.empty_like(cp), torch.empty_like(ap)
        d_cc = torch.empty_like(cc) if cc is not None else None
        d_ac = torch.empty_like(ac) if ac is not None else None
        with torch.cuda.device(cp.device):
            check(_lib.load().ntx_alpha_loss(ptr(ct), ptr(at), ptr(cp), ptr(ap), ptr(cc), ptr(ac), n, kinds[0], kinds[1],
                                             float(gamma), int(filt), int(hard), ptr(loss), ptr(d_cp), ptr(d_ap), ptr(d_cc),
                                             ptr(d_ac), stream_ptr()))
        ctx.grads = (d_cp.view_as(color_pred), d_ap.view_as(alpha_pred),
                     None if d_cc is None else d_cc.view_as(color_coarse), None if d_ac is None else d_ac.view_as(alpha_coarse))
        return loss.reshape(())

    @staticmethod
    def backward(ctx, g):
        d_cp, d_ap, d_cc, d_ac = ctx.grads
        return (None, None, g * d_cp, g * d_ap, None if d_cc is None else g * d_cc, None if d_ac is None else g * d_ac,
                None, None, None, None)


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
        kind = {mse: 0, smape: 1}
        self._kinds = (kind[self.loss], kind[self.alpha_loss]) if self.loss in kind and self.alpha_loss in kind else None

    def __call__(self, color_true, alpha_true, color_pred, alpha_pred, color_pred_coarse=None,
                 alpha_pred_coarse=None, **kwargs):
        if self._kinds is not None and color_pred.is_cuda and (color_pred_coarse is None) == (alpha_pred_coarse is None):
            return _FusedAlphaLoss.apply(color_true, alpha_true, color_pred, alpha_pred, color_pred_coarse, alpha_pred_coarse,
                                         self._kinds, self.gamma, self.filter_color_loss, self.use_hard_mask)
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
