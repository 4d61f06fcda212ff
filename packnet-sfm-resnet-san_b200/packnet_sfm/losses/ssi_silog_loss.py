"""Scale-shift-invariant + Silog supervised loss (reference: losses/ssi_silog_loss.py:53-429).

loss = ssi_weight * ssi + silog_weight * silog over the valid pixels:
  ssi   = var(d) + alpha * mean(d)^2,  d = pred - gt on inverse depths (:176-197)
  silog = sqrt(|mean(l^2) - silog_ratio2 * mean(l)^2| + 1e-8),  l = log P - log G on depths 1 / max(., 1e-6) clamped to
          [min_depth or 1e-3, max_depth or 100] (:218-262)
0 with fewer than 100 valid pixels (:398-400).  One pair of reduction kernels + one gradient kernel through
`psfm_supervised_loss[_backward]`.  With `gradient_weight > 0` (train_resnet_san_ncdb_dual_head_640x384.yaml: 0.2) the
multi-scale Sobel gradient term (:116-174) is added through `psfm_sobel_gradient_loss[_backward]`.  The reference's
diagnostic statistics (quantiles, out-of-range fractions) and its NaN guards are not reproduced -- the parts of the loss
are logged."""
from typing import Optional

from packnet_sfm.losses.loss_base import LossBase


class SSISilogLoss(LossBase):
    def __init__(self, alpha=0.85, silog_ratio=10, silog_ratio2=0.85, ssi_weight=0.7, silog_weight=0.3,
                 gradient_weight=0.0, gradient_scales=4,
                 min_depth: Optional[float] = None, max_depth: Optional[float] = None):
        super().__init__()
        self.alpha = alpha
        self.silog_ratio = silog_ratio
        self.silog_ratio2 = silog_ratio2
        self.ssi_weight = ssi_weight
        self.silog_weight = silog_weight
        self.gradient_weight = gradient_weight
        self.gradient_scales = gradient_scales
        self.min_depth = min_depth
        self.max_depth = max_depth
        self.state = None

    def set_depth_range(self, min_depth: float, max_depth: float):
        """Optionally set depth clamp range after construction."""
        self.min_depth = float(min_depth)
        self.max_depth = float(max_depth)

    def forward(self, pred_inv_depth, gt_inv_depth, mask=None, sparse=True):
        """pred_inv_depth, gt_inv_depth: [B,1,H,W]; mask: optional valid-pixel mask, intersected with gt > 0 (the
        reference's callers build it that way, losses/supervised_loss.py:314-331).  Returns the scalar loss (:366-429)."""
        from packnet_sfm.block_ops import supervised_loss_op
        cmin = 1e-3 if self.min_depth is None else float(self.min_depth)
        cmax = 100.0 if self.max_depth is None else float(self.max_depth)
        loss, st = supervised_loss_op(pred_inv_depth, gt_inv_depth, 'ssi-silog', sparse=True, mask=mask,
                                      ratio2=self.silog_ratio2, alpha=self.alpha, ssi_weight=self.ssi_weight,
                                      silog_weight=self.silog_weight, clamp_min=cmin, clamp_max=cmax)
        self.state = st
        gradient_loss = 0.0
        if self.gradient_weight > 0:
            from packnet_sfm.block_ops import sobel_gradient_loss_op
            gradient_loss, self.gradient_state = sobel_gradient_loss_op(pred_inv_depth, gt_inv_depth, mask=mask,
                                                                        scales=self.gradient_scales)
            # fewer than 100 valid pixels: the whole loss is the constant 0 (:398-400); a NaN ssi / silog part: the
            # constant 1 (:403-407; state[15] flags both); a NaN gradient term is dropped (:410-411).  All decided on the
            # device (torch.where, not a product: 0 * NaN is NaN)
            import torch
            term = self.gradient_weight * gradient_loss
            keep = (st[4] >= 100) & (st[15] == 0) & torch.isfinite(term)
            loss = loss + torch.where(keep, term, torch.zeros_like(term)).to(loss.dtype)
        self.add_metric('ssi_mean', st[10])
        self.add_metric('ssi_var', st[11])
        self.add_metric('silog1', st[12])
        self.add_metric('silog2', st[13])
        self.add_metric('silog_var', (st[12] - st[13]).abs())
        self.add_metric('ssi_component', st[8])
        self.add_metric('silog_component', st[9])
        self.add_metric('gradient_component', gradient_loss)
        self.add_metric('ssi_weight_used', self.ssi_weight)
        self.add_metric('silog_weight_used', self.silog_weight)
        self.add_metric('gradient_weight_used', self.gradient_weight)
        return loss
