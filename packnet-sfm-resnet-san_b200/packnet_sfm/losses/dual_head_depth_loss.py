"""Integer-fractional dual-head depth loss (reference: losses/dual_head_depth_loss.py:23-201).

Over the pixels with min_depth < gt < max_depth:
  integer_loss     = mean |integer_pred - floor(gt) / max_depth|
  fractional_loss  = mean |fractional_pred - (gt - floor(gt))|
  consistency_loss = mean |integer_pred * max_depth + fractional_pred - gt|
  loss = integer_weight * integer_loss + fractional_weight * fractional_loss + consistency_weight * consistency_loss
Two reduction kernels forward, one kernel backward (`psfm_dual_head_loss[_backward]`)."""
from packnet_sfm.losses.loss_base import LossBase


class DualHeadDepthLoss(LossBase):
    def __init__(self, max_depth=15.0, integer_weight=1.0, fractional_weight=10.0, consistency_weight=0.5, min_depth=0.5,
                 **kwargs):
        super().__init__()
        assert max_depth > min_depth, f"max_depth ({max_depth}) must be > min_depth ({min_depth})"
        assert max_depth > 0, f"max_depth must be positive, got {max_depth}"
        assert min_depth >= 0, f"min_depth must be non-negative, got {min_depth}"
        assert integer_weight >= 0, f"integer_weight must be non-negative, got {integer_weight}"
        assert fractional_weight > 0, f"fractional_weight must be positive, got {fractional_weight}"
        assert consistency_weight >= 0, f"consistency_weight must be non-negative, got {consistency_weight}"
        self.max_depth = max_depth
        self.min_depth = min_depth
        self.integer_weight = integer_weight
        self.fractional_weight = fractional_weight
        self.consistency_weight = consistency_weight

    def forward(self, outputs, depth_gt, return_logs=False, progress=0.0):
        """outputs: dict with ("integer", 0) and ("fractional", 0) -> [B,1,H,W] sigmoid outputs; depth_gt [B,1,H',W']
        (resized with nearest interpolation to the prediction's size).  Returns {'loss', 'integer_loss',
        'fractional_loss', 'consistency_loss'}; with no valid pixel every entry is 0."""
        from packnet_sfm.block_ops import dual_head_loss_op
        integer_pred, fractional_pred = outputs[("integer", 0)], outputs[("fractional", 0)]
        if depth_gt.shape[-2:] != integer_pred.shape[-2:]:
            from packnet_sfm.ops import resize_op
            depth_gt = resize_op(depth_gt, integer_pred.shape[-2:], mode='nearest')
        loss, parts, state = dual_head_loss_op(integer_pred, fractional_pred, depth_gt, self.max_depth, self.min_depth,
                                               self.integer_weight, self.fractional_weight, self.consistency_weight)
        if return_logs:
            self.add_metric('integer_loss', parts[0])
            self.add_metric('fractional_loss', parts[1])
            self.add_metric('consistency_loss', parts[2])
            self.add_metric('total_loss', loss)
            valid = state[3].clamp(min=1.0)
            self.add_metric('mean_depth_error', parts[2])                       # the same mean |depth_pred - gt|
            self.add_metric('integer_accuracy', state[4] / valid)               # within 1 m
            self.add_metric('fractional_rmse', (state[5] / valid).sqrt())
            # (the reference also logs a median depth error; not computed here)
        return {'loss': loss, 'integer_loss': parts[0], 'fractional_loss': parts[1], 'consistency_loss': parts[2]}
