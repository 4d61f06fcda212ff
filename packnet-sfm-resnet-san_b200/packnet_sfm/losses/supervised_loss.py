"""Supervised loss for inverse depth maps (reference: losses/supervised_loss.py).

Provided: `SupervisedLoss` with the loss functions the reference's `get_loss_func` returns for the methods
ending in 'l1', 'mse', 'silog' and 'ssi-silog' ('sparse-' and dense forms) -- what `train_resnet_san_kitti.yaml`,
`train_packnet_san_kitti.yaml`, `train_packnet_san_ddad.yaml` (`supervised_method: 'sparse-silog'`) and the NCDB yamls
(`'sparse-ssi-silog'`) run every step.  Each scale is two reduction kernels forward and one kernel backward through the C ABI
(`psfm_supervised_loss[_backward]`).  Not provided: 'berhu' (the reference's own BerHuLoss raises, :65) and the
other SSI / dual-head families (separate loss modules of the fork)."""
import torch.nn as nn

from packnet_sfm.losses.loss_base import LossBase, ProgressiveScaling
from packnet_sfm.utils.image import match_scales


class _ElementLoss(nn.Module):
    """nn.L1Loss / nn.MSELoss / SilogLoss stand-in: forward(pred, gt) -> scalar mean-type loss."""
    method = None

    def __init__(self, ratio2=0.85):
        super().__init__()
        self.ratio2 = ratio2

    def forward(self, pred, gt, mask=None, sparse=False):
        from packnet_sfm.block_ops import supervised_loss_op
        loss, self.state = supervised_loss_op(pred, gt, self.method, sparse=sparse, mask=mask, ratio2=self.ratio2)
        return loss


class L1Loss(_ElementLoss):
    method = 'l1'


class MSELoss(_ElementLoss):
    method = 'mse'


class SilogLoss(_ElementLoss):
    """sqrt(E[l^2] - ratio2 * E[l]^2), l = log pred - log gt (reference :67-82; `ratio` is unused there too)."""
    method = 'silog'

    def __init__(self, ratio=10, ratio2=0.85):
        super().__init__(ratio2)
        self.ratio = ratio


def get_loss_func(supervised_method, **kwargs):
    """Loss function of a supervised method name (reference :84-238)."""
    if supervised_method.endswith('l1'):
        return L1Loss()
    elif supervised_method.endswith('mse'):
        return MSELoss()
    elif supervised_method.endswith('ssi-silog'):
        from packnet_sfm.losses.ssi_silog_loss import SSISilogLoss
        return SSISilogLoss(min_depth=kwargs.get('min_depth', None), max_depth=kwargs.get('max_depth', None),
                            ssi_weight=kwargs.get('ssi_weight', 0.7), silog_weight=kwargs.get('silog_weight', 0.3),
                            alpha=kwargs.get('alpha', 0.85), silog_ratio=kwargs.get('silog_ratio', 10),
                            silog_ratio2=kwargs.get('silog_ratio2', 0.85),
                            gradient_weight=kwargs.get('gradient_weight', 0.0),
                            gradient_scales=kwargs.get('gradient_scales', 4))
    elif 'ssi' in supervised_method:
        raise NotImplementedError('the other SSI loss families are outside this library (method %s)' % supervised_method)
    elif supervised_method.endswith('silog'):
        return SilogLoss()
    elif supervised_method.endswith('berhu'):
        raise NotImplementedError("'berhu' raises inside the reference's own BerHuLoss and is not provided")
    else:
        raise ValueError('Unknown supervised loss {}'.format(supervised_method))


class SupervisedLoss(LossBase):
    """Supervised loss for inverse depth maps (reference :241-437).

    supervised_method: '[sparse-]l1' | '[sparse-]mse' | '[sparse-]silog'; supervised_num_scales: scales used;
    progressive_scaling: training percentage for progressive scaling (0.0 disables)."""

    def __init__(self, supervised_method='sparse-l1', supervised_num_scales=4, progressive_scaling=0.0, **kwargs):
        super().__init__()
        self.loss_func = get_loss_func(supervised_method, **kwargs)
        self.supervised_method = supervised_method
        self.n = supervised_num_scales
        self.progressive_scaling = ProgressiveScaling(progressive_scaling, self.n)

    @property
    def logs(self):
        return {'supervised_num_scales': self.n}

    def calculate_loss(self, inv_depths, gt_inv_depths, masks=None):
        """Average over the scales of the loss function on (masked) predictions (reference :292-417).
        Sparse methods: pixels with gt <= 0 (or outside `masks[i]`) are filled with 1e-6 in both maps and still
        count in the mean, as in the reference.  Per-scale 's{i}/loss', 's{i}/valid_px', 's{i}/valid_ratio' are
        logged as device scalars (the reference reads them back with .item())."""
        num_scales = self.n
        sparse = self.supervised_method.startswith('sparse')
        total_loss = 0.0
        for i in range(num_scales):
            mask = None
            if sparse and masks is not None and i < len(masks) and masks[i] is not None:
                mask = masks[i]
                if tuple(mask.shape[-2:]) != tuple(gt_inv_depths[i].shape[-2:]):
                    from packnet_sfm.ops import resize_op
                    mask = resize_op(mask.to(gt_inv_depths[i].dtype), gt_inv_depths[i].shape[-2:], mode='nearest')
            loss_i = self.loss_func(inv_depths[i], gt_inv_depths[i], mask=mask, sparse=sparse)
            if sparse:
                state = self.loss_func.state
                self.add_metric('s%d/loss' % i, loss_i)
                self.add_metric('s%d/valid_px' % i, state[4])
                self.add_metric('s%d/valid_ratio' % i, state[4] / state[6])
                if isinstance(self.loss_func, LossBase):          # e.g. SSISilogLoss: copy its parts (:358-360)
                    for k, v in self.loss_func.metrics.items():
                        self.add_metric('s%d/%s' % (i, k), v)
            total_loss = total_loss + loss_i
        return total_loss / float(num_scales)

    def forward(self, inv_depths, gt_inv_depth, return_logs=False, progress=0.0, masks=None):
        """inv_depths: list of [B,1,Hs,Ws]; gt_inv_depth: [B,1,H,W] (0 where unknown).  Returns {'loss': [1],
        'metrics': {...}} (reference :420-437)."""
        self._progress = progress
        self.n = self.progressive_scaling(progress)
        gt_inv_depths = match_scales(gt_inv_depth, inv_depths, self.n, mode='nearest', align_corners=None)
        loss = self.calculate_loss(inv_depths, gt_inv_depths, masks=masks)
        self.add_metric('supervised_loss', loss)
        return {'loss': loss.unsqueeze(0), 'metrics': self.metrics}
