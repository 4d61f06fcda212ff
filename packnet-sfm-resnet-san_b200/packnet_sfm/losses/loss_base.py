"""Loss base classes (reference: losses/loss_base.py)."""
import numpy as np
import torch
import torch.nn as nn

from packnet_sfm.utils.types import is_list


class ProgressiveScaling:
    """Drops one scale each time the training progress passes a multiple of
    `progressive_scaling`; 0.0 disables it (reference :10-49)."""

    def __init__(self, progressive_scaling, num_scales=4):
        self.num_scales = num_scales
        if progressive_scaling > 0.0:
            self.progressive_scaling = np.float32(
                [progressive_scaling * (i + 1) for i in range(num_scales - 1)] + [1.0])
        else:
            self.progressive_scaling = progressive_scaling

    def __call__(self, progress):
        # Parity note: the reference tests `is_list(self.progressive_scaling)` (:45), which is False
        # for the np.float32 array built in __init__, so it always returns `num_scales`.  The same
        # predicate is kept here so both implementations see the same number of scales.
        if is_list(self.progressive_scaling):
            return int(self.num_scales - np.searchsorted(self.progressive_scaling, progress))
        return self.num_scales


class LossBase(nn.Module):
    """Holds the logs / metrics dictionaries (reference :53-81).  No parameters, no buffers."""

    def __init__(self):
        super().__init__()
        self._logs = {}
        self._metrics = {}

    @property
    def logs(self):
        return self._logs

    @property
    def metrics(self):
        return self._metrics

    def add_metric(self, key, val):
        """Store a detached copy of `val` under `key`."""
        self._metrics[key] = val.detach() if hasattr(val, 'detach') else torch.tensor(float(val))
