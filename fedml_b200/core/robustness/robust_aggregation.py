"""Reference-compatible ``RobustAggregator`` facade (core/robustness/robust_aggregation.py:41-99).

Method names and argument shapes match the reference so the ``FedAvg_robust`` runner and user
code keep working; the maths lives in :mod:`fedml_b200.core.aggregation` (CPU / oracle) and in
the sm_100a kernels of ``csrc/robust.cu`` (device-resident stacked updates).  Krum and trimmed
mean are additions (absent from the reference).
"""
from collections import OrderedDict

import torch

from .. import aggregation as A
from ..aggregation import is_weight_param, vectorize_weight  # re-export, reference names

__all__ = ["RobustAggregator", "is_weight_param", "vectorize_weight", "load_model_weight_diff"]


def load_model_weight_diff(local_state_dict, weight_diff, global_state_dict):
    """``w_t + clipped(w_local - w_t)`` scattered back over weight entries."""
    sd = local_state_dict.state_dict() if hasattr(local_state_dict, "state_dict") and \
        not isinstance(local_state_dict, dict) else local_state_dict
    out, off = OrderedDict(), 0
    for k, v in sd.items():
        if is_weight_param(k):
            out[k] = weight_diff[off:off + v.numel()].view(v.size()) + global_state_dict[k]
            off += v.numel()
        else:
            out[k] = v
    return out


class RobustAggregator(object):
    def __init__(self, args):
        self.defense_type = getattr(args, "defense_type", None)
        self.norm_bound = float(getattr(args, "norm_bound", 5.0))
        self.stddev = float(getattr(args, "stddev", 0.025))
        self.trim_ratio = float(getattr(args, "trim_ratio", 0.1))
        self.krum_f = int(getattr(args, "krum_f", 0))
        self.krum_m = int(getattr(args, "krum_m", 1))

    def norm_diff_clipping(self, local_state_dict, global_state_dict):
        return A.norm_diff_clip(local_state_dict, global_state_dict, self.norm_bound)

    def add_noise(self, local_weight, device=None, generator=None):
        if isinstance(local_weight, dict):
            return A.add_gaussian_noise(local_weight, self.stddev, generator)
        noise = torch.randn(local_weight.size(), device=local_weight.device, generator=generator)
        return local_weight + noise * self.stddev

    def coordinate_median_agg(self, model_list):
        return A.coordinate_median(model_list)

    def trimmed_mean_agg(self, model_list, trim_ratio=None):
        return A.trimmed_mean(model_list, self.trim_ratio if trim_ratio is None else trim_ratio)

    def krum_agg(self, model_list, f=None, m=None):
        return A.krum(model_list, self.krum_f if f is None else f, self.krum_m if m is None else m)

    def defend(self, model_list, global_state_dict, generator=None):
        """Apply ``defense_type`` to ``[(n_i, state_dict)]`` and return the aggregate."""
        kind = (self.defense_type or "none").lower()
        if kind in ("norm_diff_clipping", "weak_dp"):
            out = []
            for n, sd in model_list:
                sd = self.norm_diff_clipping(sd, global_state_dict)
                if kind == "weak_dp":
                    sd = A.add_gaussian_noise(sd, self.stddev, generator)
                out.append((n, sd))
            return A.fedavg_aggregate(out)
        if kind in ("median", "coordinate_median"):
            return self.coordinate_median_agg(model_list)
        if kind in ("trimmed_mean", "trimmedmean"):
            return self.trimmed_mean_agg(model_list)
        if kind in ("krum", "multi_krum", "multikrum"):
            return self.krum_agg(model_list)
        return A.fedavg_aggregate(model_list)
