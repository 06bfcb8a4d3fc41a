"""TurboAggregate-style secure FedAvg in the single-process simulator.

The reference's version is plain FedAvg plus a finite-field toolbox whose protocol body is ``pass``
(single_process/turboaggregate/TA_trainer.py:99-116): note that it divides by the *global* train count.
Here each round's updates really go through the pairwise-mask secure aggregation of
``core/security/secagg.py`` (quantise -> mask -> modular sum -> unmask -> de-quantise), so the server only ever
sees the weighted sum; ``args.secagg_dropout`` simulates clients that vanish after masking (mask recovery
path).  The result equals FedAvg up to the 1/scale quantisation step."""
import torch

from ...core import aggregation as A
from ...core.security.secagg import secure_weighted_mean
from .fedavg_api import FedAvgAPI


class TurboAggregateAPI(FedAvgAPI):
    def _aggregate(self, w_locals, w_global=None):
        scale = float(getattr(self.args, "secagg_scale", 2 ** 16))
        n_drop = int(getattr(self.args, "secagg_dropout", 0) or 0)
        dropped = list(range(len(w_locals) - n_drop, len(w_locals))) if n_drop else []
        vecs = [A.vectorize_weight(sd, weights_only=False) for _, sd in w_locals]
        weights = [float(n) for n, _ in w_locals]
        mean = secure_weighted_mean(vecs, weights, scale=scale, dropped=dropped)
        return A.unvectorize_like(mean.to(vecs[0].device), w_locals[0][1], weights_only=False)
