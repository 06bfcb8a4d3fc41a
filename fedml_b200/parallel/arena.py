"""Flat parameter arena <-> ``state_dict``.

The reference moves models around as ``OrderedDict[str, Tensor]`` on the CPU and aggregates key by
key in Python.  Here a model is one flat fp32 vector with a
fixed layout, so aggregation / DP / server-optimizer are single streaming kernels:

    [ trainable parameters | float buffers (BN running stats) | integer buffers stored as fp32 ]
      <---- P_train ---->   (each entry padded to a multiple of 4 floats = 16 B)

``P_train`` is exactly the region the reference calls "weight params" (everything except
``running_mean / running_var / num_batches_tracked``, robust_aggregation.py:33-38) and the region
FedOpt applies its server optimizer to (``named_parameters``, fedopt_api.py:173-186).  Int64
``num_batches_tracked`` counters are averaged as floats and cast back on export, which is what
the reference's promotion + ``load_state_dict`` does.
"""
from __future__ import annotations

from collections import OrderedDict
from dataclasses import dataclass
from typing import Dict, Iterable, List, Optional

import torch
import torch.nn as nn


@dataclass
class Segment:
    name: str
    shape: torch.Size
    dtype: torch.dtype
    offset: int        # floats from the start of the flat vector
    numel: int
    kind: str          # "param" | "buffer" | "int_buffer"


def _pad4(n: int) -> int:
    return (n + 3) // 4 * 4


class StackedViews(OrderedDict):
    arena = None
    layout = None


class ArenaLayout:
    def __init__(self, model_or_state, param_names: Optional[Iterable[str]] = None):
        if isinstance(model_or_state, nn.Module):
            state = model_or_state.state_dict()
            pnames = [n for n, _ in model_or_state.named_parameters()]
        else:
            state = model_or_state
            pnames = list(param_names) if param_names is not None else \
                [k for k in state if state[k].dtype.is_floating_point and "running_" not in k]
        pset = set(pnames)
        self.segments: List[Segment] = []
        off = 0
        for kind in ("param", "buffer", "int_buffer"):
            for name, t in state.items():
                k = "param" if name in pset else ("buffer" if t.dtype.is_floating_point else "int_buffer")
                if k != kind:
                    continue
                self.segments.append(Segment(name, t.shape, t.dtype, off, t.numel(), k))
                off += _pad4(t.numel())
            if kind == "param":
                self.P_train = off
        self.P = off
        self.order = list(state.keys())
        self.by_name: Dict[str, Segment] = {s.name: s for s in self.segments}

    # ------------------------------------------------------------------ flat <-> dict
    def flatten(self, state: Dict[str, torch.Tensor], out: Optional[torch.Tensor] = None, device=None) -> torch.Tensor:
        device = device if device is not None else (out.device if out is not None else "cpu")
        flat = torch.zeros(self.P, dtype=torch.float32, device=device) if out is None else out
        for s in self.segments:
            flat[s.offset:s.offset + s.numel].copy_(state[s.name].reshape(-1).to(torch.float32), non_blocking=True)
        return flat

    def unflatten(self, flat: torch.Tensor, clone: bool = True) -> "OrderedDict[str, torch.Tensor]":
        out = OrderedDict()
        for name in self.order:
            s = self.by_name[name]
            v = flat[s.offset:s.offset + s.numel].view(s.shape)
            if s.kind == "int_buffer":
                v = v.round().to(s.dtype)
            elif s.dtype != torch.float32:
                v = v.to(s.dtype)
            elif clone:
                v = v.clone()
            out[name] = v
        return out

    def views(self, flat: torch.Tensor) -> "OrderedDict[str, torch.Tensor]":
        """Zero-copy fp32 views (int buffers stay fp32 views)."""
        return OrderedDict((s.name, flat[s.offset:s.offset + s.numel].view(s.shape)) for s in self.segments)

    def stacked_views(self, stack: torch.Tensor) -> "OrderedDict[str, torch.Tensor]":
        """[C, P(+pad)] client stack -> {name: [C, *shape]} strided zero-copy views.  The returned dict also
        carries ``.arena`` (the stack) and ``.layout`` so kernels that walk the whole arena can find it."""
        C = stack.shape[0]
        out = StackedViews((s.name, stack[:, s.offset:s.offset + s.numel].view(C, *s.shape)) for s in self.segments)
        out.arena, out.layout = stack, self
        return out

    def param_segments(self) -> List[Segment]:
        return [s for s in self.segments if s.kind == "param"]


def bind_module_to_flat(model: nn.Module, layout: ArenaLayout, flat: torch.Tensor, grad_flat: torch.Tensor = None):
    """Re-point every parameter / float buffer of ``model`` at views of ``flat`` (and its ``.grad`` at
    ``grad_flat``), so eager training updates the arena in place — no state_dict copies per round."""
    views = layout.views(flat)
    gviews = layout.views(grad_flat) if grad_flat is not None else None
    named_params = dict(model.named_parameters())
    for name, p in named_params.items():
        v = views[name]
        v.copy_(p.detach().to(torch.float32))
        p.data = v
        if gviews is not None:
            p.grad = gviews[name]
    for name, b in model.named_buffers():
        s = layout.by_name.get(name)
        if s is None or s.kind != "buffer":
            continue
        v = views[name]
        v.copy_(b.detach().to(torch.float32))
        mod_name, _, leaf = name.rpartition(".")
        mod = model.get_submodule(mod_name) if mod_name else model
        # re-point the existing buffer at the arena through .data: the buffer keeps its OWN autograd version counter.
        # Installing the view itself would make every buffer share the arena's counter, and cuDNN BatchNorm saves its
        # running statistics for backward: the next layer's in-place running-stat update would then invalidate them
        # ("modified by an inplace operation").
        if isinstance(mod._buffers.get(leaf), torch.Tensor):
            mod._buffers[leaf].data = v
        else:
            mod._buffers[leaf] = v
    return model
