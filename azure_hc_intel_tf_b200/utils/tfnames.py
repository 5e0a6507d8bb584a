"""Mapping between PyTorch parameters and tf_cnn_benchmarks variable names / layouts.

tf_cnn_benchmarks checkpoints (``--train_dir``; not enabled by the reference — no such flag at
benchmark-scripts/run-tf-sing-ucx-openmpi.sh:62-81 — but required by the product definition) name
variables by graph scope: ``v0/cg/conv0/conv2d/kernel``, ``v0/cg/conv0/batchnorm0/{gamma,beta,
moving_mean,moving_variance}``, ``v0/cg/resnet_v1<i>/conv<j>/...``, ``v0/cg/affine0/{weights,biases}``,
momentum slots ``<var>/Momentum`` and ``global_step``. The names themselves are attached
to each parameter by models/builder.py when the layer is created; this module is the single place that
knows the outer prefix and the layout conversions, so either can be corrected in one spot.

Layouts: TF conv kernels are HWIO (R,S,C,K), PyTorch's logical layout is (K,C,R,S); TF dense weights are
(in,out), PyTorch's are (out,in).
"""
from __future__ import annotations

from typing import Dict, Iterator, Tuple

import torch

TOWER_PREFIX = "v0"          # per-tower variable managers (independent / replicated / horovod)
MOMENTUM_SLOT = "Momentum"
GLOBAL_STEP = "global_step"


def to_tf_layout(t: torch.Tensor, layout: str) -> torch.Tensor:
    if layout == "HWIO":
        return t.permute(2, 3, 1, 0).contiguous()
    if layout == "HWC1":                       # depthwise [C,1,R,S] -> [R,S,C,1]
        return t.permute(2, 3, 0, 1).contiguous()
    if layout == "IO":
        return t.t().contiguous()
    return t.contiguous()


def from_tf_layout(t: torch.Tensor, layout: str) -> torch.Tensor:
    if layout == "HWIO":
        return t.permute(3, 2, 0, 1)
    if layout == "HWC1":
        return t.permute(2, 3, 0, 1)
    if layout == "IO":
        return t.t()
    return t


def retarget(name: str, prefix: str = TOWER_PREFIX) -> str:
    """Swap the outer scope (builder emits ``v0/...``)."""
    head, _, rest = name.partition("/")
    return f"{prefix}/{rest}" if head in ("v0", "v") else name


def named_tf_parameters(model) -> Iterator[Tuple[str, torch.nn.Parameter]]:
    seen = set()
    for _, p in model.named_parameters():
        if id(p) in seen or getattr(p, "tf_skip", False):
            continue
        seen.add(id(p))
        yield retarget(p.tf_name), p


def named_tf_buffers(model) -> Iterator[Tuple[str, torch.Tensor, object, str]]:
    for m in model.modules():
        tfb = getattr(m, "tf_buffers", None)
        if tfb:
            for attr, name in tfb.items():
                yield retarget(name), getattr(m, attr), m, attr


def variables(model, optimizer=None, global_step: int = 0) -> Dict[str, torch.Tensor]:
    """All checkpoint variables keyed by tf name, in TF layouts, on CPU."""
    out: Dict[str, torch.Tensor] = {}
    for name, p in named_tf_parameters(model):
        out[name] = to_tf_layout(p.detach().float().cpu(), p.tf_layout)
        for slot, t in _slots_of(optimizer, p).items():
            out[f"{name}/{slot}"] = to_tf_layout(t.detach().float().cpu(), p.tf_layout)
    for name, buf, _m, _a in named_tf_buffers(model):
        out[name] = buf.detach().float().cpu().clone()
    out[GLOBAL_STEP] = torch.tensor(int(global_step), dtype=torch.int64)
    return out


def _slots_of(optimizer, p) -> Dict[str, torch.Tensor]:
    """Optimizer slot variables of ``p`` keyed by TF slot name: ``Momentum`` (MomentumOptimizer), ``RMSProp`` /
    ``RMSProp_1`` (RMSPropOptimizer's rms and momentum), ``Adam`` / ``Adam_1`` (m and v)."""
    if optimizer is None:
        return {}
    opt = getattr(optimizer, "_opt", optimizer)
    if hasattr(opt, "slots_of"):
        return opt.slots_of(p) if hasattr(p, "flat_index") else {}
    st = getattr(opt, "state", {}).get(p)
    if st and st.get("momentum_buffer") is not None:
        return {MOMENTUM_SLOT: st["momentum_buffer"]}
    return {}


def assign(model, optimizer, values: Dict[str, torch.Tensor]) -> int:
    """Restore variables produced by :func:`variables`. Returns the global step."""
    missing = []
    with torch.no_grad():
        for name, p in named_tf_parameters(model):
            if name not in values:
                missing.append(name)
                continue
            p.copy_(from_tf_layout(torch.as_tensor(values[name]), p.tf_layout).to(p.dtype))
            for slot, t in _slots_of(optimizer, p).items():
                if f"{name}/{slot}" in values:
                    t.copy_(from_tf_layout(torch.as_tensor(values[f"{name}/{slot}"]), p.tf_layout).to(t.dtype))
        for name, buf, _m, _a in named_tf_buffers(model):
            if name in values:
                buf.copy_(torch.as_tensor(values[name]).to(buf.dtype))
            else:
                missing.append(name)
    if missing:
        raise KeyError(f"checkpoint lacks {len(missing)} variables, e.g. {missing[:3]}")
    opt = getattr(optimizer, "_opt", optimizer)
    if hasattr(opt, "sync_shadow"):
        opt.sync_shadow()
    return int(values.get(GLOBAL_STEP, 0))
