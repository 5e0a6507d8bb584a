"""tf_cnn_benchmarks-style flag surface.

The reference launcher passes 19 flags to tf_cnn_benchmarks.py
(benchmark-scripts/run-tf-sing-ucx-openmpi.sh:62-81). Every one of them is accepted here with the
same spelling (``--flag=value``, absl-style booleans ``TRUE``/``False``). CPU-affinity knobs of the
Intel stack (``--kmp_*``, ``--num_intra_threads``, ``--num_inter_threads``, ``--mkl``) are accepted
and reported as inert on a GPU. B200-specific additions: ``--fabric``, ``--kernels``,
``--fusion_threshold``, ``--use_cuda_graph``, ``--lars``.
"""
from __future__ import annotations

import argparse
import os
from dataclasses import dataclass, field
from typing import List, Optional

REFERENCE_FLAGS = [
    "batch_size", "num_warmup_batches", "num_batches", "model", "num_intra_threads",
    "num_inter_threads", "kmp_blocktime", "kmp_affinity", "display_every", "data_format",
    "optimizer", "forward_only", "device", "mkl", "variable_update", "horovod_device",
    "local_parameter_device", "data_dir", "data_name",
]

INERT_ON_GPU = ["num_intra_threads", "num_inter_threads", "kmp_blocktime", "kmp_affinity", "mkl",
                "local_parameter_device", "horovod_device"]

VARIABLE_UPDATES = ["horovod", "independent", "replicated", "parameter_server",
                    "distributed_replicated", "distributed_all_reduce", "collective_all_reduce"]
SUPPORTED_VARIABLE_UPDATES = ["horovod", "independent"]
OPTIMIZERS = ["momentum", "sgd", "rmsprop", "adam", "lars"]
FABRICS = ["nvl", "nccl", "gloo", "sock", "ib", "auto"]
KERNELS = ["b200", "cudnn", "auto"]


def absl_bool(v) -> bool:
    if isinstance(v, bool):
        return v
    s = str(v).strip().lower()
    if s in ("1", "true", "t", "yes", "y"):
        return True
    if s in ("0", "false", "f", "no", "n"):
        return False
    raise argparse.ArgumentTypeError(f"not a boolean: {v!r}")


@dataclass
class Params:
    # --- the reference's 19
    batch_size: int = 0
    num_warmup_batches: Optional[int] = None
    num_batches: int = 100
    model: str = "resnet50"
    num_intra_threads: int = 0
    num_inter_threads: int = 0
    kmp_blocktime: int = 0
    kmp_affinity: str = ""
    display_every: int = 10
    data_format: str = "NCHW"
    optimizer: str = "momentum"
    forward_only: bool = False
    device: str = "gpu"
    mkl: bool = False
    variable_update: str = "independent"
    horovod_device: str = ""
    local_parameter_device: str = "gpu"
    data_dir: Optional[str] = None
    data_name: str = "imagenet"
    # --- other tf_cnn_benchmarks flags that matter for the workload
    num_gpus: int = 1
    momentum: float = 0.9
    weight_decay: float = 4e-5
    init_learning_rate: Optional[float] = None
    use_fp16: bool = True          # bf16 on B200 (the name is kept for CLI parity)
    train_dir: Optional[str] = None
    save_model_steps: int = 0
    save_model_secs: int = 0
    trace_file: Optional[str] = None
    tfprof_file: Optional[str] = None
    graph_file: Optional[str] = None
    print_training_accuracy: bool = False
    seed: int = 1234
    # --- B200 additions
    fabric: str = "auto"
    kernels: str = "auto"
    fusion_threshold: int = 0      # 0 → HOROVOD_FUSION_THRESHOLD env, else 16 MiB on fabric nvl / 128 MiB elsewhere
    use_cuda_graph: Optional[bool] = None      # None → on for --kernels=b200 on a GPU (the measured configuration)
    lr_schedule: str = "constant"  # constant | warmup_poly (LARS style) | piecewise (tf_cnn_benchmarks ResNet style)
    num_learning_rate_warmup_epochs: float = 0.0
    num_epochs_total: float = 90.0
    horovod_timeline: Optional[str] = None     # filled from HOROVOD_TIMELINE
    lars: bool = False
    lars_eta: float = 0.001
    timeline: Optional[str] = None
    json_summary: Optional[str] = None
    notes: List[str] = field(default_factory=list)


def build_parser() -> argparse.ArgumentParser:
    p = argparse.ArgumentParser(
        prog="cnn_benchmarks",
        description="B200-native tf_cnn_benchmarks-style CNN training benchmark", allow_abbrev=False)
    d = Params()
    p.add_argument("--batch_size", type=int, default=d.batch_size, help="per-worker batch; 0 = model default")
    p.add_argument("--num_warmup_batches", type=int, default=None)
    p.add_argument("--num_batches", type=int, default=d.num_batches)
    p.add_argument("--model", type=str, default=d.model)
    p.add_argument("--num_intra_threads", type=int, default=0)
    p.add_argument("--num_inter_threads", type=int, default=0)
    p.add_argument("--kmp_blocktime", type=int, default=0)
    p.add_argument("--kmp_affinity", type=str, default="")
    p.add_argument("--display_every", type=int, default=d.display_every)
    p.add_argument("--data_format", type=str, default=d.data_format, choices=["NCHW", "NHWC"])
    p.add_argument("--optimizer", type=str, default=d.optimizer, choices=OPTIMIZERS)
    p.add_argument("--forward_only", type=absl_bool, default=False, nargs="?", const=True)
    p.add_argument("--device", type=str, default=d.device, choices=["cpu", "gpu"])
    p.add_argument("--mkl", type=absl_bool, default=False, nargs="?", const=True)
    p.add_argument("--variable_update", type=str, default=d.variable_update, choices=VARIABLE_UPDATES)
    p.add_argument("--horovod_device", type=str, default="")
    p.add_argument("--local_parameter_device", type=str, default="gpu")
    p.add_argument("--data_dir", type=str, default=None)
    p.add_argument("--data_name", type=str, default=d.data_name)
    p.add_argument("--num_gpus", type=int, default=1)
    p.add_argument("--momentum", type=float, default=d.momentum)
    p.add_argument("--weight_decay", type=float, default=d.weight_decay)
    p.add_argument("--init_learning_rate", type=float, default=None)
    p.add_argument("--use_fp16", type=absl_bool, default=True, nargs="?", const=True)
    p.add_argument("--train_dir", type=str, default=None)
    p.add_argument("--save_model_steps", type=int, default=0)
    p.add_argument("--save_model_secs", type=int, default=0)
    p.add_argument("--trace_file", type=str, default=None)
    p.add_argument("--tfprof_file", type=str, default=None)
    p.add_argument("--graph_file", type=str, default=None)
    p.add_argument("--print_training_accuracy", type=absl_bool, default=False, nargs="?", const=True)
    p.add_argument("--seed", type=int, default=d.seed)
    p.add_argument("--fabric", type=str, default=d.fabric, choices=FABRICS)
    p.add_argument("--kernels", type=str, default=d.kernels, choices=KERNELS)
    p.add_argument("--fusion_threshold", type=int, default=0)
    p.add_argument("--use_cuda_graph", type=absl_bool, default=None, nargs="?", const=True)
    p.add_argument("--lr_schedule", type=str, default=d.lr_schedule, choices=["constant", "warmup_poly", "piecewise"])
    p.add_argument("--num_learning_rate_warmup_epochs", type=float, default=0.0)
    p.add_argument("--num_epochs_total", type=float, default=d.num_epochs_total)
    p.add_argument("--lars", type=absl_bool, default=False, nargs="?", const=True)
    p.add_argument("--lars_eta", type=float, default=d.lars_eta)
    p.add_argument("--timeline", type=str, default=None)
    p.add_argument("--json_summary", type=str, default=None)
    return p


NVL_FUSION_THRESHOLD = 16 << 20


def resolve_fusion_threshold(p: Params, fabric: Optional[str] = None) -> int:
    """--fusion_threshold, else HOROVOD_FUSION_THRESHOLD (run-tf-sing-ucx-openmpi.sh:105), else the fabric's default:
    Horovod's 128 MiB for the library fabrics (the reference's setting), 16 MiB on the NVLink fabric. There a bucket
    costs one kernel launch and no negotiation, so buckets are sized by backward time rather than by message
    overhead: ResNet-50's 102 MB of gradients become seven buckets and the first ones are reduced and applied while
    the earlier layers are still in backward; with 128 MiB the whole model is ONE bucket that can only start after
    the last gradient exists."""
    if p.fusion_threshold > 0:
        return p.fusion_threshold
    env = os.environ.get("HOROVOD_FUSION_THRESHOLD")
    if env:
        return int(env)
    return NVL_FUSION_THRESHOLD if fabric == "nvl" else 134217728


def resolve_fusion_tail(fabric: Optional[str] = None) -> int:
    """Bytes of the gradient stream's END that get a bucket of their own (B200_FUSION_TAIL; default 1 MiB on the NVLink
    fabric, 0 = Horovod's plain greedy fill elsewhere). The last bucket's collective runs after backward has ended,
    so whatever it takes is exposed: ResNet-50's last megabyte is conv0 + the first residual block."""
    env = os.environ.get("B200_FUSION_TAIL")
    if env not in (None, ""):
        return int(env)
    return (1 << 20) if fabric == "nvl" else 0


def parse(argv=None) -> Params:
    ns = build_parser().parse_args(argv)
    p = Params(**vars(ns))
    p.horovod_timeline = os.environ.get("HOROVOD_TIMELINE") or None
    validate(p)
    return p


def validate(p: Params) -> None:
    if p.variable_update not in SUPPORTED_VARIABLE_UPDATES:
        raise SystemExit(
            f"--variable_update={p.variable_update} is a parameter-server / in-graph replication mode of "
            f"tf_cnn_benchmarks that the reference never uses (it passes horovod); supported: "
            f"{SUPPORTED_VARIABLE_UPDATES}")
    if p.data_name not in ("imagenet", "cifar10"):
        raise SystemExit("--data_name must be imagenet or cifar10 (synthetic data of that shape)")
    from ..models import dataset_of, available_models
    if p.model in available_models() and dataset_of(p.model) != p.data_name:
        want = dataset_of(p.model)
        p.notes.append(f"--model={p.model} is a {want} model: --data_name={p.data_name} replaced by {want}")
        p.data_name = want
    if p.data_dir:
        p.notes.append(f"--data_dir={p.data_dir} ignored: no dataset on this box, synthetic ImageNet is used")
        p.data_dir = None
    if p.fabric == "ib":
        p.notes.append("--fabric=ib maps to nvl (NVLink 5 / NVSwitch replaces the InfiniBand path)")
        p.fabric = "nvl"
    if p.fabric == "sock":
        p.fabric = "gloo"
    if p.device == "cpu":
        if p.fabric in ("auto",):
            p.fabric = "gloo"
        p.kernels = "cudnn" if p.kernels == "auto" else p.kernels
        p.use_fp16 = False if p.device == "cpu" else p.use_fp16
    if p.device == "gpu":
        inert = [f for f in INERT_ON_GPU if getattr(p, f) not in (0, "", False, "gpu", "cpu")]
        if inert:
            p.notes.append("CPU-affinity flags accepted and ignored on GPU: " + ", ".join("--" + f for f in inert))
    if p.lars:
        p.optimizer = "lars"
    if p.num_batches <= 0:
        raise SystemExit("--num_batches must be positive")
    if p.display_every <= 0:
        p.display_every = 10
