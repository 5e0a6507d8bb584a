"""tf_cnn_benchmarks-format reporting (header, per-interval line, final total).

The reference asks for a line every 10 steps (``--display_every=10``,
benchmark-scripts/run-tf-sing-ucx-openmpi.sh:71) and users ``tee`` the output into
``~/logs/tfmn-<N>n-<B>b-real-<fabric>-r<run>.log`` (:9). The formats below follow tf_cnn_benchmarks:

    Step    Img/sec total_loss
    10      images/sec: 2510.3 +/- 12.1 (jitter = 30.5)     7.952
    ----------------------------------------------------------------
    total images/sec: 20050.12
    ----------------------------------------------------------------
"""
from __future__ import annotations

import math
import statistics
from typing import List, Sequence

RULE = "-" * 64


def header_lines(model: str, dataset: str, mode: str, global_batch: int, per_device_batch: int,
                 num_batches: int, num_epochs: float, devices: Sequence[str], data_format: str,
                 optimizer: str, variables: str, extra=()) -> List[str]:
    out = [f"Model:       {model}",
           f"Dataset:     {dataset}",
           f"Mode:        {mode}",
           "SingleSess:  False",
           f"Batch size:  {global_batch} global",
           f"             {per_device_batch} per device",
           f"Num batches: {num_batches}",
           f"Num epochs:  {num_epochs:.2f}",
           f"Devices:     {list(devices)}",
           f"Data format: {data_format}",
           f"Optimizer:   {optimizer}",
           f"Variables:   {variables}"]
    out += list(extra)
    out.append("==========")
    return out


def table_header() -> str:
    return "Step\tImg/sec\ttotal_loss"


def jitter(step_times: Sequence[float]) -> float:
    """tf_cnn_benchmarks' jitter: 1.4826 × median absolute deviation of the per-step speeds' times."""
    if len(step_times) < 2:
        return 0.0
    med = statistics.median(step_times)
    return 1.4826 * statistics.median(abs(t - med) for t in step_times)


def step_line(step: int, batch_size: int, step_times: Sequence[float], loss: float, top1=None, top5=None) -> str:
    """``step_times`` are the seconds of every timed step so far (tf_cnn_benchmarks reports the running
    mean speed, its standard error and the jitter of the speeds)."""
    speeds = [batch_size / t for t in step_times if t > 0]
    if not speeds:
        return f"{step}\timages/sec: 0.0 +/- 0.0 (jitter = 0.0)\t{loss:.3f}"
    mean = sum(speeds) / len(speeds)
    if len(speeds) > 1:
        var = sum((s - mean) ** 2 for s in speeds) / (len(speeds) - 1)
        unc = math.sqrt(var) / math.sqrt(len(speeds))
    else:
        unc = 0.0
    line = f"{step}\timages/sec: {mean:.1f} +/- {unc:.1f} (jitter = {jitter(speeds):.1f})\t{loss:.3f}"
    if top1 is not None:
        line += f"\t{top1:.3f}\t{top5:.3f}"
    return line


def total_lines(images_per_sec: float) -> List[str]:
    return [RULE, f"total images/sec: {images_per_sec:.2f}", RULE]


def log_name(num_nodes: int, batch: int, fabric: str, run: int = 1, data: str = "synth") -> str:
    """The reference's log-file naming convention (run-tf-sing-ucx-openmpi.sh:9)."""
    return f"tfmn-{num_nodes}n-{batch}b-{data}-{fabric}-r{run}.log"
