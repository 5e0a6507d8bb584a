"""Learning-rate schedules (the role of tf_cnn_benchmarks' ``get_learning_rate`` / ``--num_learning_rate_warmup_epochs``
behind the reference's ``--optimizer=momentum``, benchmark-scripts/run-tf-sing-ucx-openmpi.sh:73).

A schedule is a plain function ``step -> lr`` evaluated on the host once per step; ``MomentumSGD.set_lr`` writes the
value into the device scalar that the update kernels read (``lr_dev``), so the schedule keeps working when the whole
training step is replayed from a captured CUDA graph.

* ``constant``     — the initial rate throughout (what a 150-step benchmark run effectively sees);
* ``piecewise``    — linear warm-up from 0 over ``warmup_epochs``, then the ImageNet ResNet staircase: ×0.1 at epochs
  30, 60, 80 and ×0.01·0.1 at 90 (tf_cnn_benchmarks' ResNet default);
* ``warmup_poly``  — linear warm-up, then polynomial (power 2) decay to zero at ``total_epochs``: the schedule LARS
  large-batch training is normally run with (without warm-up LARS at batch 2048+ diverges in the first steps).
"""
from __future__ import annotations

from typing import Callable

IMAGENET_TRAIN = 1281167


def steps_per_epoch(global_batch: int, num_examples: int = IMAGENET_TRAIN) -> float:
    return num_examples / float(global_batch)


def make(kind: str, base_lr: float, global_batch: int, warmup_epochs: float = 0.0, total_epochs: float = 90.0,
         num_examples: int = IMAGENET_TRAIN) -> Callable[[int], float]:
    spe = steps_per_epoch(global_batch, num_examples)
    warm = warmup_epochs * spe

    def warmup(step: int):
        return base_lr * (step + 1) / warm if step < warm else None

    if kind == "constant":
        if warm <= 0:
            return lambda step: base_lr
        return lambda step: (lambda w: base_lr if w is None else w)(warmup(step))
    if kind == "piecewise":
        bounds = [(30, 1.0), (60, 0.1), (80, 0.01), (90, 0.001)]

        def piecewise(step: int) -> float:
            w = warmup(step)
            if w is not None:
                return w
            epoch = step / spe
            for upto, mult in bounds:
                if epoch < upto:
                    return base_lr * mult
            return base_lr * 1e-4
        return piecewise
    if kind == "warmup_poly":
        total = max(total_epochs * spe, warm + 1)

        def poly(step: int) -> float:
            w = warmup(step)
            if w is not None:
                return w
            frac = min(1.0, (step - warm) / (total - warm))
            return base_lr * (1.0 - frac) ** 2
        return poly
    raise ValueError(f"unknown learning-rate schedule {kind!r}")
