"""Checkpoint save / resume with tf_cnn_benchmarks variable naming (``--train_dir``).

tf_cnn_benchmarks writes ``model.ckpt-<global_step>`` plus a ``checkpoint`` index file via
``tf.train.Saver``; under Horovod only rank 0 writes. TensorFlow is not installed here, so
the container is a self-describing ``.npz`` keyed by the TF variable names, in TF layouts (HWIO / in×out).
The index file keeps TF's text format so tools that look for ``model_checkpoint_path`` still work.
"""
from __future__ import annotations

import os
import re
import tempfile
from typing import Optional

import numpy as np
import torch

from . import tfnames

_INDEX = "checkpoint"


def _path(train_dir: str, step: int) -> str:
    return os.path.join(train_dir, f"model.ckpt-{step}.npz")


def save(train_dir: str, model, optimizer, global_step: int, rank: int = 0, keep: int = 5) -> Optional[str]:
    if rank != 0:
        return None
    os.makedirs(train_dir, exist_ok=True)
    vals = tfnames.variables(model, optimizer, global_step)
    arrays = {k: v.numpy() for k, v in vals.items()}
    path = _path(train_dir, global_step)
    fd, tmp = tempfile.mkstemp(dir=train_dir, suffix=".tmp.npz")
    os.close(fd)
    np.savez(tmp, **arrays)
    os.replace(tmp if tmp.endswith(".npz") else tmp + ".npz", path)
    existing = sorted(all_steps(train_dir))
    for old in existing[:-keep] if keep > 0 else []:
        try:
            os.remove(_path(train_dir, old))
        except OSError:
            pass
    kept = sorted(all_steps(train_dir))
    with open(os.path.join(train_dir, _INDEX), "w") as f:
        f.write(f'model_checkpoint_path: "model.ckpt-{global_step}"\n')
        for s in kept:
            f.write(f'all_model_checkpoint_paths: "model.ckpt-{s}"\n')
    return path


def all_steps(train_dir: str):
    out = []
    if not os.path.isdir(train_dir):
        return out
    for fn in os.listdir(train_dir):
        m = re.fullmatch(r"model\.ckpt-(\d+)\.npz", fn)
        if m:
            out.append(int(m.group(1)))
    return out


def latest(train_dir: str) -> Optional[str]:
    idx = os.path.join(train_dir, _INDEX)
    if os.path.exists(idx):
        with open(idx) as f:
            m = re.search(r'model_checkpoint_path: "model\.ckpt-(\d+)"', f.read())
        if m and os.path.exists(_path(train_dir, int(m.group(1)))):
            return _path(train_dir, int(m.group(1)))
    steps = all_steps(train_dir)
    return _path(train_dir, max(steps)) if steps else None


def load(path: str):
    with np.load(path) as z:
        return {k: torch.from_numpy(np.array(z[k])) for k in z.files}


def restore(train_dir: str, model, optimizer) -> int:
    """Restore the latest checkpoint in ``train_dir``; returns its global step, or 0 if there is none."""
    path = latest(train_dir)
    if path is None:
        return 0
    return tfnames.assign(model, optimizer, load(path))
