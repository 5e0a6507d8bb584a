"""``python -m azure_hc_intel_tf_b200.cli.cnn_benchmarks`` — the tf_cnn_benchmarks.py role.

The reference runs ``python tf_cnn_benchmarks.py $TF_ARGS`` under mpirun
(benchmark-scripts/run-tf-sing-ucx-openmpi.sh:99-109) with the 19 flags at :62-81: 50 warm-up batches,
100 timed batches, a line every 10 steps, momentum SGD, Horovod variable update. This module accepts the
same flags, prints the same header / step lines / ``total images/sec`` summary, and drives the B200 stack.

Timing is on the device: one CUDA event per step boundary on the compute stream; the per-rank time of the
timed region is reduced with MAX over ranks before images/sec is computed.
"""
from __future__ import annotations

import json
import sys
import time
from typing import List, Optional

import torch

from .. import hvd, models, ops
from ..ops import data as data_mod
from ..utils import ckpt, flags as flags_mod, log, lr_schedule
from ..utils.timeline import Timeline


class BenchmarkCNN:
    def __init__(self, params: flags_mod.Params, data_source: str = "device"):
        self.p = params
        p = params
        self.use_hvd = p.variable_update == "horovod"
        want_cuda = p.device == "gpu" and torch.cuda.is_available()
        if p.device == "gpu" and not torch.cuda.is_available():
            raise SystemExit("--device=gpu but no CUDA device is visible (use --device=cpu for the plumbing path)")
        fabric = p.fabric
        if fabric == "auto":
            fabric = "nvl" if want_cuda else "gloo"
        if self.use_hvd:
            hvd.init(fabric=fabric, device="cuda" if want_cuda else "cpu")
            self.world, self.rank = hvd.size(), hvd.rank()
            self.device = hvd.device()
            self.fabric = hvd.fabric()
        else:
            self.world, self.rank = 1, 0
            self.device = torch.device("cuda", torch.cuda.current_device()) if want_cuda else torch.device("cpu")
            self.fabric = "none"
        self.kernels = p.kernels
        if self.kernels == "auto":
            self.kernels = "b200" if want_cuda else "cudnn"
        if self.kernels == "b200" and not want_cuda:
            raise SystemExit("--kernels=b200 needs a B200; use --kernels=cudnn on CPU")
        self.dtype = torch.bfloat16 if (p.use_fp16 and want_cuda) else torch.float32
        if self.kernels == "cudnn" and want_cuda:
            # the stock-library path gets the library's own tuning: per-shape algorithm search instead of heuristics
            torch.backends.cudnn.benchmark = True
        torch.manual_seed(p.seed)
        self.model = models.create_model(p.model, kernels=self.kernels)
        self.batch = p.batch_size or self.model.default_batch_size
        self.global_batch = self.batch * self.world
        self.model.to(self.device)
        self.model.train(not p.forward_only)
        self.warmup = p.num_warmup_batches if p.num_warmup_batches is not None else max(10, min(50, p.num_batches))
        self.lr = p.init_learning_rate if p.init_learning_rate is not None else self.model.get_learning_rate(self.global_batch)
        # --timeline=<file>, or Horovod's own switch HOROVOD_TIMELINE=<file> that the reference's launcher would
        # forward to every rank with -x (run-tf-sing-ucx-openmpi.sh:104-106)
        tl_path = p.timeline or p.horovod_timeline
        self.timeline = Timeline(tl_path, self.rank) if tl_path else None
        # whole-step CUDA graph: on by default for the sm_100a kernel path (the configuration bench.py measures); a
        # timeline needs per-bucket events, which a replayed graph cannot give, so it switches the graph off
        self.use_graph = (p.use_cuda_graph if p.use_cuda_graph is not None else
                          (self.kernels == "b200" and want_cuda and self.timeline is None and not p.forward_only))
        self.opt = None if p.forward_only else self._make_optimizer()
        self.global_step = 0
        if p.train_dir:
            self.global_step = ckpt.restore(p.train_dir, self.model, self.opt)
            if self.global_step and self.rank == 0:
                print(f"Restored checkpoint at global_step {self.global_step} from {p.train_dir}")
        if self.use_hvd and self.world > 1:
            # Horovod's broadcast_global_variables(0)
            hvd.broadcast_parameters(self.model.state_dict(), root_rank=0)
            if self.opt is not None:
                hvd.broadcast_optimizer_state(self.opt, root_rank=0)
        src = data_mod.HostSynthetic if data_source == "host" else data_mod.DeviceSynthetic
        self.data = src(self.batch, self.model.image_size, self.model.cfg.num_classes, self.device, self.dtype,
                        seed=p.seed + self.rank, own_kernels=self.kernels == "b200")
        self._graph = None
        self._last_save = time.time()
        import os as _os
        self._nvtx = self.device.type == "cuda" and _os.environ.get("B200_NVTX", "") not in ("", "0")

    # ------------------------------------------------------------------ set-up
    def _make_optimizer(self):
        p = self.p
        named = [(n, q) for n, q in self.model.named_parameters() if q.requires_grad]
        thr = flags_mod.resolve_fusion_threshold(p, self.fabric if self.world > 1 else None)
        native = not (self.kernels == "cudnn" and self.fabric == "nccl")
        if native:
            # every --optimizer runs on the plan-ordered flat buffers with its own update kernel (csrc/kernels/optim.cu)
            mu = 0.0 if p.optimizer in ("sgd", "adam") else p.momentum
            rule = p.optimizer if p.optimizer in ("rmsprop", "adam") else "momentum"
            opt = hvd.MomentumSGD(named, lr=self.lr, momentum=mu, weight_decay=p.weight_decay,
                                  lars=p.optimizer == "lars", lars_eta=p.lars_eta, rule=rule,
                                  lowp=self.kernels == "b200", fusion_threshold=thr,
                                  fusion_tail=flags_mod.resolve_fusion_tail(self.fabric if self.world > 1 else None))
            if p.lr_schedule != "constant" or p.num_learning_rate_warmup_epochs > 0:
                n_img = 50000 if p.data_name == "cifar10" else lr_schedule.IMAGENET_TRAIN
                opt.lr_schedule = lr_schedule.make(p.lr_schedule, self.lr, self.global_batch,
                                                   p.num_learning_rate_warmup_epochs, p.num_epochs_total, n_img)
        else:
            # stock-library baseline arm: torch.optim on the fp32 parameters
            ps = [q for _, q in named]
            if p.optimizer in ("momentum", "sgd", "lars"):
                opt = torch.optim.SGD(ps, lr=self.lr, momentum=0.0 if p.optimizer == "sgd" else p.momentum,
                                      weight_decay=p.weight_decay, **({"fused": True} if self.device.type == "cuda" else {}))
            elif p.optimizer == "rmsprop":
                opt = torch.optim.RMSprop(ps, lr=self.lr, alpha=0.9, momentum=p.momentum, eps=1.0,
                                          weight_decay=p.weight_decay)
            else:
                opt = torch.optim.Adam(ps, lr=self.lr, weight_decay=p.weight_decay)
        if self.use_hvd:
            opt = hvd.DistributedOptimizer(opt, named_parameters=named, fusion_threshold=thr, timeline=self.timeline)
        return opt

    # ------------------------------------------------------------------ one step
    def forward_loss(self, images, labels):
        if self.kernels == "cudnn" and self.dtype == torch.bfloat16:
            with torch.autocast("cuda", dtype=torch.bfloat16):
                logits = self.model(images)
        else:
            logits = self.model(images)
        return ops.softmax_cross_entropy(logits, labels, self.model.cfg), logits

    def train_step(self, images, labels):
        if self.p.forward_only:
            with torch.no_grad():
                loss, _ = self.forward_loss(images, labels)
            return loss
        nvtx = self._nvtx
        self.opt.zero_grad(set_to_none=not hasattr(self.opt, "plan"))
        if nvtx:
            torch.cuda.nvtx.range_push("forward")
        loss, _ = self.forward_loss(images, labels)
        if nvtx:
            torch.cuda.nvtx.range_pop()
            torch.cuda.nvtx.range_push("backward+allreduce")
        loss.backward()
        if nvtx:
            torch.cuda.nvtx.range_pop()
            torch.cuda.nvtx.range_push("optimizer")
        self.opt.step()
        if nvtx:
            torch.cuda.nvtx.range_pop()
        self.global_step += 1
        return loss

    def step(self):
        images, labels = self.data.next()
        if self.use_graph and self.device.type == "cuda":
            return self._graph_step(images, labels)
        return self.train_step(images, labels)

    # ------------------------------------------------------------------ whole-step CUDA graph
    def _graph_step(self, images, labels):
        """Replay the captured step (forward, backward, fused gradient exchange and update: every launch of the
        step, including the side-stream collectives) with this step's batch. The first call runs three eager steps
        (lazy initialisation, plan handshake, allocator warm-up) and then captures. The learning rate is a device
        scalar the update kernels read at run time, so a schedule keeps working across replays; momentum and weight
        decay are launch scalars baked into the graph (changing them needs a re-capture: ``self._graph = None``)."""
        opt = getattr(self.opt, "_opt", self.opt)
        if self._graph is not None and hasattr(opt, "apply_schedule"):
            opt.apply_schedule()          # host → device scalar, ordered before the replay on this stream
        if self._graph is None:
            self._g_images = images.clone(memory_format=torch.preserve_format)
            self._g_labels = labels.clone()
            side = torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                for _ in range(3):
                    self.train_step(self._g_images, self._g_labels)
            torch.cuda.current_stream().wait_stream(side)
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            from ..ops import b200 as _b200ops
            before = _b200ops.launch_count
            with torch.cuda.graph(g):
                self._g_loss = self.train_step(self._g_images, self._g_labels)
            # every launcher of ours ran exactly once during capture: that is the number of our kernels per replay
            self.graph_launches = _b200ops.launch_count - before
            self._graph = g
        if images.data_ptr() != self._g_images.data_ptr():
            self._g_images.copy_(images, non_blocking=True)
            self._g_labels.copy_(labels, non_blocking=True)
        self._graph.replay()
        self.global_step += 1
        if hasattr(opt, "lr_dev"):
            opt.steps += 1                # the captured step() ran on the device only; keep the schedule's clock going
        return self._g_loss

    # ------------------------------------------------------------------ reporting
    def print_info(self):
        if self.rank != 0:
            return
        p = self.p
        dev = "gpu" if self.device.type == "cuda" else "cpu"
        devices = [f"horovod/{dev}:{i}" for i in range(self.world)] if self.use_hvd else [f"/{dev}:0"]
        n_img = 50000 if p.data_name == "cifar10" else 1281167       # training-set sizes tf_cnn_benchmarks reports epochs against
        extra = []
        if self.use_hvd:
            extra.append(f"Horovod on:  {dev}")
            extra.append(f"Fabric:      {self.fabric}")
        extra.append(f"Kernels:     {self.kernels}")
        extra.append(f"Precision:   {'bf16' if self.dtype == torch.bfloat16 else 'fp32'}")
        lines = log.header_lines(self.model.name, f"{p.data_name} (synthetic)",
                                 "forward-only" if p.forward_only else "training", self.global_batch, self.batch,
                                 p.num_batches, p.num_batches * self.global_batch / n_img, devices,
                                 "NHWC" if self.device.type == "cuda" else p.data_format, p.optimizer,
                                 p.variable_update, extra)
        for note in p.notes:
            print("Note:", note)
        print("\n".join(lines))

    def total_loss(self, xent: torch.Tensor) -> float:
        v = float(xent.detach().float().item())
        if not self.p.forward_only and self.p.weight_decay > 0:
            v += float(ops.l2_loss([q for q in self.model.parameters() if q.requires_grad and getattr(q, "decay", True)],
                                   self.p.weight_decay).item())
        return v

    def _sync(self):
        if self.device.type == "cuda":
            torch.cuda.synchronize()

    def _barrier(self):
        if self.use_hvd and self.world > 1:
            hvd.barrier()

    def run(self) -> dict:
        p = self.p
        self.print_info()
        cuda = self.device.type == "cuda"
        if self.rank == 0:
            print("Running warm up")
        for _ in range(self.warmup):
            loss = self.step()
        self._sync()
        self._barrier()
        self._sync()
        # the profiled step is a real training step with its collectives, so EVERY rank runs it; rank 0 writes
        if p.graph_file and self.rank == 0:
            self._write_graph(p.graph_file)
        if p.trace_file or p.tfprof_file:
            self._write_trace(p.trace_file, p.tfprof_file)
        if self.rank == 0:
            print("Done warm up")
            print(log.table_header())
        events: List = []
        wall: List[float] = []
        step_times: List[float] = []
        losses = []

        def mark():
            if cuda:
                e = torch.cuda.Event(enable_timing=True)
                e.record()
                events.append(e)
            else:
                wall.append(time.perf_counter())

        mark()
        last_reported = 0
        for i in range(1, p.num_batches + 1):
            loss = self.step()
            mark()
            if i % p.display_every == 0 or i == p.num_batches:
                self._sync()
                if cuda:
                    step_times += [events[j].elapsed_time(events[j + 1]) / 1e3 for j in range(last_reported, i)]
                else:
                    step_times += [wall[j + 1] - wall[j] for j in range(last_reported, i)]
                last_reported = i
                tl = self.total_loss(loss)
                losses.append(tl)
                if self.rank == 0:
                    print(log.step_line(i, self.global_batch, step_times, tl), flush=True)
            self._maybe_save()
        self._sync()
        total_s = sum(step_times)
        if self.use_hvd and self.world > 1:
            t = torch.tensor([total_s], dtype=torch.float64, device=self.device if self.fabric != "gloo" else "cpu")
            torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
            total_s = float(t.item())
        self._barrier()
        ips = p.num_batches * self.global_batch / total_s if total_s > 0 else 0.0
        if self.rank == 0:
            print("\n".join(log.total_lines(ips)), flush=True)
        if p.train_dir:
            ckpt.save(p.train_dir, self.model, self.opt, self.global_step, self.rank)
        summary = {"model": self.model.name, "images_per_sec": ips, "ms_per_step": 1e3 * total_s / p.num_batches,
                   "global_batch": self.global_batch, "world": self.world, "fabric": self.fabric,
                   "kernels": self.kernels, "final_loss": losses[-1] if losses else None}
        if p.json_summary and self.rank == 0:
            with open(p.json_summary, "w") as f:
                json.dump(summary, f)
        if self.timeline is not None:
            self.timeline.close()
        return summary

    def _write_trace(self, trace_path: Optional[str], tfprof_path: Optional[str]):
        """--trace_file: Chrome trace of one training step; --tfprof_file: per-op time table of the same step
        (tf_cnn_benchmarks' flags of the same names; the reference sets neither — SURVEY.md §5.1). CPU ops, CUDA
        kernels and NVTX ranges of exactly one step. Called on every rank (the step contains collectives); only
        rank 0 profiles and writes."""
        if self.rank != 0:
            self.step()
            self._sync()
            return
        from torch.profiler import ProfilerActivity, profile
        acts = [ProfilerActivity.CPU] + ([ProfilerActivity.CUDA] if self.device.type == "cuda" else [])
        was = self._nvtx
        self._nvtx = self.device.type == "cuda"
        with profile(activities=acts) as prof:
            self.step()
            self._sync()
        self._nvtx = was
        if trace_path:
            prof.export_chrome_trace(trace_path)
            print(f"Wrote trace of one step to {trace_path}")
        if tfprof_path:
            key = "device_time_total" if self.device.type == "cuda" else "self_cpu_time_total"
            evs = sorted(prof.key_averages(), key=lambda e: -getattr(e, key, 0.0))
            with open(tfprof_path, "w") as f:
                f.write(f"# per-op profile of one training step: model={self.model.name} batch={self.batch} "
                        f"device={self.device.type}; times in microseconds, sorted by {key}\n")
                f.write(f"{'calls':>7s} {'device_us':>12s} {'self_cpu_us':>12s}  op\n")
                for e in evs:
                    dev_us = float(getattr(e, "device_time_total", 0.0) or 0.0)
                    cpu_us = float(getattr(e, "self_cpu_time_total", 0.0) or 0.0)
                    if dev_us <= 0.0 and cpu_us <= 0.0:
                        continue
                    f.write(f"{e.count:7d} {dev_us:12.1f} {cpu_us:12.1f}  {e.key}\n")
            print(f"Wrote per-op profile of one step to {tfprof_path}")

    def _write_graph(self, path: str):
        """--graph_file: a text description of the model graph (tf_cnn_benchmarks writes the GraphDef; here: one line
        per layer in module order with its geometry, then every variable with its tf_cnn_benchmarks name, checkpoint
        layout and shape)."""
        with open(path, "w") as f:
            f.write(f"# model {self.model.name}: input {self.model.image_size}x{self.model.image_size}x3, "
                    f"{self.model.cfg.num_classes} classes, kernels={self.kernels}\n# layers\n")
            for name, mod in self.model.named_modules():
                if len(list(mod.children())) == 0:
                    f.write(f"{name or '.'}: {mod.__class__.__name__}({mod.extra_repr()})\n")
            f.write("# variables (tf name, layout, shape, trainable)\n")
            total = 0
            for pname, prm in self.model.named_parameters():
                if getattr(prm, "tf_skip", False):
                    continue
                total += prm.numel() if prm.requires_grad else 0
                f.write(f"{getattr(prm, 'tf_name', pname)} {getattr(prm, 'tf_layout', '-')} {tuple(prm.shape)} "
                        f"{'trainable' if prm.requires_grad else 'frozen'}\n")
            f.write(f"# trainable parameters: {total}\n")
        print(f"Wrote model graph description to {path}")

    def _maybe_save(self):
        p = self.p
        if not p.train_dir or self.p.forward_only:
            return
        due = (p.save_model_steps and self.global_step % p.save_model_steps == 0) or \
              (p.save_model_secs and time.time() - self._last_save >= p.save_model_secs)
        if due:
            ckpt.save(p.train_dir, self.model, self.opt, self.global_step, self.rank)
            self._last_save = time.time()


def main(argv: Optional[List[str]] = None) -> int:
    params = flags_mod.parse(argv)
    bench = BenchmarkCNN(params)
    try:
        bench.run()
    finally:
        if bench.use_hvd:
            hvd.shutdown()
    return 0


if __name__ == "__main__":
    sys.exit(main())
