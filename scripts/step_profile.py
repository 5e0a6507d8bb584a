"""Kernel-time breakdown of a training step (torch.profiler / CUPTI). Usage: step_profile.py [model] [batch] [kernels]"""
import sys
import torch
sys.path.insert(0, ".")
from azure_hc_intel_tf_b200.cli.cnn_benchmarks import BenchmarkCNN
from azure_hc_intel_tf_b200.utils import flags
model = sys.argv[1] if len(sys.argv) > 1 else "resnet50"
batch = sys.argv[2] if len(sys.argv) > 2 else "256"
kern = sys.argv[3] if len(sys.argv) > 3 else "b200"
p = flags.parse([f"--model={model}", f"--batch_size={batch}", "--variable_update=horovod", "--device=gpu", f"--kernels={kern}",
                 "--fabric=" + ("nccl" if kern == "cudnn" else "auto")])
b = BenchmarkCNN(p)
for _ in range(5):
    b.step()
torch.cuda.synchronize()
import time
t0 = time.perf_counter()
for _ in range(10):
    b.step()
torch.cuda.synchronize()
print(f"step wall {(time.perf_counter() - t0) * 100:.2f} ms")
t0 = time.perf_counter()
for _ in range(10):
    b.step()
t1 = time.perf_counter()
torch.cuda.synchronize()
print(f"host-side launch time per step {(t1 - t0) * 100:.2f} ms (GPU-bound if smaller than the step)")
from torch.profiler import profile, ProfilerActivity
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for _ in range(3):
        b.step()
    torch.cuda.synchronize()
ev = [e for e in prof.key_averages() if e.device_time_total > 0 and e.device_type.name == "CUDA"]
tot = sum(e.device_time_total for e in ev) / 3
print(f"sum of kernel time per step: {tot / 1e3:.2f} ms")
for e in sorted(ev, key=lambda e: -e.device_time_total)[:28]:
    print(f"{e.device_time_total / 3 / 1e3:8.3f} ms  {e.count // 3:5d}x  {e.key[:110]}")
