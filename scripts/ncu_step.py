"""Two eager ResNet-50 training steps at the benchmark shape (batch 256, bf16): the ncu capture target for every kernel of
the step. usage: python scripts/ncu_step.py [model] [batch]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from azure_hc_intel_tf_b200.cli.cnn_benchmarks import BenchmarkCNN
from azure_hc_intel_tf_b200.utils import flags
model = sys.argv[1] if len(sys.argv) > 1 else "resnet50"
batch = sys.argv[2] if len(sys.argv) > 2 else "256"
p = flags.parse([f"--model={model}", f"--batch_size={batch}", "--variable_update=horovod", "--device=gpu", "--kernels=b200",
                 "--use_cuda_graph=False"])
b = BenchmarkCNN(p)
for _ in range(2):
    loss = b.step()
torch.cuda.synchronize()
print("ok", float(loss))
