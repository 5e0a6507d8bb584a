"""Every model family of the zoo takes two training steps through the sm_100a kernel path (tf_cnn_benchmarks models
selected with --model; the reference hard-codes resnet50 at run-tf-sing-ucx-openmpi.sh:34,66). Small batches: this is a
coverage check (no layer shape falls off the kernels, the loss is finite, nothing goes to the library), not a benchmark."""
import pytest
import torch

pytestmark = pytest.mark.gpu

MODELS = ["resnet50", "resnet50_v1.5", "resnet50_v2", "vgg11", "inception3", "inception4", "googlenet", "alexnet",
          "overfeat", "lenet", "trivial", "resnet20", "resnet20_v2", "densenet40_k12", "mobilenet", "nasnet", "official_resnet18", "official_resnet34"]


@pytest.mark.parametrize("name", MODELS)
def test_two_training_steps_on_the_kernel_path(name):
    from azure_hc_intel_tf_b200.cli.cnn_benchmarks import BenchmarkCNN
    from azure_hc_intel_tf_b200.ops import b200, conv_impl
    from azure_hc_intel_tf_b200.utils import flags
    p = flags.parse([f"--model={name}", "--batch_size=8", "--variable_update=horovod", "--device=gpu", "--kernels=b200",
                     "--use_cuda_graph=False"])
    b = BenchmarkCNN(p)
    conv_impl.path_counts.clear()
    b200.launch_count = 0
    for _ in range(2):
        loss = b.step()
    torch.cuda.synchronize()
    v = float(loss.detach())
    # (the 'trivial' model is one affine layer on raw pixels: its first losses are in the thousands)
    assert v == v and 0.0 < v < (1e5 if name == "trivial" else 100.0), v
    assert b200.launch_count > 0
    paths = dict(conv_impl.path_counts)
    # layers whose channel counts are multiples of 8 never leave the tensor-core kernels
    lib = {k: n for k, n in paths.items() if k.endswith("_lib") and not k.startswith("gemm")}
    if name not in ("densenet40_k12", "resnet20", "resnet20_v2", "lenet", "trivial"):      # growth 12 / 16-channel stems
        assert not lib, paths
