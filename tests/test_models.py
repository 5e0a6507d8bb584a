"""Model zoo shape / parameter-count checks on CPU (tf_cnn_benchmarks models selected with --model; the reference
hard-codes resnet50 at run-tf-sing-ucx-openmpi.sh:34)."""
import json
import subprocess
import sys

import pytest
import torch

from azure_hc_intel_tf_b200.models import available_models, create_model

EXPECTED_PARAMS = {          # trainable parameters with 1001 classes
    "resnet50": 25559081, "resnet101": 44551209, "resnet152": 60194857, "vgg16": 138361641, "vgg19": 143671337,
    "inception3": 23819401, "googlenet": 6999577, "alexnet": 61842345,
}


@pytest.mark.parametrize("name,count", sorted(EXPECTED_PARAMS.items()))
def test_parameter_counts(name, count):
    m = create_model(name)
    assert sum(p.numel() for p in m.parameters() if p.requires_grad) == count


def test_resnet50_has_161_trainable_tensors_and_53_convs():
    from azure_hc_intel_tf_b200.models.builder import Conv
    m = create_model("resnet50")
    assert sum(1 for p in m.parameters() if p.requires_grad) == 161
    assert sum(1 for x in m.modules() if isinstance(x, Conv)) == 53


@pytest.mark.parametrize("name", ["resnet50", "googlenet", "lenet", "trivial"])
def test_forward_backward_shapes(name):
    m = create_model(name)
    x = torch.randn(2, 3, m.image_size, m.image_size).contiguous(memory_format=torch.channels_last)
    y = m(x)
    assert y.shape == (2, 1001)
    y.sum().backward()
    assert all(p.grad is not None for p in m.parameters() if p.requires_grad)


def test_registry_and_unknown_model():
    assert {"resnet50", "vgg16", "inception3", "resnet50_v1.5", "overfeat"} <= set(available_models())
    with pytest.raises(ValueError):
        create_model("no_such_model")


def test_bench_reference_arm_reports_unavailable():
    out = subprocess.run([sys.executable, "bench.py", "--impl", "reference"], capture_output=True, text=True, check=True)
    d = json.loads(out.stdout.strip().splitlines()[-1])
    assert d["impl"] == "reference" and "unavailable" in d
