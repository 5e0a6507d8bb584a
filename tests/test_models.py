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
    "resnet50_v2": 25551529, "inception4": 42649769, "mobilenet": 3506153,
    # NASNet-A: the published Keras / slim sizes are 5,289,978 (mobile) and 88,753,150 (large) trainable parameters
    # with 1000 classes; the 1001st class adds 1056 + 1 and 4032 + 1
    "nasnet": 5289978 + 1057, "nasnetlarge": 88753150 + 4033,
    # TF official-models ResNets: the published 11,689,512 / 21,797,672 / 25,557,032 (1000 classes) + the 1001st class
    "official_resnet18": 11689512 + 513, "official_resnet34": 21797672 + 513, "official_resnet50": 25557032 + 2049,
    # CIFAR-10 models (11 classes): the published sizes are 0.27 M / 0.85 M / 1.7 M (ResNet-20/56/110) and
    # 1.0 M / 7.0 M / 27.2 M (DenseNet 40-k12 / 100-k12 / 100-k24)
    "resnet20": 269787, "resnet56": 853083, "resnet110": 1728027, "resnet20_v2": 269787,
    "densenet40_k12": 1020171, "densenet100_k12": 6980811, "densenet100_k24": 27251403,
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


@pytest.mark.parametrize("name", ["resnet50", "googlenet", "lenet", "trivial", "resnet50_v2", "resnet20", "resnet32_v2",
                                  "densenet40_k12", "inception4", "mobilenet", "nasnet", "official_resnet18"])
def test_forward_backward_shapes(name):
    m = create_model(name)
    x = torch.randn(2, 3, m.image_size, m.image_size).contiguous(memory_format=torch.channels_last)
    y = m(x)
    assert y.shape == (2, m.cfg.num_classes)
    assert m.cfg.num_classes == (11 if m.image_size == 32 else 1001)
    y.sum().backward()
    assert all(p.grad is not None for p in m.parameters() if p.requires_grad)


def test_cifar_models_select_the_cifar_input_through_the_cli_flags():
    from azure_hc_intel_tf_b200.utils import flags
    p = flags.parse(["--model=resnet56", "--batch_size=8", "--variable_update=horovod", "--device=cpu"])
    assert p.data_name == "cifar10" and any("cifar10" in n for n in p.notes)
    p = flags.parse(["--model=resnet50", "--batch_size=8", "--variable_update=horovod", "--device=cpu"])
    assert p.data_name == "imagenet"


def test_tf_variable_names_of_the_preactivation_models():
    """ResNet v2 / DenseNet use standalone batch-norms; their scopes follow the builder's counters like the conv ones."""
    m = create_model("resnet20_v2")
    names = {getattr(p, "tf_name", None) for p in m.parameters()}
    assert "v0/cg/conv0/conv2d/kernel" in names
    assert any(n and n.startswith("v0/cg/resnet_v20/batchnorm") and n.endswith("/gamma") for n in names)
    assert "v0/cg/affine0/weights" in names
    assert None not in names


def test_registry_and_unknown_model():
    assert {"resnet50", "vgg16", "inception3", "resnet50_v1.5", "overfeat"} <= set(available_models())
    with pytest.raises(ValueError):
        create_model("no_such_model")


def test_bench_reference_arm_reports_unavailable():
    out = subprocess.run([sys.executable, "bench.py", "--impl", "reference"], capture_output=True, text=True, check=True)
    d = json.loads(out.stdout.strip().splitlines()[-1])
    assert d["impl"] == "reference" and "unavailable" in d


def test_mobilenet_depthwise_layers_and_checkpoint_layout(tmp_path):
    """17 depthwise 3x3 layers with ReLU6; their filters go to a checkpoint in TF's [kh, kw, C, 1] layout and back."""
    from azure_hc_intel_tf_b200.models.builder import Conv
    from azure_hc_intel_tf_b200.utils import tfnames
    m = create_model("mobilenet_v2")
    dw = [c for c in m.modules() if isinstance(c, Conv) and c.groups > 1]
    assert len(dw) == 17 and all(c.groups == c.cin == c.cout and c.k == (3, 3) and c.activation == "relu6" for c in dw)
    assert sum(1 for c in m.modules() if isinstance(c, Conv)) == 52
    w = dw[3].weight.detach()
    assert dw[3].weight.tf_layout == "HWC1" and dw[3].weight.tf_name.endswith("depthwise_conv2d/depthwise_kernel")
    t = tfnames.to_tf_layout(w, "HWC1")
    assert t.shape == (3, 3, w.shape[0], 1)
    assert torch.equal(tfnames.from_tf_layout(t, "HWC1"), w)
    # linear bottleneck: the projection has no activation and takes the block input as the fused residual
    blk = m.blocks[2]
    assert blk.use_res and blk.project.activation is None and not m.blocks[1].use_res


def test_relu6_and_grouped_conv_on_the_stock_path_match_torch():
    import torch.nn.functional as F
    from azure_hc_intel_tf_b200.models.builder import BuildConfig, ConvNetBuilder
    torch.manual_seed(0)
    cfg = BuildConfig(kernels="cudnn", use_batch_norm=True, bn_scale=True)
    b = ConvNetBuilder(cfg, 16, (9, 9))
    c = b.conv(16, 3, 2, "SAME", activation="relu6", groups=16, bias=None)
    x = torch.randn(2, 16, 9, 9) * 4
    y = c(x)
    (pt, pb), (pl, pr) = c.pad
    ref = F.conv2d(F.pad(x, (pl, pr, pt, pb)), c.weight, None, 2, 0, 1, 16)
    ref = F.relu6(F.batch_norm(ref, None, None, c.bn.gamma, c.bn.beta, True, 0.0, c.bn.eps))
    assert y.shape == (2, 16, 5, 5) and torch.allclose(y, ref, atol=1e-5)
    r = torch.randn_like(ref)
    c.activation = None
    assert torch.allclose(c(x, residual=r), F.batch_norm(F.conv2d(F.pad(x, (pl, pr, pt, pb)), c.weight, None, 2, 0, 1, 16),
                                                         None, None, c.bn.gamma, c.bn.beta, True, 0.0, c.bn.eps) + r, atol=1e-5)


def test_nasnet_cell_structure():
    """16 cells for the mobile model (2 stem reductions, 3 x 4 normal, 2 reductions), channel widths 6F / 4F, separable
    convolutions with 3x3 / 5x5 / 7x7 depthwise halves; the large model keeps p across a reduction cell."""
    from azure_hc_intel_tf_b200.models.builder import Conv
    m = create_model("nasnet")
    assert [c.reduction for c in m.cells] == [True, True] + [False] * 4 + [True] + [False] * 4 + [True] + [False] * 4
    assert [c.out[0] for c in m.cells] == [44, 88] + [264] * 4 + [352] + [528] * 4 + [704] + [1056] * 4
    assert m.cells[-1].out[1] == (7, 7) and not any(c.keep_p for c in m.cells)
    assert m.cells[0].adjust.mode == "same" and m.cells[1].adjust.mode == "reduce" and m.cells[3].adjust.mode == "project"
    ks = {c.k for c in m.modules() if isinstance(c, Conv) and c.groups > 1}
    assert ks == {(3, 3), (5, 5), (7, 7)}
    big = create_model("nasnetlarge")
    assert len(big.cells) == 2 + 3 * 6 + 2 and sum(c.keep_p for c in big.cells) == 2 and big.image_size == 331
