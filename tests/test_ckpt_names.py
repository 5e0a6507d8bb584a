"""tf_cnn_benchmarks variable naming + checkpoint round trip. The reference never enables
checkpoints (no --train_dir at run-tf-sing-ucx-openmpi.sh:62-81); the product definition requires them."""
import os

import torch

from azure_hc_intel_tf_b200 import hvd
from azure_hc_intel_tf_b200.models import create_model
from azure_hc_intel_tf_b200.utils import ckpt, tfnames


def test_resnet50_names():
    m = create_model("resnet50")
    names = [n for n, _ in tfnames.named_tf_parameters(m)]
    assert names[0] == "v0/cg/conv0/conv2d/kernel"
    assert "v0/cg/conv0/batchnorm0/gamma" in names and "v0/cg/conv0/batchnorm0/beta" in names
    assert "v0/cg/resnet_v10/conv1/conv2d/kernel" in names
    assert "v0/cg/resnet_v115/conv52/conv2d/kernel" in names       # 16th block, 53rd conv
    assert "v0/cg/affine0/weights" in names and "v0/cg/affine0/biases" in names
    assert len(names) == 161 and len(set(names)) == 161
    bufs = [n for n, *_ in tfnames.named_tf_buffers(m)]
    assert "v0/cg/conv0/batchnorm0/moving_mean" in bufs and len(bufs) == 2 * 53


def test_layouts():
    m = create_model("resnet50")
    v = tfnames.variables(m, None, 7)
    assert tuple(v["v0/cg/conv0/conv2d/kernel"].shape) == (7, 7, 3, 64)        # HWIO
    assert tuple(v["v0/cg/affine0/weights"].shape) == (2048, 1001)             # in x out
    assert int(v["global_step"]) == 7
    w = torch.randn(8, 4, 3, 3)
    assert torch.equal(tfnames.from_tf_layout(tfnames.to_tf_layout(w, "HWIO"), "HWIO"), w)


def test_inception_has_no_gamma_and_vgg_has_biases():
    inc = [n for n, _ in tfnames.named_tf_parameters(create_model("inception3"))]
    assert not any(n.endswith("/gamma") for n in inc) and any(n.endswith("/beta") for n in inc)
    assert any("/incept0/" in n for n in inc)
    vgg = [n for n, _ in tfnames.named_tf_parameters(create_model("vgg16"))]
    assert "v0/cg/conv0/biases" in vgg and "v0/cg/affine2/weights" in vgg and len(vgg) == 32


def test_round_trip_with_momentum(tmp_path):
    torch.manual_seed(0)
    m = create_model("googlenet")
    named = [(n, p) for n, p in m.named_parameters() if p.requires_grad]
    opt = hvd.MomentumSGD(named, lr=0.01, momentum=0.9, weight_decay=1e-4)
    x = torch.randn(2, 3, 224, 224)
    for _ in range(2):
        opt.zero_grad()
        m(x).square().mean().backward()
        opt.step()
    path = ckpt.save(str(tmp_path), m, opt, 2)
    assert os.path.basename(path) == "model.ckpt-2.npz"
    assert 'model_checkpoint_path: "model.ckpt-2"' in open(tmp_path / "checkpoint").read()
    saved = ckpt.load(path)
    assert "v0/cg/conv0/conv2d/kernel/Momentum" in saved and saved["v0/cg/conv0/conv2d/kernel/Momentum"].abs().sum() > 0

    torch.manual_seed(1)
    m2 = create_model("googlenet")
    opt2 = hvd.MomentumSGD([(n, p) for n, p in m2.named_parameters() if p.requires_grad], lr=0.01)
    step = ckpt.restore(str(tmp_path), m2, opt2)
    assert step == 2
    for (n1, p1), (n2, p2) in zip(m.named_parameters(), m2.named_parameters()):
        assert torch.equal(p1, p2), n1
    assert torch.equal(opt.mom, opt2.mom)
    # a resumed run continues identically
    for mm, oo in ((m, opt), (m2, opt2)):
        oo.zero_grad()
        mm(x).square().mean().backward()
        oo.step()
    assert torch.allclose(m.head.weight, m2.head.weight)


def test_keep_and_latest(tmp_path):
    m = create_model("trivial")
    for s in (1, 2, 3, 4):
        ckpt.save(str(tmp_path), m, None, s, keep=2)
    assert sorted(ckpt.all_steps(str(tmp_path))) == [3, 4]
    assert ckpt.latest(str(tmp_path)).endswith("model.ckpt-4.npz")
    assert ckpt.save(str(tmp_path), m, None, 9, rank=1) is None       # only rank 0 writes
