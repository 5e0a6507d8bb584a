"""Depthwise convolution kernels (csrc/kernels/depthwise.cu) and the ReLU6 mode of the BN kernels vs plain PyTorch fp32
from the same bf16 inputs; MobileNet-v2 (tf_cnn_benchmarks --model=mobilenet, run-tf-sing-ucx-openmpi.sh:66) end to end."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu
DEV = "cuda"


class M:
    def __init__(self, c, k, stride, pad, in_hw):
        self.cin = self.cout = self.groups = c
        self.k, self.stride, self.pad = (k, k), (stride, stride), pad
        h, w = in_hw
        self.out_hw = ((h + sum(pad[0]) - k) // stride + 1, (w + sum(pad[1]) - k) // stride + 1)

    def extra_repr(self):
        return f"depthwise {self.cin} k={self.k} s={self.stride} pad={self.pad}"


def rnd(n, h, w, c, seed):
    g = torch.Generator().manual_seed(seed)
    return torch.randn(n, h, w, c, generator=g).to(DEV).to(torch.bfloat16).permute(0, 3, 1, 2)


def rel_err(a, b):
    return ((a.float() - b.float()).norm() / (b.float().norm() + 1e-12)).item()


CASES = [
    # (n, h, w, c, k, stride, pad)      the MobileNet-v2 shapes (SAME padding), plus odd sizes and a 2x2 filter
    (2, 112, 112, 32, 3, 1, ((1, 1), (1, 1))),
    (2, 112, 112, 96, 3, 2, ((0, 1), (0, 1))),
    (2, 56, 56, 144, 3, 1, ((1, 1), (1, 1))),
    (2, 56, 56, 144, 3, 2, ((0, 1), (0, 1))),
    (3, 28, 28, 192, 3, 2, ((0, 1), (0, 1))),
    (3, 14, 14, 384, 3, 1, ((1, 1), (1, 1))),
    (2, 14, 14, 576, 3, 2, ((0, 1), (0, 1))),
    (5, 7, 7, 960, 3, 1, ((1, 1), (1, 1))),
    (2, 9, 13, 1040, 3, 1, ((1, 1), (1, 1))),       # more than 128 channel vectors: the per-thread channel loop
    (2, 15, 11, 40, 3, 2, ((1, 1), (1, 1))),
    (2, 10, 10, 64, 2, 2, ((0, 0), (0, 0))),
    # generic kernels (taps staged in shared memory): NASNet's 5x5 / 7x7 separable convolutions, a 3x3 at stride 3
    (2, 28, 28, 88, 5, 1, ((2, 2), (2, 2))),
    (2, 28, 28, 88, 7, 2, ((2, 3), (2, 3))),
    (3, 14, 14, 176, 7, 1, ((3, 3), (3, 3))),
    (2, 9, 9, 352, 5, 2, ((2, 2), (2, 2))),
    (2, 11, 7, 1040, 5, 1, ((2, 2), (2, 2))),       # five channel chunks of 32 vectors
    (2, 13, 13, 48, 3, 3, ((1, 1), (1, 1))),
]


@pytest.mark.parametrize("n,h,w,c,k,stride,pad", CASES)
def test_depthwise_fprop_dgrad_wgrad(n, h, w, c, k, stride, pad):
    from azure_hc_intel_tf_b200.ops import conv_umma
    m = M(c, k, stride, pad, (h, w))
    x = rnd(n, h, w, c, 1)
    g = torch.Generator().manual_seed(2)
    wt = (torch.randn(c, 1, k, k, generator=g) / k).to(DEV).to(torch.bfloat16)
    assert conv_umma.supports_depthwise(x.shape, m)
    (pt, pb), (pl, pr) = pad
    xr = F.pad(x.float(), (pl, pr, pt, pb)).requires_grad_(True)
    wr = wt.float().requires_grad_(True)
    yr = F.conv2d(xr, wr, None, stride, 0, 1, c)
    y = conv_umma.dw_fprop(x, wt, m)
    assert y.shape == yr.shape and y.is_contiguous(memory_format=torch.channels_last)
    assert rel_err(y, yr) < 5e-3
    dy = rnd(n, m.out_hw[0], m.out_hw[1], c, 3)
    yr.backward(dy.float())
    dxr = xr.grad[:, :, pt:pt + h, pl:pl + w]
    dx = conv_umma.dw_dgrad(dy, wt, m, x.shape)
    assert rel_err(dx, dxr) < 5e-3
    dw = conv_umma.dw_wgrad(x, dy, m)
    assert dw.shape == (c, 1, k, k) and dw.dtype == torch.float32
    assert rel_err(dw, wr.grad) < 2e-3
    # fixed-order fold: the filter gradient is bitwise reproducible
    assert torch.equal(dw, conv_umma.dw_wgrad(x, dy, m))


@pytest.mark.parametrize("residual", [False, True])
def test_bn_relu6_forward_backward(residual):
    """relu=2 of bn_apply clamps to [0, 6]; the backward passes the gradient where 0 < f < 6 (one mask bit per element)."""
    from azure_hc_intel_tf_b200 import ops
    from azure_hc_intel_tf_b200.models.builder import BatchNorm, BuildConfig
    cfg = BuildConfig(kernels="b200", bn_scale=True)
    bn = BatchNorm(cfg, 64, "bn").to(DEV)
    with torch.no_grad():
        bn.gamma.uniform_(2.0, 4.0)          # spread so that both clamps are hit
        bn.beta.uniform_(1.0, 3.0)
    y = rnd(4, 14, 14, 64, 5).requires_grad_(True)
    res = rnd(4, 14, 14, 64, 6).requires_grad_(True) if residual else None
    z = ops.batch_norm(y, bn, relu=2, residual=res)
    dz = rnd(4, 14, 14, 64, 7)
    z.backward(dz)
    yf = y.detach().float().requires_grad_(True)
    rf = res.detach().float().requires_grad_(True) if residual else None
    gam, bet = bn.gamma.detach().clone().requires_grad_(True), bn.beta.detach().clone().requires_grad_(True)
    pre = F.batch_norm(yf, None, None, gam, bet, True, 0.0, bn.eps)
    if residual:
        pre = pre + rf
    zr = F.relu6(pre)
    frac_hi = (zr >= 6.0).float().mean().item()
    frac_lo = (zr <= 0.0).float().mean().item()
    assert frac_hi > 0.02 and frac_lo > 0.02, (frac_lo, frac_hi)
    zr.backward(dz.float())
    assert float(z.max()) <= 6.0 and float(z.min()) >= 0.0
    assert rel_err(z, zr) < 1e-2
    assert rel_err(y.grad, yf.grad) < 3e-2
    assert rel_err(bn.gamma.grad, gam.grad) < 2e-2 and rel_err(bn.beta.grad, bet.grad) < 2e-2
    if residual:
        assert rel_err(res.grad, rf.grad) < 2e-2


def test_mobilenet_trains_on_own_kernels_and_loss_falls():
    from azure_hc_intel_tf_b200.cli.cnn_benchmarks import BenchmarkCNN
    from azure_hc_intel_tf_b200.ops import b200, conv_impl
    from azure_hc_intel_tf_b200.utils import flags
    p = flags.parse(["--model=mobilenet", "--batch_size=16", "--variable_update=horovod", "--device=gpu", "--kernels=b200",
                     "--use_cuda_graph=False", "--init_learning_rate=0.02"])
    b = BenchmarkCNN(p)
    conv_impl.path_counts.clear()
    b200.launch_count = 0
    losses = [float(b.step().detach()) for _ in range(12)]      # synthetic data: one batch, so the loss must fall
    torch.cuda.synchronize()
    assert all(v == v and v < 100.0 for v in losses), losses
    assert min(losses[-3:]) < losses[0], losses
    paths = dict(conv_impl.path_counts)
    assert paths.get("fprop_depthwise") == 17 * 12 and paths.get("dgrad_depthwise") == 17 * 12 and paths.get("wgrad_depthwise") == 17 * 12, paths
    assert not {k: n for k, n in paths.items() if k.endswith("_lib") and not k.startswith("gemm")}, paths


def test_standalone_relu_node():
    from azure_hc_intel_tf_b200.ops import b200
    x = rnd(3, 9, 11, 40, 11).requires_grad_(True)
    z = b200.relu(x)
    dz = rnd(3, 9, 11, 40, 12)
    z.backward(dz)
    assert torch.equal(z, torch.relu(x.detach())) and z.data_ptr() != x.data_ptr()
    assert torch.equal(x.grad, torch.where(x.detach() > 0, dz, torch.zeros_like(dz)))


@pytest.mark.parametrize("c,k,stride", [(40, 3, 1), (96, 3, 2), (88, 5, 2), (1040, 7, 1)])
def test_depthwise_kernels_stay_inside_their_buffers(c, k, stride):
    """Guard bands (compute-sanitizer is closed on this pool): every output and the wgrad workspace sit between two
    sentinel regions that must come back untouched, on odd sizes that leave partial chunks at the row ends."""
    import ctypes as C
    from azure_hc_intel_tf_b200.ops import conv_umma
    from azure_hc_intel_tf_b200.ops.b200 import _check, _ptr, _stream, lib
    n, h, w = 2, 13, 11
    pad = ((k // 2, k // 2), (k // 2, k // 2))
    m = M(c, k, stride, pad, (h, w))
    g = conv_umma.geom((n, c, h, w), m)
    p, q = m.out_hw
    x = rnd(n, h, w, c, 21).contiguous(memory_format=torch.channels_last)
    dy = rnd(n, p, q, c, 22).contiguous(memory_format=torch.channels_last)
    wt = (torch.randn(c, 1, k, k) / k).to(DEV).to(torch.bfloat16)
    guard = 4096

    def banded(numel, dtype, fill):
        buf = torch.full((numel + 2 * guard,), fill, dtype=dtype, device=DEV)
        return buf, buf[guard:guard + numel]

    def intact(buf, numel, fill):
        return bool((buf[:guard] == fill).all()) and bool((buf[guard + numel:] == fill).all())

    ybuf, y = banded(n * p * q * c, torch.bfloat16, 7.0)
    _check(lib().b200_dwconv_fprop(_ptr(x), _ptr(wt), _ptr(y), C.byref(g), _stream()), "dwconv_fprop")
    dxbuf, dx = banded(n * h * w * c, torch.bfloat16, 7.0)
    _check(lib().b200_dwconv_dgrad(_ptr(dy), _ptr(wt), _ptr(dx), C.byref(g), _stream()), "dwconv_dgrad")
    ws_bytes = lib().b200_dwconv_wgrad_workspace_bytes(C.byref(g))
    wsbuf, ws = banded(ws_bytes // 4, torch.float32, 7.0)
    dwbuf, dw = banded(c * k * k, torch.float32, 7.0)
    _check(lib().b200_dwconv_wgrad(_ptr(x), _ptr(dy), _ptr(dw), C.byref(g), _ptr(ws), _stream()), "dwconv_wgrad")
    torch.cuda.synchronize()
    assert intact(ybuf, y.numel(), 7.0) and intact(dxbuf, dx.numel(), 7.0)
    assert intact(wsbuf, ws.numel(), 7.0) and intact(dwbuf, dw.numel(), 7.0)
    # and the results are the same as through the allocating wrappers
    assert torch.equal(y.view(n, p, q, c).permute(0, 3, 1, 2), conv_umma.dw_fprop(x, wt, m))
    assert torch.equal(dx.view(n, h, w, c).permute(0, 3, 1, 2), conv_umma.dw_dgrad(dy, wt, m, x.shape))
    assert torch.equal(dw.view(c, 1, k, k), conv_umma.dw_wgrad(x, dy, m).reshape(c, 1, k, k))
