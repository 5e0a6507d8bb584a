"""tcgen05 implicit-GEMM convolution / GEMM kernels vs F.conv2d / matmul in fp32 from the same bf16 inputs
(K1–K3, K8)."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu
DEV = "cuda"


class M:
    def __init__(self, cin, cout, k, stride, pad, in_hw):
        self.cin, self.cout = cin, cout
        self.k, self.stride = (k, k) if isinstance(k, int) else k, (stride, stride)
        self.pad = pad
        h, w = in_hw
        self.out_hw = ((h + sum(pad[0]) - self.k[0]) // stride + 1, (w + sum(pad[1]) - self.k[1]) // stride + 1)

    def extra_repr(self):
        return f"{self.cin}->{self.cout} k={self.k} s={self.stride} pad={self.pad}"


def rnd(shape_nhwc, seed, scale=1.0):
    g = torch.Generator().manual_seed(seed)
    n, h, w, c = shape_nhwc
    return (torch.randn(n, h, w, c, generator=g) * scale).to(DEV).to(torch.bfloat16).permute(0, 3, 1, 2)


def wrnd(k, c, r, s, seed):
    g = torch.Generator().manual_seed(seed)
    w = torch.randn(k, r, s, c, generator=g) * (1.0 / (c * r * s) ** 0.5)
    return w.to(DEV).to(torch.bfloat16).permute(0, 3, 1, 2)      # [K,C,R,S] logical, KRSC storage


def ref_conv(x, w, m):
    (pt, pb), (pl, pr) = m.pad
    return F.conv2d(F.pad(x.float(), (pl, pr, pt, pb)), w.float(), None, m.stride)


def rel_err(a, b):
    return ((a.float() - b.float()).norm() / (b.float().norm() + 1e-12)).item()


CASES = [
    # (n, h, w, cin, cout, k, stride, pad)
    (2, 56, 56, 64, 64, 1, 1, ((0, 0), (0, 0))),
    (2, 56, 56, 64, 256, 1, 1, ((0, 0), (0, 0))),
    (3, 56, 56, 64, 64, 3, 1, ((1, 1), (1, 1))),
    (2, 28, 28, 128, 128, 3, 1, ((1, 1), (1, 1))),
    (2, 14, 14, 256, 256, 3, 1, ((1, 1), (1, 1))),
    (5, 7, 7, 512, 512, 3, 1, ((1, 1), (1, 1))),
    (2, 56, 56, 256, 128, 1, 2, ((0, 0), (0, 0))),
    (2, 56, 56, 256, 512, 1, 2, ((0, 0), (0, 0))),
    (3, 7, 7, 512, 2048, 1, 1, ((0, 0), (0, 0))),
    (2, 28, 28, 128, 128, 3, 2, ((1, 1), (1, 1))),      # ResNet v1.5 strided 3x3
    (2, 17, 17, 128, 192, (1, 7), 1, ((0, 0), (3, 3))),  # Inception 1x7
    (2, 17, 17, 128, 192, (7, 1), 1, ((3, 3), (0, 0))),
    (2, 35, 35, 64, 96, 3, 1, ((1, 1), (1, 1))),
    (2, 224, 224, 64, 64, 3, 1, ((1, 1), (1, 1))),      # VGG conv1_2
    # thin output side (fewer than 64 output channels): the 64-wide boxes are zero-filled / clipped past the edge
    (2, 35, 35, 192, 48, 1, 1, ((0, 0), (0, 0))),       # Inception-v3 5x5 reduce
    (2, 28, 28, 192, 16, 1, 1, ((0, 0), (0, 0))),       # GoogLeNet 5x5 reduce
    (2, 14, 14, 512, 24, 1, 1, ((0, 0), (0, 0))),       # GoogLeNet: Cout % 16 != 0
    # thin input side (fewer than 64 input channels, not the first layer): half-empty k-blocks through TMA zero-fill
    (2, 149, 149, 32, 32, 3, 1, ((0, 0), (0, 0))),      # Inception-v3 stem Conv2d_2a
    (2, 147, 147, 32, 64, 3, 1, ((1, 1), (1, 1))),      # Inception-v3 stem Conv2d_2b
    (2, 35, 35, 48, 64, 5, 1, ((2, 2), (2, 2))),        # Inception-v3 5x5 branch
    (2, 28, 28, 16, 32, 5, 1, ((2, 2), (2, 2))),        # GoogLeNet 5x5 branch
    # CTA-pair (cta_group::2) mode with an odd number of 128-row tiles: the last pair's second CTA has no rows
    (3, 14, 14, 512, 256, 1, 1, ((0, 0), (0, 0))),      # fprop in pair mode (8 k-blocks, 256-wide tile), 5 M tiles
    (3, 14, 14, 256, 512, 1, 1, ((0, 0), (0, 0))),      # dgrad in pair mode, 5 M tiles
]


@pytest.mark.parametrize("n,h,w,cin,cout,k,stride,pad", CASES)
def test_fprop(n, h, w, cin, cout, k, stride, pad):
    from azure_hc_intel_tf_b200.ops import conv_umma
    m = M(cin, cout, k, stride, pad, (h, w))
    x, wt = rnd((n, h, w, cin), 1), wrnd(cout, cin, m.k[0], m.k[1], 2)
    assert conv_umma.supports_fprop(x, wt, m)
    y, stats = conv_umma.fprop(x, wt, m, want_stats=True)
    torch.cuda.synchronize()
    ref = ref_conv(x, wt, m)
    assert y.shape == ref.shape
    assert rel_err(y, ref) < 1e-2
    sums = conv_umma.stats_reduce(stats)
    yf = y.float()
    assert torch.allclose(sums[0], yf.sum((0, 2, 3)), rtol=1e-3, atol=1e-2 * yf.abs().sum((0, 2, 3)).max().item() / 100)
    assert torch.allclose(sums[1], (yf * yf).sum((0, 2, 3)), rtol=1e-3, atol=1e-3)


STRIDED_DGRAD = [
    # general strided filters: one stride-1 launch per stride phase (b200_conv_dgrad_strided)
    (2, 56, 56, 128, 128, 3, 2, ((0, 1), (0, 1))),      # TF 'SAME' on an even size: asymmetric padding
    (2, 35, 35, 288, 384, 3, 2, ((0, 0), (0, 0))),      # Inception-v3 Mixed_6a reduction (VALID)
    (2, 17, 17, 192, 320, 3, 2, ((0, 0), (0, 0))),      # Inception-v3 Mixed_7a
    (2, 27, 27, 64, 192, 5, 2, ((2, 2), (2, 2))),       # 5x5 / 2
    (2, 30, 30, 64, 64, 7, 2, ((3, 3), (3, 3))),        # 7x7 / 2 with 64 input channels
    (2, 20, 20, 64, 128, 3, 3, ((1, 1), (1, 1))),       # stride 3: nine phases, some with a single tap
    (2, 16, 16, 64, 128, 1, 2, ((1, 1), (1, 1))),       # padded strided 1x1: three of four phases have no tap (zeros)
    (2, 14, 14, 256, 256, (3, 1), 2, ((1, 1), (0, 0))),
]


@pytest.mark.parametrize("n,h,w,cin,cout,k,stride,pad", CASES + STRIDED_DGRAD)
def test_dgrad(n, h, w, cin, cout, k, stride, pad):
    from azure_hc_intel_tf_b200.ops import conv_umma
    m = M(cin, cout, k, stride, pad, (h, w))
    x, wt = rnd((n, h, w, cin), 1), wrnd(cout, cin, m.k[0], m.k[1], 2)
    dy = rnd((n, m.out_hw[0], m.out_hw[1], cout), 3)
    assert conv_umma.supports_dgrad(dy, wt, m, x.shape)
    dx = conv_umma.dgrad(dy, wt, m, x.shape)
    torch.cuda.synchronize()
    xf = x.float().requires_grad_()
    ref_conv(xf, wt, m).backward(dy.float())
    assert rel_err(dx, xf.grad) < 1e-2


@pytest.mark.parametrize("mm,nn,kk", [(256, 1001, 2048), (64, 4096, 25088), (256, 4096, 4096), (100, 16, 64),
                                      (300, 64, 152), (64, 1001, 1056), (130, 72, 8)])   # partial last k-block
def test_gemm_nt(mm, nn, kk):
    from azure_hc_intel_tf_b200.ops import conv_umma
    g = torch.Generator().manual_seed(0)
    a = (torch.randn(mm, kk, generator=g)).to(DEV).to(torch.bfloat16)
    b = (torch.randn(nn, kk, generator=g) / kk ** 0.5).to(DEV).to(torch.bfloat16)
    out = conv_umma.gemm_nt(a, b, out_fp32=True)
    torch.cuda.synchronize()
    ref = a.float() @ b.float().t()
    assert rel_err(out, ref) < 2e-3
    out16 = conv_umma.gemm_nt(a, b)
    assert rel_err(out16, ref) < 1e-2


@pytest.mark.parametrize("n,h,w,cin,cout,k,stride,pad", CASES + [
    (16, 56, 56, 64, 64, 3, 1, ((1, 1), (1, 1))),
    (16, 56, 56, 256, 64, 1, 1, ((0, 0), (0, 0))),
    (8, 14, 14, 1024, 256, 1, 1, ((0, 0), (0, 0))),
    (8, 28, 28, 512, 1024, 1, 2, ((0, 0), (0, 0))),
])
def test_wgrad(n, h, w, cin, cout, k, stride, pad):
    from azure_hc_intel_tf_b200.ops import conv_umma
    m = M(cin, cout, k, stride, pad, (h, w))
    x, wt = rnd((n, h, w, cin), 1), wrnd(cout, cin, m.k[0], m.k[1], 2)
    dy = rnd((n, m.out_hw[0], m.out_hw[1], cout), 3, 0.1)
    assert conv_umma.supports_wgrad(x, dy, m)
    dw = conv_umma.wgrad(x, dy, m)
    torch.cuda.synchronize()
    wf = wt.float().requires_grad_()
    ref_conv(x, wf, m).backward(dy.float())
    assert dw.shape == wf.grad.shape
    assert rel_err(dw, wf.grad) < 5e-3
    # writing into a caller-provided fp32 view (the fusion-buffer path) gives the same result
    out = torch.full((cout, m.k[0], m.k[1], cin), 7.0, device=DEV).permute(0, 3, 1, 2)
    conv_umma.wgrad(x, dy, m, out=out)
    assert rel_err(out, wf.grad) < 5e-3


@pytest.mark.parametrize("b,i,o", [(256, 2048, 1001), (64, 25088, 4096), (32, 4096, 4096)])
def test_dense_layer_gemms(b, i, o):
    from azure_hc_intel_tf_b200.ops import conv_umma
    g = torch.Generator().manual_seed(0)
    x = torch.randn(b, i, generator=g).to(DEV).to(torch.bfloat16)
    w = (torch.randn(o, i, generator=g) / i ** 0.5).to(DEV).to(torch.bfloat16)
    ld = (o + 7) // 8 * 8
    dy = torch.zeros(b, ld, device=DEV, dtype=torch.bfloat16)
    dy[:, :o] = (torch.randn(b, o, generator=g) * 0.1).to(DEV).to(torch.bfloat16)
    dw = conv_umma.dense_wgrad(x, dy, o)
    ref_dw = dy[:, :o].float().t() @ x.float()
    assert rel_err(dw, ref_dw) < 5e-3
    if i % 64 == 0:
        dx = conv_umma.gemm_nn(dy, w, k=o)
        ref_dx = dy[:, :o].float() @ w.float()
        assert rel_err(dx, ref_dx) < 1e-2


@pytest.mark.parametrize("n,h,w,cout,k,stride,pad", [
    (4, 224, 224, 64, 7, 2, ((3, 3), (3, 3))),      # ResNet conv0
    (2, 224, 224, 64, 3, 1, ((1, 1), (1, 1))),      # VGG conv1_1
    (2, 299, 299, 32, 3, 2, ((0, 0), (0, 0))),      # Inception-v3 stem
    (2, 227, 227, 64, 11, 4, ((0, 0), (0, 0))),     # AlexNet
])
def test_first_layer_patch_matrix_path(n, h, w, cout, k, stride, pad):
    from azure_hc_intel_tf_b200.ops import conv_umma
    m = M(3, cout, k, stride, pad, (h, w))
    x, wt = rnd((n, h, w, 3), 1), wrnd(cout, 3, m.k[0], m.k[1], 2)
    assert conv_umma.supports_first_layer(x, m)
    y, stats, patches = conv_umma.first_layer_fprop(x, wt, m, True)
    ref = ref_conv(x, wt, m)
    assert y.shape == ref.shape and rel_err(y, ref) < 1e-2
    from azure_hc_intel_tf_b200.ops.b200 import lib
    sums = stats.sum(0)
    assert torch.allclose(sums[0], y.float().sum((0, 2, 3)), rtol=2e-3, atol=1.0)
    dy = rnd((n, m.out_hw[0], m.out_hw[1], cout), 3, 0.1)
    dw = conv_umma.first_layer_wgrad(patches, dy, m)
    wf = wt.float().requires_grad_()
    ref_conv(x, wf, m).backward(dy.float())
    assert dw.shape == wf.grad.shape and rel_err(dw, wf.grad) < 5e-3
