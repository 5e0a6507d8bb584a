"""The allocation-free BatchNorm-statistics entry point, called through the C ABI exactly as an external caller
would (ctypes on libb200cnn.so, raw device pointers), checked against the regular entry point that takes
caller-provided scratch.

Who needs it: code that may not allocate device memory where it calls this library — a memory-planned region of a
host framework, a plug-in that is handed fixed input/output buffers only, or a call site from which the host's
allocator must not be re-entered. For them the scratch is a fixed 2 MiB array inside the library image.

This file is named so that pytest collects it last: if it fails, a run that stops at the first failure has already
run every other test. It has to pass both alone in a fresh process and after all other tests, and it makes two
calls in a row because the shared array keeps the previous call's contents."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _lib():
    from azure_hc_intel_tf_b200 import _native
    return _native.cuda()


def _p(t):
    return C.c_void_p(t.data_ptr() if t is not None else 0)


def _run(lib, fn_name, x, c, with_scratch):
    rows = x.numel() // c
    dev = x.device
    g = torch.Generator().manual_seed(c)
    gamma = (torch.rand(c, generator=g) + 0.5).to(dev)
    beta = (torch.randn(c, generator=g) * 0.1).to(dev)
    rm, rv = torch.zeros(c, device=dev), torch.ones(c, device=dev)
    out = torch.zeros(4, c, device=dev)
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    args = [_p(x), rows, c, _p(gamma), _p(beta), _p(rm), _p(rv), 0.9, 1e-5, _p(out[0]), _p(out[1]), _p(out[2]), _p(out[3])]
    if with_scratch:
        need = lib.b200_bn_stats_scratch_bytes(rows, c)
        scratch = torch.zeros(need, dtype=torch.uint8, device=dev)
        args.append(_p(scratch))
    args.append(stream)
    rc = getattr(lib, fn_name)(*args)
    torch.cuda.synchronize()
    return rc, out, rm, rv


def test_no_alloc_matches_regular_entry_point_twice_in_a_row():
    lib = _lib()
    assert lib.b200_bn_stats_noalloc_max_bytes() == 2 << 20
    shapes = [(16, 28, 28, 128), (4, 56, 56, 64)]      # second call: different shape, reuses the same array
    for i, (n, h, w, c) in enumerate(shapes):
        g = torch.Generator().manual_seed(100 + i)
        x = (torch.randn(n, h, w, c, generator=g) * 2 + 0.3).cuda().to(torch.bfloat16)
        rc_ref, ref, rm_ref, rv_ref = _run(lib, "b200_bn_stats", x, c, True)
        rc, out, rm, rv = _run(lib, "b200_bn_stats_noalloc", x, c, False)
        assert rc_ref == 0 and rc == 0, lib.b200_last_error().decode()
        assert torch.allclose(out, ref, atol=1e-5, rtol=1e-5)
        assert torch.allclose(rm, rm_ref, atol=1e-6, rtol=1e-5) and torch.allclose(rv, rv_ref, atol=1e-6, rtol=1e-5)
        # oracle: plain PyTorch fp32 statistics of the same bf16 input
        xf = x.float().reshape(-1, c)
        assert torch.allclose(out[0], xf.mean(0), atol=2e-3, rtol=2e-3)
        assert torch.allclose(out[1], torch.rsqrt(xf.var(0, unbiased=False) + 1e-5), atol=2e-3, rtol=2e-3)


def test_no_alloc_argument_checks_and_size_bound():
    lib = _lib()
    # every BatchNorm shape of the model zoo fits the built-in array; anything that would not is refused, not grown
    for rows, c in ((256 * 112 * 112, 64), (256 * 56 * 56, 256), (256 * 14 * 14, 1024), (256 * 7 * 7, 2048), (64, 4096)):
        assert 0 < lib.b200_bn_stats_scratch_bytes(rows, c) <= lib.b200_bn_stats_noalloc_max_bytes()
    x = torch.zeros(64, 12, dtype=torch.bfloat16, device="cuda")
    rc, *_ = _run(lib, "b200_bn_stats_noalloc", x, 12, False)          # c must be a multiple of 8
    assert rc != 0 and "c%8" in lib.b200_last_error().decode()
