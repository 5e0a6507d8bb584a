"""The C ABI is the contract between the kernel libraries and everything above them: the headers under include/, the
ctypes signature tables and the symbols the built libraries export must agree (a stale or half-built library is an
error, not a reason to fall back), and the host-callable planning entry points of the conv library must be
self-consistent. Runs without a GPU: the kernel library is cross-compiled and only its host functions are called."""
import ctypes as C
import pathlib
import re

import pytest

ROOT = pathlib.Path(__file__).resolve().parents[1]


def _declared(header: str):
    text = (ROOT / "include" / header).read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b((?:b200|hvd)_[a-z0-9_]+)\s*\(", text)))


def _cuda_lib():
    from azure_hc_intel_tf_b200 import _native
    return _native.cuda()


def test_every_declared_kernel_entry_point_is_exported_and_has_a_signature():
    from azure_hc_intel_tf_b200 import _cuda_sigs
    lib = _cuda_lib()
    known = set(_cuda_sigs.SIGS) | set(_cuda_sigs.OPTIONAL)
    comm_src = (ROOT / "azure_hc_intel_tf_b200" / "parallel" / "symm.py").read_text()
    for header in ("b200cnn.h", "b200conv.h", "b200comm.h"):
        names = _declared(header)
        assert len(names) >= 5, header
        for name in names:
            assert hasattr(lib, name), f"{name} is declared in include/{header} but libb200cnn.so does not export it"
            if header == "b200comm.h":
                assert f"lib.{name}." in comm_src, f"{name}: no ctypes signature in parallel/symm.py"
            else:
                assert name in known, f"{name}: no ctypes signature in _cuda_sigs.py"


def test_signature_tables_name_only_declared_functions():
    from azure_hc_intel_tf_b200 import _cuda_sigs
    declared = set(_declared("b200cnn.h")) | set(_declared("b200conv.h"))
    extra = (set(_cuda_sigs.SIGS) | set(_cuda_sigs.OPTIONAL)) - declared
    assert not extra, f"signatures without a declaration in include/: {sorted(extra)}"


def test_host_runtime_exports_its_header():
    from azure_hc_intel_tf_b200 import _native
    lib = _native.host()
    for name in _declared("hvd_core.h"):
        assert hasattr(lib, name), name


def _geom(n, h, w, c, k, r, s, stride, pad):
    from azure_hc_intel_tf_b200._cuda_sigs import ConvGeom
    p = (h + 2 * pad - r) // stride + 1
    q = (w + 2 * pad - s) // stride + 1
    g = ConvGeom()
    g.n, g.h, g.w, g.c, g.k, g.r, g.s = n, h, w, c, k, r, s
    g.stride_h = g.stride_w = stride
    g.pad_t = g.pad_b = g.pad_l = g.pad_r = pad
    g.p, g.q = p, q
    return g


FPROP, DGRAD, WGRAD = 0, 1, 2


def test_conv_support_matrix():
    lib = _cuda_lib()
    ok = lambda g, op: bool(lib.b200_conv_supported(C.byref(g), op))
    code = lambda g, op: int(lib.b200_conv_supported(C.byref(g), op))
    # ResNet / VGG body layers: everything on the tensor-core path
    for g in (_geom(256, 56, 56, 64, 256, 1, 1, 1, 0), _geom(256, 56, 56, 64, 64, 3, 3, 1, 1),
              _geom(64, 224, 224, 64, 64, 3, 3, 1, 1), _geom(256, 7, 7, 512, 2048, 1, 1, 1, 0)):
        assert ok(g, FPROP) and ok(g, DGRAD) and ok(g, WGRAD)
    # thin output side (1x1 reductions): taken; Cout only has to keep rows 16-byte aligned
    thin = _geom(128, 28, 28, 192, 16, 1, 1, 1, 0)
    assert ok(thin, FPROP) and ok(thin, DGRAD) and ok(thin, WGRAD)
    assert ok(_geom(128, 14, 14, 512, 24, 1, 1, 1, 0), FPROP)
    assert not ok(_geom(128, 14, 14, 512, 20, 1, 1, 1, 0), FPROP)
    # thin input side after the first layer (Inception stem): on the same kernels through TMA zero-fill
    stem = _geom(128, 149, 149, 32, 32, 3, 3, 1, 0)
    assert ok(stem, FPROP) and ok(stem, WGRAD) and ok(stem, DGRAD)
    assert lib.b200_conv_min_channels() == 8
    # strided dgrad: unpadded 1x1 in one launch (1), every other strided filter by stride phases (2)
    assert code(_geom(256, 56, 56, 256, 512, 1, 1, 2, 0), DGRAD) == 1
    assert code(_geom(256, 28, 28, 128, 128, 3, 3, 2, 1), DGRAD) == 2
    assert code(_geom(256, 56, 56, 64, 64, 3, 3, 1, 1), DGRAD) == 1
    assert ok(_geom(256, 28, 28, 128, 128, 3, 3, 2, 1), FPROP) and ok(_geom(256, 28, 28, 128, 128, 3, 3, 2, 1), WGRAD)


@pytest.mark.parametrize("shape", [(256, 56, 56, 64, 256, 1, 1, 1, 0), (256, 56, 56, 64, 64, 3, 3, 1, 1),
                                   (256, 14, 14, 256, 256, 3, 3, 1, 1), (256, 14, 14, 1024, 256, 1, 1, 1, 0),
                                   (3, 14, 14, 512, 256, 1, 1, 1, 0), (2, 7, 7, 512, 2048, 1, 1, 1, 0),
                                   (1, 8, 8, 64, 48, 1, 1, 1, 0)])
def test_conv_statistics_scratch_plan_is_consistent(shape):
    """The fprop epilogue leaves per-CTA BatchNorm partials: 8 scratch rows + 1 folded result row per CTA of the
    persistent grid. The two host functions that describe that layout to the caller must agree with each other and
    with the grid rules (one CTA per SM, whole CTA pairs, never more CTAs than tiles)."""
    lib = _cuda_lib()
    g = _geom(*shape)
    rows, first = lib.b200_conv_stats_rows(C.byref(g)), lib.b200_conv_stats_first_result_row(C.byref(g))
    assert rows > 0 and first > 0 and rows % 9 == 0 and first % 8 == 0
    grid = rows // 9
    assert first == grid * 8
    assert 1 <= grid <= 148
    m_tiles = -(-(g.n * g.p * g.q) // 128)
    assert grid <= max(1, m_tiles) * max(1, -(-g.k // 64)) + 1


def test_wgrad_workspace_is_bounded_and_deterministic_split():
    lib = _cuda_lib()
    for shape in [(256, 56, 56, 64, 64, 3, 3, 1, 1), (256, 14, 14, 256, 1024, 1, 1, 1, 0), (256, 7, 7, 512, 512, 3, 3, 1, 1)]:
        g = _geom(*shape)
        ws = lib.b200_conv_wgrad_workspace_bytes(C.byref(g))
        out_bytes = g.k * g.r * g.s * g.c * 4
        assert ws % out_bytes == 0                      # a whole number of per-slice partials
        assert 0 <= ws <= 96 << 20                      # capped (kWorkspaceCap)
    assert lib.b200_dense_wgrad_workspace_bytes(256, 2048, 1001) >= 0
