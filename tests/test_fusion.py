"""Static fusion planner + readiness tracker (C++: csrc/core/hvd_core.cpp). The reference's counterpart is
Horovod's fusion buffer with HOROVOD_FUSION_THRESHOLD=134217728 (run-tf-sing-ucx-openmpi.sh:105)."""
import pytest
from hypothesis import given, settings, strategies as st

from azure_hc_intel_tf_b200.hvd.fusion import ALIGN_ELEMS, DEFAULT_THRESHOLD, FusionPlan


def test_default_threshold_is_the_references():
    assert DEFAULT_THRESHOLD == 134217728


def test_resnet50_fits_one_bucket():
    from azure_hc_intel_tf_b200.models import create_model
    m = create_model("resnet50")
    numels = [p.numel() for p in m.parameters() if p.requires_grad][::-1]
    plan = FusionPlan(numels, [4] * len(numels), [0] * len(numels), [1] * len(numels))
    assert len(plan.buckets) == 1 and len(plan.segments) == 161
    assert plan.buckets[0].numel * 4 <= DEFAULT_THRESHOLD


def test_vgg16_fc6_is_chunked():
    # 25088x4096 fp32 = 411 MB > 128 MiB: Horovod would send it unfused; the planner splits it
    numels = [4096 * 1001, 4096, 4096 * 4096, 4096, 25088 * 4096, 512 * 512 * 9]
    plan = FusionPlan(numels, [4] * 6, [0] * 6, [1] * 6)
    segs = [s for s in plan.segments if s.tensor == 4]
    assert len(segs) >= 4
    assert sum(s.numel for s in segs) == 25088 * 4096
    for b in plan.buckets:
        assert b.numel * b.elem_size <= DEFAULT_THRESHOLD
    dt, start, n = plan.tensor_range(4)     # still one contiguous flat range
    assert n == 25088 * 4096


def test_dtype_segregation():
    plan = FusionPlan([100, 100, 100], [4, 2, 4], [0, 1, 0], [0, 0, 0])
    assert [b.dtype_id for b in plan.buckets] == [0, 1, 0]


def test_hash_is_deterministic_and_sensitive():
    a = FusionPlan([10, 20, 30], [4] * 3, [0] * 3, [1] * 3)
    b = FusionPlan([10, 20, 30], [4] * 3, [0] * 3, [1] * 3)
    c = FusionPlan([10, 21, 30], [4] * 3, [0] * 3, [1] * 3)
    assert a.hash == b.hash and a.hash != c.hash


def test_rejects_bad_input():
    with pytest.raises(ValueError):
        FusionPlan([0], [4], [0], [0])
    with pytest.raises(ValueError):
        FusionPlan([10], [4], [0], [0], align_elems=3)


@settings(max_examples=60, deadline=None)
@given(st.lists(st.integers(1, 50000), min_size=1, max_size=40), st.integers(12, 18))
def test_plan_invariants(numels, log2thr):
    thr = 1 << log2thr
    plan = FusionPlan(numels, [4] * len(numels), [0] * len(numels), [1] * len(numels), thr, 64)
    covered = [0] * len(numels)
    for s in plan.segments:
        assert s.bucket_offset % 64 == 0
        assert s.bucket_offset + s.numel <= plan.buckets[s.bucket].numel
        assert s.tensor_offset == covered[s.tensor]
        covered[s.tensor] += s.numel
    assert covered == list(numels)
    cap = max(64, thr // 4 // 64 * 64)
    for b in plan.buckets:
        assert b.numel % 64 == 0 and b.numel <= cap
    order = [s.tensor for s in plan.segments]
    assert order == sorted(order)                      # readiness order is preserved
    for t in range(len(numels)):
        assert plan.tensor_range(t)[2] == numels[t]   # contiguous in the flat layout


def test_tracker_counts_down_and_rearms():
    plan = FusionPlan([1000, 1000, 1000, 1000], [4] * 4, [0] * 4, [1] * 4, threshold_bytes=2 * ALIGN_ELEMS * 4)
    assert len(plan.buckets) == 2
    tr = plan.tracker()
    tr.begin_step()
    assert tr.ready(0) == [] and tr.pending() == 2
    assert tr.ready(1) == [0] and tr.pending() == 1
    with pytest.raises(RuntimeError):
        tr.ready(1)
    assert tr.ready(3) == [] and tr.ready(2) == [1] and tr.pending() == 0
    assert tr.oldest_inflight_seconds() >= 0.0
    tr.done(0), tr.done(1)
    assert tr.oldest_inflight_seconds() == 0.0
    tr.begin_step()
    assert tr.pending() == 2 and tr.ready(0) == []


def test_tail_bucket_is_cut_at_a_tensor_boundary():
    """The last bucket's collective cannot overlap with backward, so the planner can cut the end of the gradient
    stream off into a small bucket of its own (fabric nvl default: 1 MiB)."""
    from azure_hc_intel_tf_b200.hvd.fusion import FusionPlan
    numels = [1000, 5000, 300000, 2000000, 100000, 64, 9408]
    plain = FusionPlan(numels, [4] * 7, [0] * 7, [1] * 7, 4 << 20, 1024, 0)
    tail = FusionPlan(numels, [4] * 7, [0] * 7, [1] * 7, 4 << 20, 1024, 1 << 20)
    assert len(tail.buckets) == len(plain.buckets) + 1
    last = tail.buckets[-1]
    segs = [s for s in tail.segments if s.bucket == len(tail.buckets) - 1]
    assert [s.tensor for s in segs] == [4, 5, 6] and all(s.tensor_offset == 0 for s in segs)
    assert last.numel * 4 <= 1 << 20
    assert tail.total_elems[0] >= plain.total_elems[0]
    # every tensor is still one contiguous range of the flat buffer
    for t in range(7):
        _, start, n = tail.tensor_range(t)
        assert n == numels[t] and start % 1024 == 0
    assert tail.hash != plain.hash
