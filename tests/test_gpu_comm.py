"""NVLink fabric: hand-written P2P allreduce / fused allreduce+SGD / broadcast kernels on 2+ GPUs, against the
arithmetic answer, against NCCL, and against a single-process run (SURVEY.md §4 "GPU" bullet; K-AR1/2/4, K-BC,
K-BAR). Reference role: Horovod's MPI_Allreduce on the fusion buffer (run-tf-sing-ucx-openmpi.sh:105) + TF
ApplyMomentum (:73)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _ngpu():
    return torch.cuda.device_count() if torch.cuda.is_available() else 0


needs2 = pytest.mark.skipif(_ngpu() < 2, reason="needs at least 2 GPUs")


@needs2
def test_p2p_allreduce_matches_nccl_and_is_rank_consistent():
    from azure_hc_intel_tf_b200.utils import testing, testing_workers as W
    world = min(_ngpu(), 8)
    sizes = [1024, 4096, 100_000, 1 << 20, (1 << 22) + 4096]
    outs = testing.run_ranks(W.w_nvl_allreduce, world, "nvl", args=(sizes, 3), timeout=600)
    for rank_out in outs:
        for dtype, n, ok, same in rank_out:
            assert ok and same, (dtype, n, ok, same)


@needs2
def test_flag_protocol_stress_with_random_delays():
    """No host synchronisation, random per-rank delays, sizes hopping across the algorithm switch points."""
    from azure_hc_intel_tf_b200.utils import testing, testing_workers as W
    world = min(_ngpu(), 8)
    outs = testing.run_ranks(W.w_nvl_flag_stress, world, "nvl", args=(3000, 7), timeout=600)
    assert outs == [0] * world, outs


@needs2
@pytest.mark.parametrize("lars", [False, True, "rmsprop", "adam"])
def test_fused_allreduce_sgd_training_matches_single_process(lars):
    from azure_hc_intel_tf_b200.utils import testing, testing_workers as W
    world = 2
    outs = testing.run_ranks(W.w_nvl_fused_training, world, "nvl", args=(4, lars), timeout=600)
    (w0, nb, used0), (w1, _, used1) = outs
    assert used0 and used1 and nb > 1
    for a, b in zip(w0, w1):
        assert torch.equal(a, b)
    ref = W.single_process_reference(world, lars if isinstance(lars, str) else ("lars" if lars else "fused"), 4)
    # Adam's first steps move every weight by about lr * sign(g): rounding-level differences in near-zero gradients
    # (GPU vs CPU convolution, reduction order) are amplified to a fraction of lr = 0.01, so its bound is on that scale
    tol = dict(atol=2e-3, rtol=1e-2) if lars == "adam" else dict(atol=5e-6, rtol=2e-4)
    for a, r in zip(w0, ref):
        assert torch.allclose(a, r, **tol), float((a - r).abs().max())
        if lars == "adam":
            assert float((a - r).abs().median()) < 2e-5


@needs2
@pytest.mark.parametrize("algo,lars", [("twoshot", False), ("nvls", False), ("auto", False), ("twoshot", True), ("nvls", True)])
def test_fused_update_on_large_buckets_is_rank_consistent_and_matches_unfused(algo, lars):
    """Buckets of 1.5 / 3 / 9.4 MiB-class tensors (above the one-shot range: the two-shot and NVLS kernels real models
    use), one rank delayed per launch, no host synchronisation between buckets. A fused update that touches a block
    before its all-gather has landed shows up as replicas that differ or as a mismatch against the unfused result."""
    from azure_hc_intel_tf_b200.utils import testing, testing_workers as W
    world = min(_ngpu(), 8)
    sizes = [2_359_296, 393_216, 1_048_576 + 512, 2_359_296, 65_536, 1000, 589_824]
    outs = testing.run_ranks(W.w_nvl_fused_buckets, world, "nvl", args=(sizes, algo, lars, 3), timeout=600)
    for o in outs:
        if isinstance(o, str) and o.startswith("skip"):
            pytest.skip(o)
        assert o == [], o


@needs2
def test_allgather_p2p_and_host_collectives_on_own_kernels():
    from azure_hc_intel_tf_b200.utils import testing, testing_workers as W
    world = min(_ngpu(), 8)
    outs = testing.run_ranks(W.w_nvl_misc, world, "nvl", args=(), timeout=600)
    for o in outs:
        assert o == [], o
