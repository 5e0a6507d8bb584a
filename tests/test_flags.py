"""The 19 flags of the reference's TF_ARGS block (run-tf-sing-ucx-openmpi.sh:62-81) parse verbatim."""
import pytest

from azure_hc_intel_tf_b200.utils import flags

REFERENCE_ARGV = [
    "--batch_size=64", "--num_warmup_batches=50", "--num_batches=100", "--model=resnet50",
    "--num_intra_threads=11", "--num_inter_threads=2", "--kmp_blocktime=1",
    "--kmp_affinity=granularity=fine,noverbose,compact,1,0", "--display_every=10", "--data_format=NCHW",
    "--optimizer=momentum", "--forward_only=False", "--device=cpu", "--mkl=TRUE", "--variable_update=horovod",
    "--horovod_device=cpu", "--local_parameter_device=cpu",
    "--data_dir=/mnt/shared/tensorflow/ilsvrc2012_tfrecords_20of1024", "--data_name=imagenet"]


def test_reference_argv_parses():
    p = flags.parse(REFERENCE_ARGV)
    assert p.batch_size == 64 and p.num_warmup_batches == 50 and p.num_batches == 100
    assert p.model == "resnet50" and p.optimizer == "momentum" and p.variable_update == "horovod"
    assert p.forward_only is False and p.mkl is True and p.display_every == 10
    assert p.data_dir is None and any("synthetic" in n for n in p.notes)
    assert p.fabric == "gloo" and p.kernels == "cudnn"      # --device=cpu


def test_every_reference_flag_is_declared():
    dests = {a.dest for a in flags.build_parser()._actions}
    assert set(flags.REFERENCE_FLAGS) <= dests and len(flags.REFERENCE_FLAGS) == 19


@pytest.mark.parametrize("v,exp", [("TRUE", True), ("False", False), ("true", True), ("0", False), ("1", True)])
def test_absl_bool(v, exp):
    assert flags.absl_bool(v) is exp


def test_unsupported_variable_update_rejected():
    with pytest.raises(SystemExit):
        flags.parse(["--variable_update=parameter_server"])


def test_fabric_aliases():
    assert flags.parse(["--fabric=ib"]).fabric == "nvl"
    assert flags.parse(["--fabric=sock"]).fabric == "gloo"


def test_fusion_threshold_env(monkeypatch):
    p = flags.parse([])
    monkeypatch.delenv("HOROVOD_FUSION_THRESHOLD", raising=False)
    assert flags.resolve_fusion_threshold(p) == 134217728
    monkeypatch.setenv("HOROVOD_FUSION_THRESHOLD", "1048576")
    assert flags.resolve_fusion_threshold(p) == 1048576
    p.fusion_threshold = 4096
    assert flags.resolve_fusion_threshold(p) == 4096
