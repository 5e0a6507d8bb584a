"""Topology discovery degrades gracefully without a GPU (the reference's counterpart is lscpu / ibv_devinfo probing,
run-tf-sing-ucx-openmpi.sh:37-38, prep-cluster.sh:23)."""
import json

from azure_hc_intel_tf_b200.parallel import topology


def test_discover_without_gpu_is_safe():
    t = topology.discover()
    assert isinstance(t.report(), list) and json.loads(t.to_json())["gpus"] is not None


def test_numa_lookup_handles_unknown_bus():
    assert topology._numa_node("00000000:FF:1F.7") in (-1, 0, 1, 2, 3)


def test_report_formats_a_synthetic_box():
    t = topology.Topology(gpus=[topology.GpuInfo(i, "NVIDIA B200", f"0000:{i:02x}:00.0", i // 4, 179.0, 18, 18, True,
                                                 [j for j in range(8) if j != i]) for i in range(8)],
                          driver="580", all_to_all_p2p=True, nvlink_healthy=True)
    rep = "\n".join(t.report())
    assert "GPUs: 8" in rep and "nvlink 18/18 up" in rep and t.numa_of(5) == 1
