"""CPU checks of the reporting tools: the roofline arithmetic (BASELINE.json: "each fused path is reported as achieved
fraction of its roofline") and the HOROVOD_TIMELINE switch (run-tf-sing-ucx-openmpi.sh:104-106 forwards Horovod's
environment to every rank)."""
import json

from azure_hc_intel_tf_b200.bench import roofline
from azure_hc_intel_tf_b200.utils import flags, lr_schedule


def test_link_bytes_per_direction():
    s = 1 << 30
    assert roofline.link_bytes("twoshot", s, 8) == s * 7 / 8
    assert roofline.link_bytes("nvls", s, 8) == s * (1 + 1 / 8)
    assert roofline.link_bytes("oneshot", s, 4) == 3 * s
    # NVLS moves fewer bytes than two-shot's two phases together (2 * 7/8 S): that is its point
    assert roofline.link_bytes("nvls", s, 8) < 2 * roofline.link_bytes("twoshot", s, 8)


def test_roofline_report_from_files(tmp_path):
    osu = tmp_path / "osu"
    osu.mkdir()
    (osu / "p2p_2gpu.json").write_text(json.dumps({"rows": [{"push_gbs": 700.0, "pull_gbs": 650.0}]}))
    (osu / "allreduce_8gpu.json").write_text(json.dumps({"ranks": 8, "rows": [
        {"dtype": "float32", "bytes": 1 << 26, "twoshot_us": 250.0, "nvls_us": 160.0, "nccl_us": 270.0}]}))
    conv = tmp_path / "kb.json"
    conv.write_text(json.dumps({"model": "resnet50", "rows": [
        {"layer": "64x56x56->64 k3x3 s1", "count": 3, "flops": 1.0, "roof_us": 42.6, "fprop": 90.0, "fprop_lib": 57.0}]}))
    out = tmp_path / "r.md"
    assert roofline.main(["--conv", str(conv), "--osu", str(osu), "--md", str(out)]) == 0
    text = out.read_text()
    assert "NVLink 700 GB/s" in text and "| 64x56x56->64 k3x3 s1 | fprop | 90.0 | 42.6 | 0.47 |" in text
    assert "| 67108864 | nvls | 160.0 |" in text


def test_horovod_timeline_env_and_schedule_flags(monkeypatch):
    monkeypatch.setenv("HOROVOD_TIMELINE", "/tmp/tl.json")
    p = flags.parse(["--model=lenet", "--variable_update=horovod", "--device=cpu", "--lr_schedule=warmup_poly",
                     "--num_learning_rate_warmup_epochs=5"])
    assert p.horovod_timeline == "/tmp/tl.json" and p.use_cuda_graph is None and p.lr_schedule == "warmup_poly"
    f = lr_schedule.make("warmup_poly", 3.2, 8192, warmup_epochs=5, total_epochs=90)
    spe = lr_schedule.steps_per_epoch(8192)
    assert f(0) < 0.01 and abs(f(int(5 * spe)) - 3.2) < 0.05 and f(int(89.9 * spe)) < 0.01
    g = lr_schedule.make("piecewise", 0.1, 256)
    assert g(0) == 0.1 and abs(g(int(31 * lr_schedule.steps_per_epoch(256))) - 0.01) < 1e-9
    assert lr_schedule.make("constant", 0.5, 256)(12345) == 0.5
