"""Log-line formats (tf_cnn_benchmarks style; --display_every=10 at run-tf-sing-ucx-openmpi.sh:71) and the
Chrome-trace timeline writer."""
import json
import re

from azure_hc_intel_tf_b200.utils import log
from azure_hc_intel_tf_b200.utils.timeline import Timeline


def test_step_line_format():
    line = log.step_line(10, 256, [0.1, 0.1, 0.11, 0.09], 7.9521)
    assert re.fullmatch(r"10\timages/sec: \d+\.\d \+/- \d+\.\d \(jitter = \d+\.\d\)\t7\.952", line)


def test_total_lines():
    lines = log.total_lines(20050.123)
    assert lines[0] == "-" * 64 and lines[1] == "total images/sec: 20050.12" and lines[2] == lines[0]


def test_header_has_tf_cnn_benchmarks_fields():
    h = "\n".join(log.header_lines("resnet50", "imagenet (synthetic)", "training", 512, 64, 100, 0.04,
                                   ["horovod/gpu:0"], "NHWC", "momentum", "horovod"))
    for key in ("Model:", "Dataset:", "Mode:", "Batch size:", "per device", "Num batches:", "Devices:",
                "Data format:", "Optimizer:", "Variables:"):
        assert key in h


def test_log_name_convention():
    assert log.log_name(4, 64, "ib", 1, "real") == "tfmn-4n-64b-real-ib-r1.log"   # reference :9


def test_timeline_is_valid_chrome_trace(tmp_path):
    path = tmp_path / "tl.json"
    tl = Timeline(str(path), 0)
    t0 = tl.now_us()
    tl.event('bucket"0', "LAUNCH", 1, t0, 12.5)
    tl.mark("step", 0)
    tl.close()
    ev = json.loads(path.read_text())
    assert ev[0]["ph"] == "X" and ev[0]["dur"] == 12.5 and ev[0]["name"] == 'bucket"0' and ev[1]["ph"] == "i"
