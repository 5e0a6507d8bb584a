"""Launcher arithmetic and CLI surface — table-driven against the reference's
benchmark-scripts/run-tf-sing-ucx-openmpi.sh:40-50."""
import io
import os

import pytest

from azure_hc_intel_tf_b200.parallel import launcher


def ref_arith(nodes, wps, sockets, cps, inter=2):
    if wps == 0:
        cpw, wpn = cps * sockets, 1
    else:
        cpw, wpn = cps // wps, wps * sockets
    return wpn, cpw, cpw // inter, nodes * wpn


@pytest.mark.parametrize("nodes,wps,sockets,cps", [
    (4, 1, 2, 22), (2, 1, 2, 22), (1, 0, 2, 22), (1, 2, 2, 22), (1, 4, 2, 48), (8, 11, 2, 22), (1, 1, 1, 8),
    (3, 0, 4, 16), (1, 4, 2, 56)])
def test_worker_arithmetic(nodes, wps, sockets, cps):
    lay = launcher.worker_layout(nodes, wps, sockets, cps)
    assert (lay.workers_per_node, lay.cores_per_worker, lay.intra_threads, lay.total_workers) == \
        ref_arith(nodes, wps, sockets, cps)


def test_readme_examples():
    # README.md:69-73 — "4 nodes, 8 workers" and "2 nodes, 4 workers" on a 2-socket server
    assert launcher.worker_layout(4, 1, 2, 22).total_workers == 8
    assert launcher.worker_layout(2, 1, 2, 22).total_workers == 4


def test_invalid_layout():
    with pytest.raises(ValueError):
        launcher.worker_layout(0, 1, 2, 22)


def test_banner_matches_reference_keys():
    lay = launcher.worker_layout(4, 1, 2, 22)
    keys = [l.split(":")[0] for l in lay.banner()]
    assert keys == ["TOTAL_NODES", "WORKERS_PER_NODE", "TOTAL_WORKERS", "CORES_PER_WORKER", "OMP_NUM_THREADS",
                    "NUM_INTRA_THREADS", "NUM_INTER_THREADS"]


def test_parse_lscpu():
    txt = "Architecture: x86_64\nSocket(s):             2\nCore(s) per socket:    22\nThread(s) per core: 2\n"
    assert launcher.parse_lscpu(txt) == {"sockets": 2, "cores_per_socket": 22}


def test_benchmark_args_cover_the_19_reference_flags():
    from azure_hc_intel_tf_b200.utils import flags
    lay = launcher.worker_layout(1, 1, 2, 22)
    args = launcher.benchmark_args(64, "resnet50", 100, 50, lay, "nvl", "gpu")
    names = {a.split("=")[0].lstrip("-") for a in args}
    missing = set(flags.REFERENCE_FLAGS) - names - {"data_dir"}   # no dataset on this box
    assert not missing
    assert "--num_warmup_batches=50" in args and "--num_batches=100" in args and "--model=resnet50" in args


def test_dry_run_prints_command():
    buf = io.StringIO()
    rc = launcher.launch(1, 1, 64, "sock", dry_run=True, out=buf, device="cpu")
    s = buf.getvalue()
    assert rc == 0 and "TOTAL_WORKERS" in s and "cnn_benchmarks" in s and "--variable_update=horovod" in s


def test_shell_wrapper_exists_and_is_executable():
    p = os.path.join(os.path.dirname(os.path.dirname(__file__)), "launch", "run-tf-b200.sh")
    assert os.access(p, os.X_OK)


def test_environment_report_runs_without_a_gpu():
    """utils/doctor.py = the reference images' %runscript sanity print (tf-hvd-gcc-ompi-ucx-mlnx.def:45-55) plus the
    update_config.sh limits, checked instead of set. On a CPU-only box it must come out 'ready' for the plumbing path."""
    from azure_hc_intel_tf_b200.utils import doctor
    lines, fail = doctor.collect()
    text = "\n".join(lines)
    for key in ("OS", "python", "g++", "torch", "libhvd_core.so", "libb200cnn.so", "ulimit nofile"):
        assert key in text, key
    import torch
    if not torch.cuda.is_available():
        assert fail == [], fail
        assert "CPU plumbing" in text
