"""Distributed plumbing on CPU/gloo, world_size=2 — BASELINE.json configs[0]; the analogue of the
reference's single-node multi-worker mode (run-tf-sing-ucx-openmpi.sh:100 ``--oversubscribe``) over its
``sock`` fabric (:94)."""
import pytest
import torch

from azure_hc_intel_tf_b200.utils import testing, testing_workers as W


def test_api_basics_ws2():
    outs = testing.run_ranks(W.w_basic, 2, "gloo")
    for r, (s, a, g, b, z, obj) in enumerate(outs):
        assert s == [3.0] * 5 and a == [1.5] * 5 and g == 3 and b == [1.0] * 3 and z == [2.0] * 4
        assert obj == {"r": 0}


@pytest.mark.parametrize("mode", ["fused", "generic", "lars", "rmsprop", "adam"])
def test_distributed_optimizer_matches_single_process(mode):
    steps = 4
    outs = testing.run_ranks(W.w_dist_opt, 2, "gloo", args=(mode, steps))
    (w0, nb0), (w1, nb1) = outs
    assert nb0 > 1                                   # several fusion buckets were exercised
    for a, b in zip(w0, w1):
        assert torch.equal(a, b)                     # ranks end bit-identical
    ref = W.single_process_reference(2, mode, steps)
    for a, r in zip(w0, ref):
        assert torch.allclose(a, r, atol=2e-6, rtol=1e-4)


def test_plan_mismatch_is_detected():
    assert testing.run_ranks(W.w_plan_mismatch, 2, "gloo") == ["mismatch", "mismatch"]


def test_cli_resnet50_batch32_ws2_cpu():
    # BASELINE.json configs[0]: ResNet-50, batch 32/worker, world_size=2 on CPU/gloo (plumbing, shortened)
    argv = ["--model=resnet50", "--batch_size=32", "--num_batches=1", "--num_warmup_batches=0", "--device=cpu",
            "--variable_update=horovod", "--display_every=1", "--optimizer=momentum"]
    outs = testing.run_ranks(W.w_cli, 2, "gloo", args=(argv,), timeout=900)
    assert outs[0][0] > 0 and abs(outs[0][1] - outs[1][1]) < 1e-3     # same weights on both ranks


def test_trace_graph_and_tfprof_files_ws2_cpu(tmp_path):
    """--trace_file / --tfprof_file / --graph_file (tf_cnn_benchmarks hooks the reference leaves unset).
    The profiled step is a real training step with its allreduce, so every rank must run it (with only rank 0
    stepping, the ranks' collective counts diverge and the run hangs); rank 0 writes the files."""
    import json
    trace, prof, graph = tmp_path / "trace.json", tmp_path / "tfprof.txt", tmp_path / "graph.txt"
    argv = ["--model=lenet", "--batch_size=4", "--num_batches=2", "--num_warmup_batches=1", "--device=cpu",
            "--variable_update=horovod", "--display_every=1", f"--trace_file={trace}", f"--tfprof_file={prof}",
            f"--graph_file={graph}"]
    outs = testing.run_ranks(W.w_cli, 2, "gloo", args=(argv,), timeout=600)
    assert abs(outs[0][1] - outs[1][1]) < 1e-4              # ranks still in lock-step after the extra profiled step
    assert "traceEvents" in json.load(open(trace))
    text = open(prof).read()
    assert text.startswith("# per-op profile of one training step: model=lenet") and len(text.splitlines()) > 5
    g = open(graph).read()
    assert "v0/cg/conv0/conv2d/kernel HWIO (32, 3, 5, 5) trainable" in g and "v0/cg/affine1/weights IO" in g
    assert "# trainable parameters:" in g


def test_horovod_timeline_env_writes_bucket_events_ws2_cpu(tmp_path, monkeypatch):
    """HOROVOD_TIMELINE=<file> (Horovod's own switch; the reference's launcher forwards such variables with -x,
    run-tf-sing-ucx-openmpi.sh:104-106) produces a Chrome trace with one LAUNCH event per fusion bucket and step."""
    import json
    path = tmp_path / "timeline.json"
    monkeypatch.setenv("HOROVOD_TIMELINE", str(path))
    argv = ["--model=lenet", "--batch_size=4", "--num_batches=2", "--num_warmup_batches=1", "--device=cpu",
            "--variable_update=horovod", "--display_every=1", "--fusion_threshold=65536"]
    testing.run_ranks(W.w_cli, 2, "gloo", args=(argv,), timeout=600)
    ev = json.load(open(path))
    launches = [e for e in ev if e.get("cat") == "LAUNCH" or e.get("args", {}).get("phase") == "LAUNCH" or "bucket" in e.get("name", "")]
    assert len(launches) >= 3 * 2, ev[:3]


def test_lr_schedule_drives_the_flat_optimizer():
    from azure_hc_intel_tf_b200 import hvd
    net = torch.nn.Linear(8, 4)
    opt = hvd.MomentumSGD(list(net.named_parameters()), lr=1.0, momentum=0.0)
    opt.lr_schedule = lambda step: 0.5 if step == 0 else 0.25
    w0 = net.weight.detach().clone()
    net(torch.ones(2, 8)).sum().backward()
    g = net.weight.grad.detach().clone()
    opt.step()
    assert torch.allclose(net.weight.detach(), w0 - 0.5 * g)
    w1 = net.weight.detach().clone()
    opt.step()                                      # same gradient (not zeroed), second rate
    assert torch.allclose(net.weight.detach(), w1 - 0.25 * g) and opt.lr == 0.25
