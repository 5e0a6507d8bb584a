"""End-to-end loss-curve parity ("custom kernels vs cuDNN path, same seed, ~100 steps"): ResNet-50 trained
for 100 steps through the sm_100a kernel path and through the stock PyTorch/cuDNN path from the same initial weights
on the same synthetic batches. A wrong filter tap, a dropped k-block on one layer, or a BatchNorm mask bug that only
shows after a few updates bends the loss curve or the weights away from the library's; one-step checks do not see it.
Reference role: the reference trains the same model through TF + MKL-DNN (run-tf-sing-ucx-openmpi.sh:34,66,76)."""
import pytest
import torch

pytestmark = pytest.mark.gpu

STEPS = 100


def _train(kernels, steps=STEPS, batch=32, model="resnet50", nbatches=16, extra=()):
    from azure_hc_intel_tf_b200.cli.cnn_benchmarks import BenchmarkCNN
    from azure_hc_intel_tf_b200.ops import data as data_mod
    from azure_hc_intel_tf_b200.utils import flags
    p = flags.parse([f"--model={model}", f"--batch_size={batch}", "--variable_update=independent", "--device=gpu",
                     f"--kernels={kernels}", "--use_cuda_graph=False", "--seed=4321", "--optimizer=momentum",
                     "--init_learning_rate=0.004"] + list(extra))
    b = BenchmarkCNN(p)
    # sixteen different batches with random labels, cycled: the loss stays O(1) for the whole run (one reused batch is
    # memorised within 100 steps and a relative comparison of losses near zero says nothing)
    batches = [data_mod.DeviceSynthetic(batch, b.model.image_size, b.model.cfg.num_classes, b.device, b.dtype,
                                        seed=100 + i).next() for i in range(nbatches)]
    losses = []
    for i in range(steps):
        losses.append(b.train_step(*batches[i % nbatches]).detach())
    torch.cuda.synchronize()
    losses = [float(x) for x in losses]
    weights = {n: q.detach().float().cpu().clone() for n, q in b.model.named_parameters() if q.requires_grad}
    return losses, weights


def _avg10(l):
    return torch.tensor([sum(l[i:i + 10]) / 10 for i in range(0, len(l), 10)])


def _flat(w):
    return torch.cat([w[n].reshape(-1) for n in sorted(w)])


def test_resnet50_loss_curve_matches_library_path_over_100_steps():
    """Three runs from the same weights on the same sixteen batches: the sm_100a kernel path (bf16), the library path
    in bf16 (autocast) and the library path in fp32. Training is chaotic enough that two correct implementations
    drift apart over 100 steps, so the yardstick is measured in the same test: the kernel path may sit at most twice
    as far from the bf16 library run as that run sits from its own fp32 twin (plus a small floor), per 10-step
    window of the loss curve, and its final weights must be as close to the fp32 run's as the bf16 library's are."""
    la, wa = _train("b200")
    lb, wb = _train("cudnn")
    lc, wc = _train("cudnn", extra=["--use_fp16=False"])
    for l in (la, lb, lc):
        assert all(x == x and x < 50.0 for x in l), l
    assert abs(la[0] - lb[0]) < 0.01 * lb[0], (la[0], lb[0])          # same weights, same batch: same first loss
    a, b, c = _avg10(la), _avg10(lb), _avg10(lc)
    yard = (b - c).abs().max().item()
    mine = (a - b).abs().max().item()
    assert mine <= 2.0 * yard + 0.15, (mine, yard, a.tolist(), b.tolist(), c.tolist())
    # all three actually trained: the last ten steps are below the first ten
    assert a[-1] < a[0] and b[-1] < b[0] and c[-1] < c[0], (a.tolist(), b.tolist(), c.tolist())
    cos = torch.nn.functional.cosine_similarity
    fa, fb, fc = _flat(wa), _flat(wb), _flat(wc)
    lib = cos(fb, fc, dim=0).item()
    own = cos(fa, fc, dim=0).item()
    assert own > lib - 0.01 and own > 0.97, (own, lib)
    # no single large tensor went its own way (a wrong tap or a dropped k-block shows up here first)
    for n in wa:
        if wa[n].numel() >= 16384:
            cn = cos(wa[n].reshape(-1), wc[n].reshape(-1), dim=0).item()
            ln = cos(wb[n].reshape(-1), wc[n].reshape(-1), dim=0).item()
            assert cn > ln - 0.05, (n, cn, ln)
