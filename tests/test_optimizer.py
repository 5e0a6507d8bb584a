"""Flat-buffer momentum SGD / LARS against plain PyTorch loops (oracle for K10 / K-O1)."""
import copy

import torch
import torch.nn as nn

from azure_hc_intel_tf_b200 import hvd


def small_net():
    torch.manual_seed(0)
    net = nn.Sequential(nn.Conv2d(3, 8, 3, padding=1), nn.ReLU(), nn.Conv2d(8, 8, 3, padding=1), nn.Flatten(),
                        nn.Linear(8 * 8 * 8, 10))
    return net.to(memory_format=torch.channels_last)


def test_matches_torch_sgd_momentum_with_weight_decay():
    a, b = small_net(), small_net()
    b.load_state_dict(copy.deepcopy(a.state_dict()))
    ref = torch.optim.SGD(a.parameters(), lr=0.05, momentum=0.9, weight_decay=1e-3)
    opt = hvd.MomentumSGD(list(b.named_parameters()), lr=0.05, momentum=0.9, weight_decay=1e-3,
                          fusion_threshold=16384)
    assert len(opt.plan.buckets) > 1
    x = torch.randn(4, 3, 8, 8)
    for _ in range(5):
        ref.zero_grad()
        a(x).square().mean().backward()
        ref.step()
        opt.zero_grad()
        b(x).square().mean().backward()
        opt.step()
    for pa, pb in zip(a.parameters(), b.parameters()):
        assert torch.allclose(pa, pb, atol=1e-6, rtol=1e-5)


def test_params_and_grads_are_views_of_flat_buffers():
    net = small_net()
    opt = hvd.MomentumSGD(list(net.named_parameters()), lr=0.1)
    p = net[0].weight
    s, n = opt.ranges[p.flat_index]
    assert p.data_ptr() == opt.master[s:].data_ptr() and p.grad.data_ptr() == opt.grad[s:].data_ptr()
    assert p.is_contiguous(memory_format=torch.channels_last)
    net(torch.randn(2, 3, 8, 8)).sum().backward()
    assert p.grad.data_ptr() == opt.grad[s:].data_ptr()          # autograd accumulated in place
    assert opt.grad[s:s + n].abs().sum() > 0


def test_lars_matches_manual_loop():
    torch.manual_seed(0)
    net = small_net()
    ref = copy.deepcopy(net)
    lr, mu, wd, eta = 0.1, 0.9, 1e-3, 0.01
    opt = hvd.LARS(list(net.named_parameters()), lr=lr, momentum=mu, weight_decay=wd, eta=eta)
    vel = [torch.zeros_like(p) for p in ref.parameters()]
    x = torch.randn(4, 3, 8, 8)
    for _ in range(3):
        opt.zero_grad()
        net(x).square().mean().backward()
        opt.step()
        ref.zero_grad()
        ref(x).square().mean().backward()
        with torch.no_grad():
            for p, v in zip(ref.parameters(), vel):
                g = p.grad + wd * p
                if p.dim() > 1:
                    trust = eta * p.norm() / (p.grad.norm() + wd * p.norm() + 1e-9)
                    g = g * trust
                v.mul_(mu).add_(g)
                p.add_(v, alpha=-lr)
    for pa, pb in zip(ref.parameters(), net.parameters()):
        assert torch.allclose(pa, pb, atol=1e-6, rtol=1e-4)


def test_lowp_shadow_tracks_master():
    net = small_net()
    opt = hvd.MomentumSGD(list(net.named_parameters()), lr=0.1, lowp=True)
    net(torch.randn(2, 3, 8, 8)).sum().backward()
    opt.step()
    p = net[4].weight
    assert p.lp.dtype == torch.bfloat16 and torch.equal(p.lp, p.detach().to(torch.bfloat16))


def test_state_dict_round_trip():
    net = small_net()
    opt = hvd.MomentumSGD(list(net.named_parameters()), lr=0.1)
    net(torch.randn(2, 3, 8, 8)).sum().backward()
    opt.step()
    sd = opt.state_dict()
    net2 = small_net()
    opt2 = hvd.MomentumSGD(list(net2.named_parameters()), lr=0.5)
    opt2.load_state_dict(sd)
    assert torch.equal(opt.master, opt2.master) and torch.equal(opt.mom, opt2.mom) and opt2.lr == 0.1
    assert torch.equal(net[0].weight, net2[0].weight)


def _tf_reference_steps(net, rule, lr, mu, wd, x, steps):
    """Per-tensor loops with TensorFlow's ApplyRMSProp / ApplyAdam update rules (tf_cnn_benchmarks defaults)."""
    ps = list(net.parameters())
    s1 = [torch.zeros_like(p) for p in ps]
    s2 = [torch.ones_like(p) if rule == "rmsprop" else torch.zeros_like(p) for p in ps]
    for t in range(1, steps + 1):
        net.zero_grad()
        net(x).square().mean().backward()
        with torch.no_grad():
            for p, a, b in zip(ps, s1, s2):
                g = p.grad + wd * p
                if rule == "rmsprop":
                    b.mul_(0.9).add_(0.1 * g * g)
                    a.mul_(mu).add_(lr * g / torch.sqrt(b + 1.0))
                    p.sub_(a)
                else:
                    a.mul_(0.9).add_(0.1 * g)
                    b.mul_(0.999).add_(0.001 * g * g)
                    lr_t = lr * (1 - 0.999 ** t) ** 0.5 / (1 - 0.9 ** t)
                    p.sub_(lr_t * a / (torch.sqrt(b) + 1e-8))
    return s1, s2


def test_flat_rmsprop_and_adam_match_tensorflow_update_rules():
    x = torch.randn(4, 3, 8, 8)
    for rule, mk, kw in (("rmsprop", hvd.RMSProp, dict(lr=0.01, momentum=0.9, weight_decay=1e-3)),
                         ("adam", hvd.Adam, dict(lr=0.01, weight_decay=1e-3))):
        net, ref = small_net(), small_net()
        opt = mk(list(net.named_parameters()), fusion_threshold=16384, lowp=True, **kw)
        assert len(opt.plan.buckets) > 1 and opt.rule == rule
        for _ in range(5):
            opt.zero_grad()
            net(x).square().mean().backward()
            opt.step()
        s1, s2 = _tf_reference_steps(ref, rule, 0.01, kw.get("momentum", 0.0), 1e-3, x, 5)
        for pa, pb, a, b in zip(net.parameters(), ref.parameters(), s1, s2):
            assert torch.allclose(pa, pb, atol=1e-6, rtol=1e-4), rule
            slots = opt.slots_of(pa)
            first, second = ("RMSProp_1", "RMSProp") if rule == "rmsprop" else ("Adam", "Adam_1")
            assert set(slots) == {first, second}
            assert torch.allclose(slots[first], a, atol=1e-6, rtol=1e-4) and torch.allclose(slots[second], b, atol=1e-6, rtol=1e-4)
        # the alignment padding behind every tensor stays zero in the weights and the first slot
        live = torch.zeros(opt.total, dtype=torch.bool)
        for s0, n in opt.ranges:
            live[s0:s0 + n] = True
        assert (~live).any() and torch.all(opt.master[~live] == 0) and torch.all(opt.mom[~live] == 0)
        assert torch.equal(opt.shadow, opt.master.to(torch.bfloat16))
        # state round trip keeps both slots and refuses a different rule
        opt2 = mk(list(small_net().named_parameters()), fusion_threshold=16384, **kw)
        opt2.load_state_dict(opt.state_dict())
        assert torch.equal(opt2.state2, opt.state2) and opt2.steps == 5
        other = hvd.MomentumSGD(list(small_net().named_parameters()), lr=0.1, fusion_threshold=16384)
        try:
            other.load_state_dict(opt.state_dict())
            raise AssertionError("a momentum optimizer accepted rmsprop / adam state")
        except ValueError:
            pass


def test_adam_bias_correction_reaches_the_device_scalar_path():
    """effective_lr() is what the kernels read from lr_dev: base rate times sqrt(1-b2^t)/(1-b1^t) for step t."""
    net = small_net()
    opt = hvd.Adam(list(net.named_parameters()), lr=0.001)
    assert abs(opt.effective_lr() - 0.001 * (1 - 0.999) ** 0.5 / (1 - 0.9)) < 1e-12
    opt.steps = 1000
    assert abs(opt.effective_lr() - 0.001 * (1 - 0.999 ** 1001) ** 0.5 / (1 - 0.9 ** 1001)) < 1e-12
    opt.lr_schedule = lambda step: 0.5
    opt.apply_schedule()
    assert opt.lr == 0.5


def test_cli_optimizers_all_run_on_the_flat_buffers():
    from azure_hc_intel_tf_b200.cli.cnn_benchmarks import BenchmarkCNN
    from azure_hc_intel_tf_b200.utils import flags, tfnames
    for name, slots in (("rmsprop", ("RMSProp", "RMSProp_1")), ("adam", ("Adam", "Adam_1")), ("momentum", ("Momentum",))):
        b = BenchmarkCNN(flags.parse(["--model=lenet", "--batch_size=4", "--variable_update=horovod", "--device=cpu",
                                      f"--optimizer={name}"]))
        opt = getattr(b.opt, "_opt", b.opt)
        assert isinstance(opt, hvd.MomentumSGD) and opt.rule == (name if name != "momentum" else "momentum")
        l0 = float(b.step().detach())
        for _ in range(3):
            l1 = float(b.step().detach())
        assert l1 == l1 and l1 < l0 * 1.5
        v = tfnames.variables(b.model, b.opt, 4)
        w = "v0/cg/conv0/conv2d/kernel"
        assert all(f"{w}/{s}" in v for s in slots), [k for k in v if k.startswith(w)]
