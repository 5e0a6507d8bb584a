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
