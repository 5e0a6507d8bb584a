"""GPU side of hvd.MomentumSGD: flat momentum-SGD / LARS kernels from libb200cnn.so (csrc/kernels/optim.cu).
Reference role: TF ApplyMomentum under --optimizer=momentum (run-tf-sing-ucx-openmpi.sh:73)."""
from __future__ import annotations

from . import b200
from .b200 import _check, _ptr, _stream, lib


def sgd_update(opt, lo: int, hi: int, lr: float, mu: float, wd: float, grad_scale: float):
    if hi <= lo:
        return
    blk = opt.block_tensor[lo // 1024:]
    shadow = opt.shadow[lo:] if opt.shadow is not None else None
    _check(lib().b200_sgd_momentum_flat(_ptr(opt.master[lo:]), _ptr(opt.mom[lo:]), _ptr(opt.grad[lo:]), _ptr(shadow),
                                        hi - lo, _ptr(blk), _ptr(opt.tensor_decay), _ptr(opt.tensor_lars),
                                        _ptr(opt.trust) if opt.lars else _ptr(None), lr, mu, wd, grad_scale,
                                        _ptr(getattr(opt, "lr_dev", None)), _stream()), "sgd_momentum_flat")


def adaptive_update(opt, lo: int, hi: int, lr: float, wd: float, grad_scale: float):
    """RMSProp / Adam (opt.rule) on flat elements [lo, hi): one launch (csrc/kernels/optim.cu adaptive_flat_kernel).
    Reference role: TF ApplyRMSProp / ApplyAdam under --optimizer=rmsprop|adam (the flag at run-tf-sing-ucx-openmpi.sh:73)."""
    if hi <= lo:
        return
    blk = opt.block_tensor[lo // 1024:]
    shadow = opt.shadow[lo:] if opt.shadow is not None else None
    h = opt.hyper
    common = (_ptr(opt.master[lo:]), _ptr(opt.mom[lo:]), _ptr(opt.state2[lo:]), _ptr(opt.grad[lo:]), _ptr(shadow), hi - lo,
              _ptr(blk), _ptr(opt.tensor_decay))
    if opt.rule == "rmsprop":
        _check(lib().b200_rmsprop_flat(*common, lr, h["rho"], opt.param_groups[0]["momentum"], h["eps"], wd, grad_scale,
                                       _ptr(opt.lr_dev), _stream()), "rmsprop_flat")
    else:
        _check(lib().b200_adam_flat(*common, opt.effective_lr(), h["beta1"], h["beta2"], h["eps"], wd, grad_scale,
                                    _ptr(opt.lr_dev), _stream()), "adam_flat")


def lars_trust(opt, wd: float, grad_scale: float):
    import torch
    if not hasattr(opt, "_lars_blk"):
        lo = torch.tensor([s // 1024 for s, _ in opt.ranges], dtype=torch.int32, device=opt.device)
        cnt = torch.tensor([(n + 1023) // 1024 for _, n in opt.ranges], dtype=torch.int32, device=opt.device)
        part = torch.empty(2 * (opt.total // 1024), dtype=torch.float32, device=opt.device)
        opt._lars_blk = (lo, cnt, part)
    lo, cnt, part = opt._lars_blk
    _check(lib().b200_lars_trust_det(_ptr(opt.master), _ptr(opt.grad), opt.total, _ptr(part), _ptr(lo), _ptr(cnt),
                                     _ptr(opt.tensor_decay), _ptr(opt.tensor_lars), _ptr(opt.trust), len(opt.flags),
                                     opt.lars_eta, wd, opt.lars_eps, grad_scale, _stream()), "lars_trust_det")
    b200.count_launch(1)
