"""Runs the BN backward column reduction three times on one shape (ncu capture target).
usage: one_bn_reduce.py rows c"""
import sys
import torch
sys.path.insert(0, ".")
from azure_hc_intel_tf_b200.ops.b200 import _check, _ptr, _stream, lib, scratch
rows, c = int(sys.argv[1]), int(sys.argv[2])
dev = torch.device("cuda")
dz = torch.randn(rows, c, device=dev).to(torch.bfloat16)
y = torch.randn(rows, c, device=dev).to(torch.bfloat16)
mask = torch.randint(0, 256, (rows * c // 8,), dtype=torch.uint8, device=dev)
st = torch.rand(4, c, device=dev) + 0.5
dg, db = torch.empty(c, device=dev), torch.empty(c, device=dev)
for _ in range(3):
    _check(lib().b200_bn_bwd_reduce(_ptr(dz), _ptr(y), _ptr(None), _ptr(mask), rows, c, _ptr(st[0]), _ptr(st[1]), _ptr(st[2]),
                                    _ptr(st[3]), 1, _ptr(dg), _ptr(db), _ptr(scratch(dev)), _stream()), "bn_bwd_reduce")
torch.cuda.synchronize()
print("ok", float(dg.sum()), float(db.sum()))
