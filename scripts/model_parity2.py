import copy, sys
import torch, torch.nn.functional as F
sys.path.insert(0, ".")
from azure_hc_intel_tf_b200 import models, ops
model, batch = sys.argv[1], int(sys.argv[2])
torch.manual_seed(0)
ref = models.create_model(model, kernels="cudnn").cuda()
ref32 = models.create_model(model, kernels="cudnn").cuda()
mine = models.create_model(model, kernels="b200").cuda()
mine.load_state_dict(copy.deepcopy(ref.state_dict())); ref32.load_state_dict(copy.deepcopy(ref.state_dict()))
g = torch.Generator().manual_seed(11)
x = torch.randn(batch, ref.image_size, ref.image_size, 3, generator=g).cuda().bfloat16().permute(0, 3, 1, 2)
labels = torch.randint(0, 1000, (batch,), device="cuda")
for net in (ref, mine, ref32):
    net.train()
    for mod in net.modules():
        if mod.__class__.__name__ == "Dropout":
            mod.p = 0.0
with torch.autocast("cuda", dtype=torch.bfloat16):
    l_ref = ops.softmax_cross_entropy(ref(x), labels, ref.cfg)
l_32 = ops.softmax_cross_entropy(ref32(x.float()), labels, ref32.cfg)
l_mine = ops.softmax_cross_entropy(mine(x), labels, mine.cfg)
print("loss bf16-lib", l_ref.item(), "fp32-lib", l_32.item(), "mine", l_mine.item())
l_ref.backward(); l_mine.backward(); l_32.backward()
rows = []
for (n1, p1), (_, p2), (_, p3) in zip(ref.named_parameters(), mine.named_parameters(), ref32.named_parameters()):
    if p1.grad is None: continue
    g1, g2, g3 = p1.grad.float().flatten(), p2.grad.float().flatten(), p3.grad.float().flatten()
    c = lambda a, b: F.cosine_similarity(a, b, dim=0).item()
    rows.append((n1, c(g1, g3), c(g2, g3), c(g1, g2)))
print(f"{'param':40s} lib16~fp32  mine~fp32  lib16~mine")
for i, r in enumerate(rows):
    if i < 6 or i % 16 == 0 or i > len(rows) - 6:
        print(f"{r[0]:40s} {r[1]:.4f}     {r[2]:.4f}     {r[3]:.4f}")
import statistics
print("median", statistics.median(r[1] for r in rows), statistics.median(r[2] for r in rows))
