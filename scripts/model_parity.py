import copy, sys
import torch, torch.nn.functional as F
sys.path.insert(0, ".")
from azure_hc_intel_tf_b200 import models, ops
model, batch = sys.argv[1], int(sys.argv[2])
torch.manual_seed(0)
ref = models.create_model(model, kernels="cudnn").cuda()
mine = models.create_model(model, kernels="b200").cuda()
mine.load_state_dict(copy.deepcopy(ref.state_dict()))
g = torch.Generator().manual_seed(11)
x = torch.randn(batch, ref.image_size, ref.image_size, 3, generator=g).cuda().bfloat16().permute(0, 3, 1, 2)
labels = torch.randint(0, 1000, (batch,), device="cuda")
for net in (ref, mine):
    net.train()
    for mod in net.modules():
        if mod.__class__.__name__ == "Dropout":
            mod.p = 0.0
with torch.autocast("cuda", dtype=torch.bfloat16):
    l_ref = ops.softmax_cross_entropy(ref(x), labels, ref.cfg)
l_mine = ops.softmax_cross_entropy(mine(x), labels, mine.cfg)
print("loss", l_ref.item(), l_mine.item())
l_ref.backward(); l_mine.backward()
for (n1, p1), (n2, p2) in zip(ref.named_parameters(), mine.named_parameters()):
    if p1.grad is None: continue
    g1, g2 = p1.grad.float().flatten(), p2.grad.float().flatten()
    print(f"{n1:40s} cos={F.cosine_similarity(g1, g2, dim=0).item():.4f} |ref|={g1.norm().item():.3e} |mine|={g2.norm().item():.3e}")
