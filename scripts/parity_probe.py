import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch
from test_gpu_parity import _train
la, wa = _train("b200"); lb, wb = _train("cudnn")
lc, wc = _train("cudnn", extra=["--use_fp16=False"])
def avg10(l): return [round(sum(l[i:i+10])/10, 3) for i in range(0, len(l), 10)]
print("b200     ", avg10(la)); print("cudnn bf16", avg10(lb)); print("cudnn fp32", avg10(lc))
fc = torch.cat([wc[n].reshape(-1) for n in sorted(wc)])
a, b = torch.tensor(la), torch.tensor(lb)
rel = (a - b).abs() / b.abs()
print("first", la[:3], lb[:3]); print("last", la[-3:], lb[-3:])
print("rel max", rel.max().item(), "mean", rel.mean().item())
fa = torch.cat([wa[n].reshape(-1) for n in sorted(wa)]); fb = torch.cat([wb[n].reshape(-1) for n in sorted(wb)])
print("cos all b200-vs-cudnn16", torch.nn.functional.cosine_similarity(fa, fb, dim=0).item())
print("cos all cudnn16-vs-cudnn32", torch.nn.functional.cosine_similarity(fb, fc, dim=0).item(), "b200-vs-cudnn32", torch.nn.functional.cosine_similarity(fa, fc, dim=0).item())
cs = sorted((torch.nn.functional.cosine_similarity(wa[n].reshape(-1), wb[n].reshape(-1), dim=0).item(), n, wa[n].numel()) for n in wa)
print(cs[:8])
