"""Where a training step (config 2, guide frozen / trainable) spends its time: torch profiler table."""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch.profiler import profile, ProfilerActivity
cm = importlib.import_module("depth_estimation_b200.crf.crf_module")
gm = importlib.import_module("depth_estimation_b200.crf.gaussian_matrix")
synthetic = importlib.import_module("depth_estimation_b200.synthetic")
dev = torch.device("cuda", 0)
h, w, L, niters = 500, 740, 64, 5
img = torch.from_numpy(synthetic.image(h, w, 0)).permute(2, 0, 1)[None].contiguous().to(dev)
logits0 = (-torch.from_numpy(synthetic.unary(h, w, L, 0))).reshape(h, w, L).permute(2, 0, 1)[None].contiguous().to(dev)
target = torch.from_numpy(synthetic.disparity(h, w, L, 0)).float()[None, None].to(dev)
labels = torch.arange(L, dtype=torch.float32, device=dev)
train_guide = len(sys.argv) > 1 and sys.argv[1] == "guide"
guide_mod = cm.ijrgbGuide(0.1, 0.1, trainable=train_guide).to(dev)
crf = cm.CRFasRNN(cm.charb(0.05), niters=niters).to(dev)

def step():
    lg = logits0.clone().requires_grad_()
    out = crf(guide_mod(img), lg)
    depth = (torch.softmax(out, dim=1) * labels[None, :, None, None]).sum(1)[:, None]
    loss = (depth - target).abs().mean()
    loss.backward()
    gm.clear_lattice_cache()

for _ in range(3):
    step()
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
    step()
    torch.cuda.synchronize()
print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=28, max_name_column_width=60), file=sys.stderr)
