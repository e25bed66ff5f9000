"""Two fused MLP steps at --batch (for ncu launch lists / captures)."""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import pytorch_distributed_examples_b200 as pkg
B = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 2
ds = pkg.SyntheticMNIST(length=1 << 20, seed=1)
x, y = ds.batch(base_index=0, count=B)
torch.manual_seed(0)
model = pkg.Model(5, 1024).cuda()
for _ in range(reps):
    loss = model.loss_and_grads(x, y)
torch.cuda.synchronize()
print("loss", loss.item())
