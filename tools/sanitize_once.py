"""Small end-to-end pass at odd sizes: CNN step at ragged batches, MLP at shapes off the tile sizes (a target for a memory checker where one is available; on its own it just has to finish)."""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import pytorch_distributed_examples_b200 as pkg
import torch.nn.functional as F
ds = pkg.SyntheticMNIST(length=4096, seed=1)
for B in (3, 33, 130):
    net = pkg.Net().cuda().train()
    opt = pkg.FusedOptimizer(torch.optim.SGD(net.parameters(), lr=0.01, momentum=0.9))
    x, y = ds.batch(base_index=0, count=B)
    for _ in range(2):
        opt.zero_grad(); F.nll_loss(net(x), y).backward(); opt.step()
    pkg.evaluate(net, [(x, y)])
for H, Fe, B in ((1, 100, 36), (2, 36, 64), (1, 128, 30)):
    m = pkg.Model(hidden_layers=H, features=Fe).cuda()
    x, y = ds.batch(base_index=0, count=B)
    loss = F.cross_entropy(m(x), y); loss.backward()
    if B % 4 == 0:
        m.loss_and_grads(x, y)
torch.cuda.synchronize()
print("SANITIZE_PASS_DONE", float(loss))
