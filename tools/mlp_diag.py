"""Developer diagnostic: per-parameter gradient error of the native MLP vs fp64, with torch-GPU fp32 beside it."""
import os, sys
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import pytorch_distributed_examples_b200 as pkg
from oracle import mlp_oracle, synth_oracle

B = int(sys.argv[1]) if len(sys.argv) > 1 else 256
torch.manual_seed(3)
ref = mlp_oracle.OracleMLP(5, 1024)
sd = {k: v.detach().clone() for k, v in ref.state_dict().items()}
xn, yn = synth_oracle.synth_batch(1234, np.arange(B))
x, y = torch.from_numpy(xn), torch.from_numpy(yn)
torch.backends.cuda.matmul.allow_tf32 = False
def run(dtype, dev):
    net = mlp_oracle.OracleMLP(5, 1024).to(dtype).to(dev)
    net.load_state_dict({k: v.to(dtype) for k, v in sd.items()})
    acts = []
    h = x.to(dev).to(dtype).reshape(B, -1)
    h = torch.relu(net.input_layer(h)); acts.append(h)
    for l in net.hidden_layers:
        h = torch.relu(l(h)); acts.append(h)
    logits = net.final_layer(h)
    loss = torch.nn.functional.cross_entropy(logits, y.to(dev))
    loss.backward()
    return logits.detach(), loss.detach(), {k: p.grad.detach().clone() for k, p in net.named_parameters()}, [a.detach() for a in acts]
lg64, loss64, g64, a64 = run(torch.float64, "cuda")
lg32, loss32, g32, a32 = run(torch.float32, "cuda")
model = pkg.Model(5, 1024); model.load_state_dict(sd); model = model.cuda()
loss = model.loss_and_grads(x.cuda(), y.cuda())
with torch.no_grad():
    logits = model(x.cuda())
print("loss", loss.item(), loss64.item(), "logits err ours", (logits.double() - lg64).abs().max().item(), "torch32", (lg32.double() - lg64).abs().max().item(), "scale", lg64.abs().max().item())
for i, (a, b) in enumerate(zip(a32, a64)):
    print(f"layer {i}: torch fp32 mask flips vs fp64: {int(((a > 0) != (b > 0)).sum())} of {a.numel()}, act err {float((a.double()-b).abs().max()):.2e} scale {float(b.abs().max()):.2e}")
for k, p in model.named_parameters():
    s = g64[k].abs().max().item()
    print(f"{k:28s} ours {float((p.grad.double() - g64[k]).abs().max()) / s:.2e}   torch-gpu-fp32 {float((g32[k].double() - g64[k]).abs().max()) / s:.2e}   (rel. to max |g| = {s:.2e})")
