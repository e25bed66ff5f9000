"""Generate tests/golden/*.pt from the REFERENCE's own classes run under the installed torch.

Run in the build container only (needs /root/reference; the GPU box has no copy):
    python oracle/gen_golden.py
What it does (this is how the oracle is pinned):
  * extracts the ``Net`` class definition from ref: horovod/mnist_horovod.py:9-25 and
    horovod_mnist_elastic.py:16-32 with ``ast`` (the scripts cannot be imported: they need horovod and
    download MNIST at import time) and executes ONLY that class node — no reference source is copied here;
  * imports ``Model`` and ``Trainer`` from ref: pytorch_elastic/mnist_ddp_elastic.py (import-safe);
  * runs the reference forward / loss / backward / optimizer loops on seeded synthetic inputs, replays the
    dropout RNG draws to export the masks, asserts the oracle restatement reproduces the reference
    BIT-FOR-BIT on this machine, and stores inputs + reference outputs as fixtures;
  * runs a real 2-process gloo DDP job of the reference Trainer._run_batch loop over the CNN and stores
    the resulting parameters (pins "mean of per-shard gradients" for the multi-rank path);
  * stores DistributedSampler index digests for a grid of (N, W, rank, epoch, seed, drop_last).
TEST INFRASTRUCTURE ONLY.
"""
from __future__ import annotations

import ast
import hashlib
import importlib.util
import os
import sys
import warnings

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
from oracle import mlp_oracle, net_oracle, sampler_oracle, synth_oracle  # noqa: E402

REF = "/root/reference"
OUT = os.path.join(REPO, "tests", "golden")
warnings.filterwarnings("ignore", message="Implicit dimension choice")


def reference_net_class(script: str):
    """Return the reference's Net class object, built from its own source by executing only the ClassDef."""
    path = os.path.join(REF, script)
    tree = ast.parse(open(path).read(), filename=path)
    node = next(n for n in tree.body if isinstance(n, ast.ClassDef) and n.name == "Net")
    ns = {"torch": torch, "nn": nn, "F": F}
    exec(compile(ast.Module(body=[node], type_ignores=[]), path, "exec"), ns)
    return ns["Net"]


def reference_ddp_module():
    path = os.path.join(REF, "pytorch_elastic", "mnist_ddp_elastic.py")
    spec = importlib.util.spec_from_file_location("ref_mnist_ddp_elastic", path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def synth(seed, start, count):
    x, y = synth_oracle.synth_batch(seed, np.arange(start, start + count))
    return torch.from_numpy(x), torch.from_numpy(y)


def gen_forward_backward(RefNet, tag, batch, seed):
    torch.manual_seed(seed)
    ref = RefNet()
    sd = {k: v.detach().clone() for k, v in ref.state_dict().items()}
    x, y = synth(1234, 1000 * seed, batch)
    # reference training-mode forward with torch's own dropout RNG
    ref.train()
    torch.manual_seed(seed + 100)
    logp_ref = ref(x)
    loss_ref = F.nll_loss(logp_ref, y)
    loss_ref.backward()
    grads_ref = {k: p.grad.detach().clone() for k, p in ref.named_parameters()}
    # replay the RNG draws -> explicit masks -> oracle must reproduce the reference bit-for-bit
    torch.manual_seed(seed + 100)
    m2, m1 = net_oracle.draw_masks(batch)
    logp_o, loss_o, grads_o = net_oracle.forward_backward(sd, x, y, m2, m1, training=True)
    assert torch.equal(logp_o, logp_ref.detach()), "oracle forward != reference forward"
    assert torch.equal(loss_o, loss_ref.detach())
    for k in grads_ref:
        assert torch.equal(grads_o[k], grads_ref[k]), k
    # eval-mode forward (test() path, ref: horovod_mnist_elastic.py:85-102)
    ref.eval()
    with torch.no_grad():
        logp_eval = ref(x)
    o = net_oracle.OracleNet(); o.load_state_dict(sd); o.eval()
    with torch.no_grad():
        assert torch.equal(o(x), logp_eval)
    fx = dict(state_dict=sd, x=x, y=y, mask2d=m2, mask1d=m1, logp=logp_ref.detach(), loss=loss_ref.detach(),
              grads=grads_ref, logp_eval=logp_eval, argmax_eval=logp_eval.argmax(1), synth_seed=1234,
              synth_start=1000 * seed, torch_version=torch.__version__)
    torch.save(fx, os.path.join(OUT, f"fwdbwd_{tag}.pt"))
    print(f"fwdbwd_{tag}: B={batch} loss={loss_ref.item():.6f}")


def gen_train(RefNet, tag, opt_kind, steps, batch, seed, **opt_kw):
    """k steps of the reference loop (ref: horovod/mnist_horovod.py:58-64) with the script's optimizer."""
    torch.manual_seed(seed)
    ref = RefNet()
    sd0 = {k: v.detach().clone() for k, v in ref.state_dict().items()}
    opt = net_oracle.make_optimizer(opt_kind, ref.parameters(), **opt_kw)
    xs, ys = synth(1234, 5000 + 100 * seed, steps * batch)
    xs = xs.view(steps, batch, 1, 28, 28); ys = ys.view(steps, batch)
    ref.train()
    torch.manual_seed(seed + 7)
    losses = []
    for k in range(steps):
        opt.zero_grad()
        out = ref(xs[k])
        loss = F.nll_loss(out, ys[k])
        loss.backward()
        opt.step()
        losses.append(loss.detach())
    torch.manual_seed(seed + 7)
    m2s, m1s = [], []
    for k in range(steps):
        m2, m1 = net_oracle.draw_masks(batch)
        m2s.append(m2); m1s.append(m1)
    m2s, m1s = torch.stack(m2s), torch.stack(m1s)
    losses_o, sd_o = net_oracle.train_steps(sd0, xs, ys, m2s, m1s, opt_kind, world=1, **opt_kw)
    assert torch.equal(losses_o[:, 0], torch.stack(losses)), "oracle loop != reference loop"
    for k, v in ref.state_dict().items():
        assert torch.equal(sd_o[k], v), k
    fx = dict(state_dict0=sd0, xs=xs, ys=ys, masks2d=m2s, masks1d=m1s, losses=torch.stack(losses),
              state_dict=dict((k, v.detach().clone()) for k, v in ref.state_dict().items()),
              opt_kind=opt_kind, opt_kw=opt_kw, torch_version=torch.__version__)
    torch.save(fx, os.path.join(OUT, f"train_{tag}.pt"))
    print(f"train_{tag}: {opt_kind} {steps}x{batch} last loss={losses[-1].item():.6f}")


def _ddp_worker(rank, world, port, sd0, xs, ys, m2s, m1s, out_path):
    """Real gloo DDP of the reference loop body (Trainer._run_batch, ref: mnist_ddp_elastic.py:71-79)."""
    import torch.distributed as dist
    from torch.nn.parallel import DistributedDataParallel as DDP
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), LOCAL_RANK=str(rank),
                      WORLD_SIZE=str(world))
    torch.set_num_threads(1)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    ddpmod = reference_ddp_module()
    net = net_oracle.OracleNet()      # bit-identical to the reference Net (asserted in gen_forward_backward)
    net.load_state_dict(sd0)
    opt = torch.optim.SGD(net.parameters(), lr=0.01)
    trainer = ddpmod.Trainer.__new__(ddpmod.Trainer)   # bypass __init__ (needs DataLoaders); use its _run_batch
    trainer.model = DDP(net, device_ids=None)
    trainer.optimizer = opt
    trainer.model.train()
    steps = xs.shape[0]
    B = xs.shape[1] // world
    for k in range(steps):
        xk = xs[k, rank * B:(rank + 1) * B]; yk = ys[k, rank * B:(rank + 1) * B]
        m2 = m2s[k, rank * B:(rank + 1) * B]; m1 = m1s[k, rank * B:(rank + 1) * B]
        trainer.criterion = lambda out, tgt: F.nll_loss(out, tgt)
        # Trainer._run_batch calls self.model(source): route the injected masks through a closure
        fwd = trainer.model.module.forward
        trainer.model.module.forward = lambda inp, _f=fwd, _m2=m2, _m1=m1: _f(inp, _m2, _m1)
        trainer._run_batch(xk, yk)
        trainer.model.module.forward = fwd
    if rank == 0:
        torch.save({k: v.detach().clone() for k, v in net.state_dict().items()}, out_path)
    dist.destroy_process_group()


def gen_ddp_w2(seed=3, steps=3, batch=16, world=2):
    import torch.multiprocessing as mp
    torch.manual_seed(seed)
    sd0 = {k: v.detach().clone() for k, v in net_oracle.OracleNet().state_dict().items()}
    xs, ys = synth(1234, 9000, steps * batch * world)
    xs = xs.view(steps, world * batch, 1, 28, 28); ys = ys.view(steps, world * batch)
    g = torch.Generator().manual_seed(seed + 11)
    m2s = torch.stack([torch.empty(world * batch, 20).bernoulli_(0.5, generator=g) for _ in range(steps)])
    m1s = torch.stack([torch.empty(world * batch, 50).bernoulli_(0.5, generator=g) for _ in range(steps)])
    tmp = os.path.join(OUT, "_ddp_tmp.pt")
    mp.spawn(_ddp_worker, args=(world, 29613, sd0, xs, ys, m2s, m1s, tmp), nprocs=world, join=True)
    sd_ddp = torch.load(tmp); os.remove(tmp)
    # the single-process restatement of the W-rank step must agree with real DDP (fp32 sum order differs)
    _, sd_o = net_oracle.train_steps(sd0, xs, ys, m2s, m1s, "sgd", world=world, lr=0.01)
    for k in sd_ddp:
        torch.testing.assert_close(sd_o[k], sd_ddp[k], rtol=1e-6, atol=1e-7)
    fx = dict(state_dict0=sd0, xs=xs, ys=ys, masks2d=m2s, masks1d=m1s, state_dict=sd_ddp, world=world,
              opt_kind="sgd", opt_kw=dict(lr=0.01), torch_version=torch.__version__)
    torch.save(fx, os.path.join(OUT, "ddp_gloo_w2.pt"))
    print("ddp_gloo_w2: ok")


def gen_sampler():
    from torch.utils.data.distributed import DistributedSampler
    cases = []
    grid = [(60000, 1, 0, 0, 0, False), (60000, 2, 1, 0, 0, False), (60000, 4, 3, 1, 0, False),
            (60000, 8, 5, 3, 0, False), (60000, 7, 3, 2, 0, False), (60000, 7, 6, 0, 5, True),
            (10000, 2, 0, 4, 0, False), (10000, 3, 2, 0, 42, True), (1000, 8, 7, 9, 12345, False),
            (10, 4, 1, 0, 0, False), (3, 8, 6, 0, 0, False), (1, 1, 0, 0, 0, False), (625, 2, 0, 0, 1, False),
            (65536, 8, 2, 1, 0, False)]
    for n, w, r, ep, seed, dl in grid:
        s = DistributedSampler(range(n), num_replicas=w, rank=r, seed=seed, drop_last=dl)
        s.set_epoch(ep)
        idx = torch.tensor(list(s), dtype=torch.int64)
        mine = torch.from_numpy(sampler_oracle.shard_indices(n, w, r, True, seed, ep, dl))
        assert torch.equal(idx, mine), (n, w, r, ep, seed, dl)
        cases.append(dict(n=n, world=w, rank=r, epoch=ep, seed=seed, drop_last=dl, len=len(idx),
                          head=idx[:16].clone(), tail=idx[-16:].clone(),
                          sha256=hashlib.sha256(idx.numpy().tobytes()).hexdigest()))
    torch.save(dict(cases=cases, torch_version=torch.__version__), os.path.join(OUT, "sampler.pt"))
    print(f"sampler: {len(cases)} cases")


def gen_mlp(tag, hidden_layers, features, steps, batch, seed):
    """The DDP script's own Model + CrossEntropyLoss + Adam through its own Trainer._run_batch
    (ref: pytorch_elastic/mnist_ddp_elastic.py:70-75,133-159,172-174), imported from the reference file."""
    ref_mod = reference_ddp_module()
    torch.manual_seed(seed)
    ref = ref_mod.Model(hidden_layers=hidden_layers, features=features)
    sd0 = {k: v.detach().clone() for k, v in ref.state_dict().items()}
    xs, ys = synth(1234, 9000 + 100 * seed, steps * batch)
    xs = xs.view(steps, batch, 1, 28, 28); ys = ys.view(steps, batch)
    # single forward/backward
    crit = nn.CrossEntropyLoss()
    logits_ref = ref(xs[0]); loss_ref = crit(logits_ref, ys[0]); loss_ref.backward()
    grads_ref = {k: p.grad.detach().clone() for k, p in ref.named_parameters()}
    logits_o, loss_o, grads_o = mlp_oracle.forward_backward(sd0, xs[0], ys[0], hidden_layers, features)
    assert torch.equal(logits_o, logits_ref.detach()) and torch.equal(loss_o, loss_ref.detach())
    for k in grads_ref:
        assert torch.equal(grads_o[k], grads_ref[k]), k
    # the reference Trainer's loop body, driven without its DDP / dataloader plumbing
    opt = torch.optim.Adam(ref.parameters(), lr=1e-3)           # ref :173
    trainer = ref_mod.Trainer.__new__(ref_mod.Trainer)
    trainer.model, trainer.optimizer, trainer.criterion = ref, opt, crit
    losses = []
    for k in range(steps):
        trainer._run_batch(xs[k], ys[k])                         # ref :70-75 (returns nothing)
        with torch.no_grad():
            losses.append(crit(ref(xs[k]), ys[k]).detach())      # post-step loss, for the record only
    losses_o, sd_o = mlp_oracle.train_steps(sd0, xs, ys, hidden_layers, features, "adam", lr=1e-3)
    for k, v in ref.state_dict().items():
        assert torch.equal(sd_o[k], v), k
    fx = dict(state_dict0=sd0, xs=xs, ys=ys, logits=logits_ref.detach(), loss=loss_ref.detach(), grads=grads_ref,
              losses=losses_o[:, 0], state_dict={k: v.detach().clone() for k, v in ref.state_dict().items()},
              hidden_layers=hidden_layers, features=features, torch_version=torch.__version__)
    torch.save(fx, os.path.join(OUT, f"mlp_{tag}.pt"))
    print(f"mlp_{tag}: H={hidden_layers} F={features} {steps}x{batch} loss0={loss_ref.item():.6f}")


def main():
    os.makedirs(OUT, exist_ok=True)
    if "--mlp-only" in sys.argv:
        gen_mlp("h1_f128", 1, 128, steps=3, batch=32, seed=11)       # the class defaults (ref :137)
        gen_mlp("h2_f64", 2, 64, steps=3, batch=24, seed=12)
        return
    NetA = reference_net_class("horovod/mnist_horovod.py")
    NetB = reference_net_class("horovod/horovod_mnist_elastic.py")
    gen_forward_backward(NetA, "b8", 8, seed=1)
    gen_forward_backward(NetA, "b1", 1, seed=2)
    gen_forward_backward(NetB, "b48", 48, seed=3)      # ragged last batch of the W=2 epoch
    gen_forward_backward(NetA, "b64", 64, seed=4)
    gen_train(NetA, "sgd", "sgd", steps=4, batch=32, seed=5, lr=0.01)                        # mnist_horovod.py:50
    gen_train(NetB, "adamw", "adamw", steps=4, batch=32, seed=6, lr=0.01 / 2 ** 0.5)         # elastic :41 at size 2
    gen_train(NetA, "adam", "adam", steps=4, batch=32, seed=7, lr=1e-3)                      # ddp script :173
    gen_train(NetA, "sgdm", "sgd", steps=4, batch=32, seed=8, lr=0.01, momentum=0.9)         # BASELINE north_star
    gen_train(NetA, "adadelta", "adadelta", steps=4, batch=32, seed=9, lr=1.0)               # BASELINE north_star
    gen_ddp_w2()
    gen_sampler()
    gen_mlp("h1_f128", 1, 128, steps=3, batch=32, seed=11)       # the class defaults (ref :137)
    gen_mlp("h2_f64", 2, 64, steps=3, batch=24, seed=12)


if __name__ == "__main__":
    main()
