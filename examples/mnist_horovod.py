"""`horovod/mnist_horovod.py` of the reference, switched to the B200 path.

What the reference script does (ref: horovod/mnist_horovod.py:27-67): hvd.init, shard MNIST with a
DistributedSampler, wrap SGD(lr=0.01) in hvd.DistributedOptimizer, broadcast the parameters from rank 0, then 50
epochs of the five-line loop, printing the loss every 5 batches.  Here the imports change and the data is born
on the device (there is no network for the MNIST download); the loop body is the reference's.

    torchrun --nproc-per-node=8 examples/mnist_horovod.py            # reference settings: 50 epochs, batch 1024
    python examples/mnist_horovod.py --epochs 1 --train_samples 8192 # a short single-GPU run
"""
import argparse

import _common  # noqa: F401
import torch
import torch.nn.functional as F
import torch.optim as optim

from pytorch_distributed_examples_b200 import DeviceLoader, DistributedSampler, Net, SyntheticMNIST, hvd


def main():
    ap = argparse.ArgumentParser(description=__doc__.splitlines()[0])
    ap.add_argument("--epochs", type=int, default=50)             # ref :58
    ap.add_argument("--batch_size", type=int, default=1024)       # ref :44
    ap.add_argument("--train_samples", type=int, default=60000)   # MNIST's training-set size
    ap.add_argument("--log_every", type=int, default=5)           # ref :65
    ap.add_argument("--seed", type=int, default=None, help="fix the initial weights and the dropout stream (the reference does not seed)")
    args = ap.parse_args()
    if args.seed is not None:
        torch.manual_seed(args.seed)

    hvd.init()                                                                                    # ref :28
    train_dataset = SyntheticMNIST(args.train_samples)                                            # ref :38
    train_sampler = DistributedSampler(train_dataset, num_replicas=hvd.size(), rank=hvd.rank())   # ref :41-42
    train_loader = DeviceLoader(train_dataset, batch_size=args.batch_size, sampler=train_sampler) # ref :44

    model = Net().cuda()                                                                          # ref :47-48
    optimizer = optim.SGD(model.parameters(), lr=0.01)                                            # ref :50
    optimizer = hvd.DistributedOptimizer(optimizer, named_parameters=model.named_parameters())    # ref :53
    hvd.broadcast_parameters(model.state_dict(), root_rank=0)                                     # ref :56

    last = None
    for epoch in range(args.epochs):
        for batch_idx, (data, target) in enumerate(train_loader):
            optimizer.zero_grad()
            output = model(data)
            loss = F.nll_loss(output, target)
            loss.backward()
            optimizer.step()
            if batch_idx % args.log_every == 0:
                last = loss.item()
                print(f"Train Epoch: {epoch} [{batch_idx * len(data)}/{len(train_sampler)}]\tLoss: {last}", flush=True)
    return last


if __name__ == "__main__":
    main()
