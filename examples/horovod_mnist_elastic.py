"""`horovod/horovod_mnist_elastic.py` of the reference, switched to the B200 path.

The reference (ref: horovod/horovod_mnist_elastic.py:35-106) trains the CNN with AdamW(lr = 0.01 / sqrt(size)) under
`@hvd.elastic.run`, commits its `TorchState` every 30 batches, rescales the learning rate in a reset callback and
prints the test accuracy at the end.  The elastic API is used verbatim; Horovod's host-discovery driver is replaced
by `hvd.elastic.request_resize(ranks)` (any member may call it) and by the automatic recovery from a lost peer.

    torchrun --nproc-per-node=8 examples/horovod_mnist_elastic.py
    python examples/horovod_mnist_elastic.py --epochs 1 --train_samples 8192 --test_samples 2048
"""
import argparse
import math

import _common  # noqa: F401
import torch
import torch.nn.functional as F
import torch.optim as optim

from pytorch_distributed_examples_b200 import DeviceLoader, DistributedSampler, Net, SyntheticMNIST, hvd


def main():
    ap = argparse.ArgumentParser(description=__doc__.splitlines()[0])
    ap.add_argument("--epochs", type=int, default=15)                 # ref :10
    ap.add_argument("--batches_per_commit", type=int, default=30)     # ref :11
    ap.add_argument("--lr", type=float, default=0.01)                 # ref :12
    ap.add_argument("--batch_size", type=int, default=128)            # ref :52
    ap.add_argument("--train_samples", type=int, default=60000)
    ap.add_argument("--test_samples", type=int, default=10000)
    ap.add_argument("--seed", type=int, default=None, help="fix the initial weights and the dropout stream (the reference does not seed)")
    args = ap.parse_args()
    if args.seed is not None:
        torch.manual_seed(args.seed)

    hvd.init()                                                                                      # ref :35
    model = Net().cuda()                                                                            # ref :38
    optimizer = optim.AdamW(model.parameters(), lr=args.lr / math.sqrt(hvd.size()))                 # ref :41
    optimizer = hvd.DistributedOptimizer(optimizer, named_parameters=model.named_parameters())      # ref :42
    train_dataset = SyntheticMNIST(args.train_samples, seed=1234)

    def get_dataset():                                                                              # ref :44-53
        sampler = DistributedSampler(train_dataset, num_replicas=hvd.size(), rank=hvd.rank())
        return DeviceLoader(train_dataset, batch_size=args.batch_size, sampler=sampler)

    @hvd.elastic.run                                                                                # ref :55
    def train(state):
        train_loader = get_dataset()          # re-created after every reset: the shard depends on the ring size
        batch_offset = state.batch
        for state.epoch in range(state.epoch, args.epochs):
            for batch_idx, (data, target) in enumerate(train_loader):
                if batch_idx >= batch_offset:
                    optimizer.zero_grad()
                    output = model(data)
                    loss = F.nll_loss(output, target)
                    loss.backward()
                    optimizer.step()
                    if state.batch % args.batches_per_commit == 0:
                        state.commit()                                                              # ref :72
                        print(f"Worker: {hvd.rank()}/{hvd.size()} | Batch: {state.batch}/{len(train_loader)} | "
                              f"Epoch: {state.epoch} | Loss: {loss.item()}", flush=True)
                    state.batch = batch_idx
            state.batch = 0
            batch_offset = 0

    def on_state_reset():                                                                           # ref :80-82
        for param_group in optimizer.param_groups:
            param_group["lr"] = args.lr / math.sqrt(hvd.size())

    def test():                                                                                     # ref :85-102
        test_dataset = SyntheticMNIST(args.test_samples, seed=1234, first_index=args.train_samples)   # held-out samples of the same stream
        model.eval()
        correct = total = 0
        with torch.no_grad():
            for images, labels in DeviceLoader(test_dataset, batch_size=128):
                outputs = model(images)
                _, predicted = torch.max(outputs, 1)
                total += labels.size(0)
                correct += (predicted == labels).sum().item()
        model.train()
        print(f"Accuracy: {(correct / total) * 100:.2f}%", flush=True)
        return correct / total

    state = hvd.elastic.TorchState(model, optimizer, batch=0, epoch=0)                              # ref :104
    state.register_reset_callbacks([on_state_reset])                                                # ref :105
    train(state)
    return test()


if __name__ == "__main__":
    main()
