"""`pytorch_elastic/mnist_ddp_elastic.py` of the reference, switched to the B200 path.

The reference's CLI is kept (ref: pytorch_elastic/mnist_ddp_elastic.py:202-213):

    torchrun --standalone --nproc-per-node=8 examples/mnist_ddp_elastic.py 10 5 [--batch_size 128]
                                                                            ^total_epochs ^save_every

Its pieces map one to one: `ddp_setup` picks the NCCL backend, the script's `Model(hidden_layers=5, features=1024)` is
this package's tensor-core `Model` (same constructor, same state_dict keys), stock DDP stays and gets
`allreduce_hook`, `DistributedSampler` is the device-side one, snapshots keep the `snapshot.pt` format (written off
the training stream), `Trainer.test` counts on the device and reports the accuracy of the WHOLE test set.  A job that
torchrun restarts after a failure resumes from `snapshot.pt` exactly like the reference (ref :54-56,61-68).
"""
import argparse
import os
import time

import _common  # noqa: F401
import torch
import torch.nn as nn
from torch.distributed import destroy_process_group, init_process_group
from torch.nn.parallel import DistributedDataParallel as DDP

from pytorch_distributed_examples_b200 import (AsyncSnapshotWriter, DeviceLoader, DistributedSampler, HookState, Model,
                                               SyntheticMNIST, allreduce_hook, evaluate, hvd, load_snapshot)


def ddp_setup():                                                             # ref :22-27
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
    init_process_group(backend="nccl")
    hvd.init()                                                               # the package's own ring for the hook


class Trainer:                                                               # ref :30-130
    def __init__(self, model, train_data, test_data, optimizer, criterion, save_every, snapshot_path):
        self.local_rank = int(os.environ.get("LOCAL_RANK", "0"))
        self.global_rank = int(os.environ.get("RANK", "0"))
        self.train_data, self.test_data = train_data, test_data
        self.optimizer, self.criterion = optimizer, criterion
        self.save_every = save_every
        self.epochs_run = 0
        self.snapshot_path = snapshot_path
        model = model.cuda()
        if os.path.exists(snapshot_path):                                    # ref :54-56: resume before the DDP wrap
            self.epochs_run = load_snapshot(model, snapshot_path)
            print(f"Resuming training from snapshot at Epoch {self.epochs_run}", flush=True)
        self.model = DDP(model, device_ids=[self.local_rank], gradient_as_bucket_view=True)      # ref :58
        self.model.register_comm_hook(HookState(), allreduce_hook)           # bucket allreduce through libddp_step.so
        self.writer = AsyncSnapshotWriter(snapshot_path)

    def _run_batch(self, source, targets):                                   # ref :70-75, unchanged
        self.optimizer.zero_grad()
        output = self.model(source)
        loss = self.criterion(output, targets)
        loss.backward()
        self.optimizer.step()

    def _run_epoch(self, epoch):                                             # ref :78-88
        self.model.train()
        print(f"Local Rank: {self.local_rank} | Global Rank: {self.global_rank} | Epoch {epoch} | "
              f"Batchsize: {self.train_data.batch_size} | Steps: {len(self.train_data)}", flush=True)
        self.train_data.sampler.set_epoch(epoch)
        for source, targets in self.train_data:
            self._run_batch(source, targets)
        return self.test()

    def train(self, max_epochs):                                             # ref :106-114
        acc = None
        for epoch in range(self.epochs_run, max_epochs):
            acc = self._run_epoch(epoch)
            if self.local_rank == 0 and epoch % self.save_every == 0:
                self.writer.save(self.model, epoch)                          # ref :95-104, off the training stream
        self.writer.flush()
        return acc

    def test(self):                                                          # ref :117-130
        self.model.eval()
        correct, total, acc = evaluate(self.model, self.test_data, hvd._require())
        print(f"Test accuracy: {acc * 100:.2f}%", flush=True)
        return acc


def load_train_objs(args):                                                   # ref :162-175
    train_set = SyntheticMNIST(args.train_samples, seed=1234)
    test_set = SyntheticMNIST(args.test_samples, seed=1234, first_index=args.train_samples)   # held-out samples of the same stream
    model = Model(hidden_layers=args.hidden_layers, features=args.features)
    optimizer = torch.optim.Adam(model.parameters(), lr=1e-3)
    criterion = nn.CrossEntropyLoss()
    return train_set, test_set, model, optimizer, criterion


def prepare_dataloader(dataset, batch_size):                                 # ref :178-189
    return DeviceLoader(dataset, batch_size=batch_size, sampler=DistributedSampler(dataset))


def main(save_every, total_epochs, batch_size, args, snapshot_path="snapshot.pt"):       # ref :192-199
    ddp_setup()
    dataset, test_dataset, model, optimizer, criterion = load_train_objs(args)
    train_data = prepare_dataloader(dataset, batch_size)
    test_data = prepare_dataloader(test_dataset, batch_size)
    trainer = Trainer(model, train_data, test_data, optimizer, criterion, save_every, snapshot_path)
    acc = trainer.train(total_epochs)
    destroy_process_group()
    return acc


if __name__ == "__main__":
    parser = argparse.ArgumentParser(description="simple distributed training job")
    parser.add_argument("total_epochs", type=int, help="Total epochs to train the model")
    parser.add_argument("save_every", type=int, help="How often to save a snapshot")
    parser.add_argument("--batch_size", default=128, type=int, help="Input batch size on each device (default: 128)")
    # not in the reference: sizes for short runs (the defaults are the reference's)
    parser.add_argument("--train_samples", default=60000, type=int)
    parser.add_argument("--test_samples", default=10000, type=int)
    parser.add_argument("--hidden_layers", default=5, type=int)
    parser.add_argument("--features", default=1024, type=int)
    parser.add_argument("--snapshot_path", default="snapshot.pt")
    cli = parser.parse_args()
    start = time.time()
    main(cli.save_every, cli.total_epochs, cli.batch_size, cli, cli.snapshot_path)
    print(f"Execution time: {time.time() - start}")
