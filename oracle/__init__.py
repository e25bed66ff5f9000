"""CPU oracle for the data-parallel training step — TEST INFRASTRUCTURE, not product code.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline / ``--impl reference``
legs may import this package, and only as the checker or the timed CPU baseline.  Nothing under
``pytorch_distributed_examples_b200/`` imports it; the product path fails loudly without its CUDA library.

What it restates (reference = /root/reference = ArnauGabrielAtienza/pytorch_distributed_examples):
  net_oracle.py      Net.forward / loss / backward     ref: horovod/mnist_horovod.py:9-25,58-64
  optim_oracle.py    SGD / Adam / AdamW / Adadelta     ref: horovod/mnist_horovod.py:50 ; horovod_mnist_elastic.py:41 ;
                                                        pytorch_elastic/mnist_ddp_elastic.py:173
  sampler_oracle.py  DistributedSampler shard indices  ref: horovod/mnist_horovod.py:41-42 ; mnist_ddp_elastic.py:188,89
  synth_oracle.py    Philox4x32-10 synthetic MNIST     (replaces ref: horovod/mnist_horovod.py:34-38)
  dp_oracle.py       DistributedOptimizer / DDP step semantics over W ranks, elastic commit/restore
  cpu_baseline.py    the reference training loop timed on host cores

The arithmetic of the reference lives in un-vendored third-party packages:
  torch       pinned 1.12.1 by ref: horovod/Dockerfile:5,63 — the oracle runs the installed 2.11.0 (skew stated);
  horovod     unpinned (built from the checkout around the Dockerfile, ref: horovod/Dockerfile:75-79), absent here:
              its DistributedOptimizer/broadcast/elastic semantics are restated from its public API.

Pinning: the reference holds NO tests, golden vectors or known-answer values, so the
oracle is pinned against outputs of the reference's own classes run in the build container:
``oracle/gen_golden.py`` extracts ``Net`` from the reference scripts (by AST, without copying source),
imports ``Model``/``Trainer`` from mnist_ddp_elastic.py, runs them under the installed torch (incl. a real
2-process gloo DDP run) and commits inputs/outputs as ``tests/golden/*.pt``.  The Horovod wrapper itself
cannot be run (package absent): for that boundary parity is "unpinned" and rests on the restated semantics.
"""
