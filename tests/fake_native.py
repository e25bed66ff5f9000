"""Test double for the C-ABI binding so HOST LOGIC (bucket layout, optimizer wrapper, elastic state machine,
communicator bootstrap) can run on CPU under gloo.  Kernel arithmetic is delegated to plain torch CPU ops;
collectives go through torch.distributed.  Lives in tests/ — the product never imports it."""
import torch
import torch.distributed as dist

from pytorch_distributed_examples_b200 import _native


class FakeNative:
    allows_cpu = True

    def bind_device(self, device):
        pass

    def param_layout(self):
        return [0, 252, 264, 5264, 5284, 21284, 21336, 21836], [250, 10, 5000, 20, 16000, 50, 500, 10]

    # bucket: emulate the byte gather/scatter with the table's python-side mirror
    def bucket_copy(self, table, bucket, direction):
        for t, off in zip(table._tensors, table._offsets_elems):
            n = t.numel()
            if direction == 0:
                bucket[off:off + n].copy_(t.reshape(-1))
            else:
                t.copy_(bucket[off:off + n].view(t.shape))

    def fused_update(self, opt, grad, params, state1, state2, n, inv_world, hyper, step=1, step_dev=None):
        if step_dev is not None:
            step = int(step_dev.item()) + 1
        g = grad.float() * inv_world
        lr, wd = hyper.lr, hyper.weight_decay
        if opt == _native.OPT_SGD:
            if wd:
                g = g + wd * params
            if hyper.momentum:
                if step <= 1:
                    state1.copy_(g)
                else:
                    state1.mul_(hyper.momentum).add_(g, alpha=1 - hyper.dampening)
                g = g + hyper.momentum * state1 if hyper.nesterov else state1
            params.add_(g, alpha=-lr)
        elif opt in (_native.OPT_ADAM, _native.OPT_ADAMW):
            if opt == _native.OPT_ADAMW:
                params.mul_(1 - lr * wd)
            elif wd:
                g = g + wd * params
            state1.lerp_(g, 1 - hyper.beta1)
            state2.mul_(hyper.beta2).addcmul_(g, g, value=1 - hyper.beta2)
            bc1, bc2 = 1 - hyper.beta1 ** step, 1 - hyper.beta2 ** step
            denom = (state2.sqrt() / bc2 ** 0.5).add_(hyper.eps)
            params.addcdiv_(state1, denom, value=-lr / bc1)
        else:
            raise NotImplementedError
        if hyper.write_back_grad:
            grad.copy_(g)
        if hyper.bump_step_dev and step_dev is not None:
            step_dev += 1

    def step_increment(self, step_dev):
        step_dev += 1

    def scale(self, buf, s):
        buf.mul_(s)

    def sampler_num_samples(self, n, world, drop_last):
        import math
        return math.ceil((n - world) / world) if (drop_last and n % world) else math.ceil(n / world)

    # collectives over the test's gloo process group
    def comm_unique_id(self):
        return b"\x07" * _native.UID_BYTES

    def comm_init(self, uid, world, rank):
        assert uid == b"\x07" * _native.UID_BYTES
        assert world == dist.get_world_size(), "the double only forms full-world rings (sub-rings are world 1 here)"
        return dist.group.WORLD

    def sampler_indices(self, n, world, rank, shuffle, seed, epoch, drop_last, perm_scratch, out):
        from oracle import sampler_oracle
        out.copy_(torch.from_numpy(sampler_oracle.shard_indices(n, world, rank, shuffle, seed, epoch, drop_last)))

    def comm_destroy(self, comm):
        pass

    def allreduce_sum(self, comm, buf):
        dist.all_reduce(buf, group=comm)

    def allreduce_max(self, comm, buf):
        dist.all_reduce(buf, op=dist.ReduceOp.MAX, group=comm)

    def broadcast(self, comm, buf, root):
        dist.broadcast(buf, src=dist.get_global_rank(comm, root), group=comm)

    def state_snapshot(self, arena, host):
        host.copy_(arena)

    def state_restore(self, arena, host):
        arena.copy_(host)
