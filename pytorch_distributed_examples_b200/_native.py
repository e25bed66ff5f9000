"""ctypes binding of libddp_step.so — the thin C-ABI seam (include/ddp_step.h).

Every method takes torch CUDA tensors, checks them, and passes raw ``data_ptr()``s plus the current
CUDA stream to the C entry points.  There is NO CPU fallback: a missing library or a CPU tensor raises.
Host-logic unit tests substitute a test double through ``set_instance`` (tests/fake_native.py).
"""
from __future__ import annotations

import ctypes
import os
from ctypes import POINTER, c_char_p, c_double, c_float, c_int, c_int32, c_int64, c_uint64, c_void_p

import torch

_LIB_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "lib")
LIB_PATH = os.environ.get("DDP_LIB_PATH") or os.path.join(_LIB_DIR, "libddp_step.so")

OPT_SGD, OPT_ADAM, OPT_ADAMW, OPT_ADADELTA = 0, 1, 2, 3
DT_F32, DT_BF16, DT_I64, DT_U8 = 0, 1, 2, 3
ARENA_ELEMS = 21848
PARAM_COUNT = 21840
UID_BYTES = 128
IPC_HANDLE_BYTES = 64
BUCKET_CHUNK_BYTES = 16384


class Hyper(ctypes.Structure):
    _fields_ = [("lr", c_double), ("momentum", c_double), ("dampening", c_double), ("weight_decay", c_double),
                ("beta1", c_double), ("beta2", c_double), ("eps", c_double), ("rho", c_double),
                ("nesterov", c_int), ("write_back_grad", c_int), ("bump_step_dev", c_int)]


class MLPShape(ctypes.Structure):
    """ddp_mlp_shape_t: Linear(in,F) -> ReLU -> hidden_layers x [Linear(F,F) -> ReLU] -> Linear(F,out)."""
    _fields_ = [("in_features", c_int), ("features", c_int), ("hidden_layers", c_int), ("out_features", c_int)]


DDP_EINVAL = -1000

# name -> (restype, argtypes); also the list of symbols the not-gpu test checks against include/ddp_step.h
SIGNATURES = {
    "ddp_abi_version": (c_int, []),
    "ddp_last_error": (c_char_p, []),
    "ddp_set_device": (c_int, [c_int]),
    "ddp_net_param_layout": (c_int, [POINTER(c_int64), POINTER(c_int64)]),
    "ddp_net_workspace_bytes": (c_int64, [c_int64]),
    "ddp_net_scratch_bytes": (c_int64, [c_int64]),
    "ddp_net_forward": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_uint64, c_uint64, c_void_p, c_int,
                                c_int64, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "ddp_net_train_fwdbwd": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_uint64, c_uint64, c_void_p, c_float,
                                     c_int64, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "ddp_net_backward": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_float, c_int, c_int64,
                                 c_void_p, c_void_p, c_void_p, c_void_p]),
    "ddp_bucket_copy": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_void_p, c_int,
                                c_void_p]),
    "ddp_fused_update": (c_int, [c_int, c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_int64, c_float,
                                 POINTER(Hyper), c_int64, c_void_p, c_void_p]),
    "ddp_step_increment": (c_int, [c_void_p, c_void_p]),
    "ddp_scale": (c_int, [c_void_p, c_int, c_int64, c_float, c_void_p]),
    "ddp_sampler_num_samples": (c_int64, [c_int64, c_int, c_int]),
    "ddp_sampler_indices": (c_int, [c_int64, c_int, c_int, c_int, c_uint64, c_uint64, c_int, c_void_p, c_void_p,
                                    c_void_p]),
    "ddp_sampler_indices_static": (c_int, [c_int64, c_int, c_int, c_int, c_uint64, c_uint64, c_int, c_void_p, c_void_p]),
    "ddp_synth_batch": (c_int, [c_uint64, c_void_p, c_int64, c_int64, c_void_p, c_void_p, c_void_p]),
    "ddp_synth_batch_u8": (c_int, [c_uint64, c_void_p, c_int64, c_int64, c_void_p, c_void_p, c_void_p]),
    "ddp_normalize_u8": (c_int, [c_void_p, c_void_p, c_int64, c_float, c_float, c_void_p]),
    "ddp_comm_unique_id": (c_int, [c_void_p]),
    "ddp_comm_init": (c_int, [POINTER(c_void_p), c_void_p, c_int, c_int]),
    "ddp_comm_destroy": (c_int, [c_void_p]),
    "ddp_comm_abort": (c_int, [c_void_p]),
    "ddp_comm_async_error": (c_int, [c_void_p, POINTER(c_int)]),
    "ddp_allreduce_sum": (c_int, [c_void_p, c_void_p, c_int64, c_int, c_void_p]),
    "ddp_allreduce_max": (c_int, [c_void_p, c_void_p, c_int64, c_int, c_void_p]),
    "ddp_broadcast": (c_int, [c_void_p, c_void_p, c_int64, c_int, c_void_p]),
    "ddp_peer_create": (c_int, [POINTER(c_void_p), c_int, c_int, c_int64]),
    "ddp_peer_local_handle": (c_int, [c_void_p, c_void_p]),
    "ddp_peer_connect": (c_int, [c_void_p, c_void_p]),
    "ddp_peer_destroy": (c_int, [c_void_p]),
    "ddp_peer_allreduce": (c_int, [c_void_p, c_void_p, c_int64, c_void_p]),
    "ddp_peer_allreduce_update": (c_int, [c_void_p, c_void_p, c_int64, c_int, c_void_p, c_void_p, c_void_p, POINTER(Hyper),
                                          c_int64, c_void_p, c_void_p]),
    "ddp_net_train_step_peer": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_uint64, c_uint64, c_void_p, c_int64,
                                        c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_void_p,
                                        POINTER(Hyper), c_int64, c_void_p, c_int, c_void_p]),
    "ddp_sym_flag_bytes": (c_int64, []),
    "ddp_sym_allreduce": (c_int, [POINTER(ctypes.c_uint64), ctypes.c_uint64, c_int64, c_int, c_int, c_int64, c_int64, c_int,
                                  c_float, c_int, c_int, c_void_p]),
    "ddp_mlp_param_count": (c_int64, [c_void_p]),
    "ddp_mlp_param_layout": (c_int, [c_void_p, POINTER(c_int64), POINTER(c_int64), c_int]),
    "ddp_mlp_workspace_bytes": (c_int64, [c_void_p, c_int64]),
    "ddp_mlp_forward": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_void_p, c_void_p, c_void_p, c_void_p]),
    "ddp_mlp_backward": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_float, c_int64, c_void_p, c_void_p, c_void_p]),
    "ddp_split_tf32": (c_int, [c_void_p, c_int64, c_int64, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "ddp_gemm_3xtf32": (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_int, c_void_p]),
    "ddp_state_snapshot": (c_int, [c_void_p, c_void_p, c_int64, c_void_p]),
    "ddp_state_restore": (c_int, [c_void_p, c_void_p, c_int64, c_void_p]),
    "ddp_prof_enable": (c_int, [c_int]),
    "ddp_prof_read": (c_int, [POINTER(c_float)]),
}
PROF_KERNELS = ("conv1_fwd", "conv2_fwd", "head_fwd", "head_bwd", "conv2_dgrad", "conv2_wgrad", "conv1_wgrad",
                "reduce_partials")

_DTYPES = {torch.float32: DT_F32, torch.bfloat16: DT_BF16, torch.int64: DT_I64, torch.uint8: DT_U8}


def load_library(path: str = LIB_PATH) -> ctypes.CDLL:
    if not os.path.exists(path):
        raise RuntimeError(
            f"libddp_step.so not found at {path}: build it with `python -c 'import __graft_entry__ as g; g.build()'`"
            " — this package has no CPU or PyTorch fallback for its CUDA path")
    lib = ctypes.CDLL(path)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    return lib


def _ptr(t, name="tensor", dtype=None):
    if t is None:
        return None
    if not isinstance(t, torch.Tensor) or not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor (this path has no CPU fallback)")
    if not t.is_contiguous():
        raise RuntimeError(f"{name} must be contiguous")
    if dtype is not None and t.dtype != dtype:
        raise RuntimeError(f"{name} must be {dtype}, got {t.dtype}")
    return t.data_ptr()


def _stream():
    return torch.cuda.current_stream().cuda_stream


class CollectiveError(RuntimeError):
    """A collective (NCCL or NVLink peer-memory) entry point failed, reported an asynchronous error, or did not
    finish before its deadline: a peer is lost or the ring is broken.  ``hvd.HorovodInternalError`` is this class;
    ``hvd.elastic.run`` answers it with restore() + re-formation (ref: horovod/horovod_mnist_elastic.py:55)."""


class Native:
    """Tensor-level façade over the C-ABI.  One instance per process."""

    def __init__(self, lib: ctypes.CDLL | None = None):
        self.lib = lib or load_library()
        if self.lib.ddp_abi_version() != 1:
            raise RuntimeError("libddp_step.so ABI version mismatch")
        self._dev_bound = set()

    # -- plumbing ---------------------------------------------------------------------------------
    def _check(self, rc: int, what: str, collective: bool = False):
        if rc != 0:
            msg = self.lib.ddp_last_error()
            text = f"{what} failed (rc={rc}): {msg.decode() if msg else ''}"
            # argument errors (DDP_EINVAL / DDP_ENOSYS, <= -1000) stay plain RuntimeErrors: they are caller bugs, not lost peers
            if collective and rc > -1000:
                raise CollectiveError(text)
            raise RuntimeError(text)

    def bind_device(self, device: int):
        self._check(self.lib.ddp_set_device(int(device)), "ddp_set_device")

    def param_layout(self):
        off = (c_int64 * 8)()
        sz = (c_int64 * 8)()
        self._check(self.lib.ddp_net_param_layout(off, sz), "ddp_net_param_layout")
        return list(off), list(sz)

    def workspace_bytes(self, batch: int) -> int:
        return int(self.lib.ddp_net_workspace_bytes(batch))

    def scratch_bytes(self, batch: int) -> int:
        return int(self.lib.ddp_net_scratch_bytes(batch))

    # -- Net ----------------------------------------------------------------------------------------
    def net_forward(self, params, x, target, mask2d, mask1d, seed, offset, training, workspace, out_logp,
                    out_loss=None, out_argmax=None, out_correct=None, scratch=None, offset_dev=None):
        batch = x.shape[0]
        rc = self.lib.ddp_net_forward(
            _ptr(params, "params", torch.float32), _ptr(x, "x", torch.float32), _ptr(target, "target", torch.int64),
            _ptr(mask2d, "mask2d", torch.float32), _ptr(mask1d, "mask1d", torch.float32),
            int(seed) & (2 ** 64 - 1), int(offset) & (2 ** 64 - 1), _ptr(offset_dev, "offset_dev", torch.int64),
            int(bool(training)), batch, _ptr(workspace, "workspace"), _ptr(out_logp, "out_logp", torch.float32),
            _ptr(out_loss, "out_loss", torch.float32), _ptr(out_argmax, "out_argmax", torch.int64),
            _ptr(out_correct, "out_correct", torch.int64), _ptr(scratch, "scratch"), _stream())
        self._check(rc, "ddp_net_forward")

    def net_train_fwdbwd(self, params, x, target, mask2d, mask1d, seed, offset, workspace, out_logp, out_loss, scratch,
                         grad_bucket, loss_scale=1.0, out_argmax=None, out_correct=None, offset_dev=None):
        """Forward + backward of one training step (fused head kernel); gradients land in ``grad_bucket``."""
        rc = self.lib.ddp_net_train_fwdbwd(
            _ptr(params, "params", torch.float32), _ptr(x, "x", torch.float32), _ptr(target, "target", torch.int64),
            _ptr(mask2d, "mask2d", torch.float32), _ptr(mask1d, "mask1d", torch.float32),
            int(seed) & (2 ** 64 - 1), int(offset) & (2 ** 64 - 1), _ptr(offset_dev, "offset_dev", torch.int64),
            float(loss_scale), x.shape[0], _ptr(workspace, "workspace"), _ptr(out_logp, "out_logp", torch.float32),
            _ptr(out_loss, "out_loss", torch.float32), _ptr(out_argmax, "out_argmax", torch.int64),
            _ptr(out_correct, "out_correct", torch.int64), _ptr(scratch, "scratch"),
            _ptr(grad_bucket, "grad_bucket", torch.float32), _stream())
        self._check(rc, "ddp_net_train_fwdbwd")

    def net_backward(self, params, x, logp, dlogp, target, loss_scale, training, workspace, scratch, grad_bucket):
        rc = self.lib.ddp_net_backward(
            _ptr(params, "params", torch.float32), _ptr(x, "x", torch.float32), _ptr(logp, "logp", torch.float32),
            _ptr(dlogp, "dlogp", torch.float32), _ptr(target, "target", torch.int64), float(loss_scale),
            int(bool(training)), x.shape[0], _ptr(workspace, "workspace"), _ptr(scratch, "scratch"),
            _ptr(grad_bucket, "grad_bucket", torch.float32), _stream())
        self._check(rc, "ddp_net_backward")

    # -- bucket -------------------------------------------------------------------------------------
    def bucket_copy(self, table, bucket, direction: int):
        """table: BucketTable (device tensors built once per layout, see bucket.py)."""
        rc = self.lib.ddp_bucket_copy(_ptr(table.ptrs), _ptr(table.offsets), _ptr(table.nbytes),
                                      _ptr(table.chunk_item), _ptr(table.chunk_start), table.nchunks,
                                      _ptr(bucket, "bucket"), direction, _stream())
        self._check(rc, "ddp_bucket_copy")

    # -- optimizer ----------------------------------------------------------------------------------
    def fused_update(self, opt, grad, params, state1, state2, n, inv_world, hyper: Hyper, step=1, step_dev=None):
        gd = _DTYPES.get(grad.dtype)
        if gd not in (DT_F32, DT_BF16):
            raise RuntimeError(f"unsupported gradient dtype {grad.dtype}")
        rc = self.lib.ddp_fused_update(opt, _ptr(grad, "grad"), gd, _ptr(params, "params", torch.float32),
                                       _ptr(state1, "state1", torch.float32), _ptr(state2, "state2", torch.float32),
                                       int(n), float(inv_world), ctypes.byref(hyper), int(step),
                                       _ptr(step_dev, "step_dev", torch.int64), _stream())
        self._check(rc, "ddp_fused_update")

    def step_increment(self, step_dev):
        self._check(self.lib.ddp_step_increment(_ptr(step_dev, "step_dev", torch.int64), _stream()),
                    "ddp_step_increment")

    def scale(self, buf, s: float):
        self._check(self.lib.ddp_scale(_ptr(buf, "buf"), _DTYPES[buf.dtype], buf.numel(), float(s), _stream()),
                    "ddp_scale")

    # -- sampler / data -----------------------------------------------------------------------------
    def sampler_num_samples(self, n, world, drop_last) -> int:
        return int(self.lib.ddp_sampler_num_samples(n, world, int(bool(drop_last))))

    def sampler_indices(self, n, world, rank, shuffle, seed, epoch, drop_last, perm_scratch, out):
        rc = self.lib.ddp_sampler_indices(n, world, rank, int(bool(shuffle)), int(seed) & (2 ** 64 - 1),
                                          int(epoch) & (2 ** 64 - 1), int(bool(drop_last)),
                                          _ptr(perm_scratch, "perm_scratch", torch.int32),
                                          _ptr(out, "out", torch.int64), _stream())
        self._check(rc, "ddp_sampler_indices")

    def sampler_indices_static(self, n, world, rank, shuffle, seed, epoch, drop_last, out):
        """Shard indices without a caller-provided permutation buffer: for callers that cannot allocate device
        memory.  Uses the 256 KiB scratch array built into the library (n <= 65536); calls must not overlap."""
        rc = self.lib.ddp_sampler_indices_static(n, world, rank, int(bool(shuffle)), int(seed) & (2 ** 64 - 1),
                                                 int(epoch) & (2 ** 64 - 1), int(bool(drop_last)),
                                                 _ptr(out, "out", torch.int64), _stream())
        self._check(rc, "ddp_sampler_indices_static")

    def synth_batch(self, seed, indices, base_index, batch, out_x, out_y):
        rc = self.lib.ddp_synth_batch(int(seed) & (2 ** 64 - 1), _ptr(indices, "indices", torch.int64),
                                      int(base_index), int(batch), _ptr(out_x, "out_x", torch.float32),
                                      _ptr(out_y, "out_y", torch.int64), _stream())
        self._check(rc, "ddp_synth_batch")

    def synth_batch_u8(self, seed, indices, base_index, batch, out_px, out_y):
        rc = self.lib.ddp_synth_batch_u8(int(seed) & (2 ** 64 - 1), _ptr(indices, "indices", torch.int64),
                                         int(base_index), int(batch), _ptr(out_px, "out_px", torch.uint8),
                                         _ptr(out_y, "out_y", torch.int64), _stream())
        self._check(rc, "ddp_synth_batch_u8")

    def normalize_u8(self, pixels, out, mean=0.1307, std=0.3081):
        """ToTensor + Normalize on the device: out = (u8/255 - mean)/std."""
        if out.numel() != pixels.numel():
            raise RuntimeError("normalize_u8: size mismatch")
        rc = self.lib.ddp_normalize_u8(_ptr(pixels, "pixels", torch.uint8), _ptr(out, "out", torch.float32),
                                       pixels.numel(), float(mean), float(std), _stream())
        self._check(rc, "ddp_normalize_u8")

    # -- collectives --------------------------------------------------------------------------------
    def comm_unique_id(self) -> bytes:
        buf = ctypes.create_string_buffer(UID_BYTES)
        self._check(self.lib.ddp_comm_unique_id(buf), "ddp_comm_unique_id")
        return buf.raw

    def comm_init(self, uid: bytes, world: int, rank: int):
        handle = c_void_p()
        buf = ctypes.create_string_buffer(uid, UID_BYTES)
        self._check(self.lib.ddp_comm_init(ctypes.byref(handle), buf, world, rank), "ddp_comm_init", collective=True)
        return handle

    def comm_destroy(self, comm):
        self._check(self.lib.ddp_comm_destroy(comm), "ddp_comm_destroy")

    def comm_abort(self, comm):
        self._check(self.lib.ddp_comm_abort(comm), "ddp_comm_abort")

    def comm_async_error(self, comm) -> int:
        """ncclCommGetAsyncError: 0 healthy, 7 in progress, else broken."""
        res = c_int(0)
        self._check(self.lib.ddp_comm_async_error(comm, ctypes.byref(res)), "ddp_comm_async_error", collective=True)
        return int(res.value)

    def allreduce_sum(self, comm, buf):
        self._check(self.lib.ddp_allreduce_sum(comm, _ptr(buf, "buf"), buf.numel(), _DTYPES[buf.dtype], _stream()),
                    "ddp_allreduce_sum", collective=True)

    def allreduce_max(self, comm, buf):
        self._check(self.lib.ddp_allreduce_max(comm, _ptr(buf, "buf"), buf.numel(), _DTYPES[buf.dtype], _stream()),
                    "ddp_allreduce_max", collective=True)

    def broadcast(self, comm, buf, root: int):
        self._check(self.lib.ddp_broadcast(comm, _ptr(buf, "buf"), buf.numel() * buf.element_size(), root, _stream()),
                    "ddp_broadcast", collective=True)

    # -- NVLink peer-memory allreduce (+ fused update) ----------------------------------------------
    def peer_create(self, world: int, rank: int, max_elems: int):
        handle = c_void_p()
        self._check(self.lib.ddp_peer_create(ctypes.byref(handle), world, rank, int(max_elems)), "ddp_peer_create")
        return handle

    def peer_local_handle(self, peer) -> bytes:
        buf = ctypes.create_string_buffer(IPC_HANDLE_BYTES)
        self._check(self.lib.ddp_peer_local_handle(peer, buf), "ddp_peer_local_handle")
        return buf.raw

    def peer_connect(self, peer, handles: bytes):
        buf = ctypes.create_string_buffer(handles, len(handles))
        self._check(self.lib.ddp_peer_connect(peer, buf), "ddp_peer_connect", collective=True)

    def peer_destroy(self, peer):
        self._check(self.lib.ddp_peer_destroy(peer), "ddp_peer_destroy")

    def peer_allreduce(self, peer, bucket):
        self._check(self.lib.ddp_peer_allreduce(peer, _ptr(bucket, "bucket", torch.float32), bucket.numel(), _stream()),
                    "ddp_peer_allreduce", collective=True)

    def net_train_step_peer(self, params, x, target, seed, offset, workspace, out_logp, out_loss, scratch, grad_bucket, peer,
                            opt, state1, state2, hyper: Hyper, step=1, step_dev=None, offset_dev=None, fused_head=False):
        """The whole training step of one rank (forward, loss, backward, NVLink peer exchange, averaged update)."""
        rc = self.lib.ddp_net_train_step_peer(
            _ptr(params, "params", torch.float32), _ptr(x, "x", torch.float32), _ptr(target, "target", torch.int64), None, None,
            int(seed) & (2 ** 64 - 1), int(offset) & (2 ** 64 - 1), _ptr(offset_dev, "offset_dev", torch.int64), x.shape[0],
            _ptr(workspace, "workspace"), _ptr(out_logp, "out_logp", torch.float32), _ptr(out_loss, "out_loss", torch.float32),
            _ptr(scratch, "scratch"), _ptr(grad_bucket, "grad_bucket", torch.float32), peer, int(opt),
            _ptr(state1, "state1", torch.float32), _ptr(state2, "state2", torch.float32), ctypes.byref(hyper), int(step),
            _ptr(step_dev, "step_dev", torch.int64), int(bool(fused_head)), _stream())
        self._check(rc, "ddp_net_train_step_peer", collective=True)

    def peer_allreduce_update(self, peer, opt, bucket, params, state1, state2, n, hyper: Hyper, step=1, step_dev=None):
        rc = self.lib.ddp_peer_allreduce_update(peer, _ptr(bucket, "bucket", torch.float32), int(n), opt,
                                                _ptr(params, "params", torch.float32),
                                                _ptr(state1, "state1", torch.float32), _ptr(state2, "state2", torch.float32),
                                                ctypes.byref(hyper), int(step), _ptr(step_dev, "step_dev", torch.int64),
                                                _stream())
        self._check(rc, "ddp_peer_allreduce_update", collective=True)

    # -- two-shot allreduce in symmetric memory ------------------------------------------------------
    def sym_flag_bytes(self) -> int:
        return int(self.lib.ddp_sym_flag_bytes())

    def sym_allreduce(self, replicas, multicast, flag_offset, world, rank, byte_offset, n, dtype, scale, max_ctas=0, unroll=0):
        """In-place scale * sum over ranks of n elements at byte_offset of a symmetric buffer (see ddp_step.h)."""
        arr = (ctypes.c_uint64 * world)(*[int(a) for a in replicas])
        rc = self.lib.ddp_sym_allreduce(arr, int(multicast or 0), int(flag_offset), int(world), int(rank), int(byte_offset),
                                        int(n), _DTYPES[dtype], float(scale), int(max_ctas), int(unroll), _stream())
        self._check(rc, "ddp_sym_allreduce", collective=True)

    # -- MLP of the DDP script (tensor-core 3xTF32 GEMMs) ---------------------------------------------
    def mlp_param_layout(self, shape: "MLPShape"):
        """[(offset, size)] in floats, reference state_dict order, and the padded arena length."""
        n = 2 * (shape.hidden_layers + 2)
        offs, sizes = (c_int64 * n)(), (c_int64 * n)()
        rc = self.lib.ddp_mlp_param_layout(ctypes.byref(shape), offs, sizes, n)
        if rc != n:
            self._check(rc if rc < 0 else DDP_EINVAL, "ddp_mlp_param_layout")
        total = int(self.lib.ddp_mlp_param_count(ctypes.byref(shape)))
        return [(int(offs[i]), int(sizes[i])) for i in range(n)], total

    def mlp_workspace_bytes(self, shape: "MLPShape", batch: int) -> int:
        v = int(self.lib.ddp_mlp_workspace_bytes(ctypes.byref(shape), int(batch)))
        if v < 0:
            self._check(DDP_EINVAL, "ddp_mlp_workspace_bytes")
        return v

    def mlp_forward(self, shape, params, x, target, workspace, out_logits=None, out_loss=None):
        batch = x.shape[0]
        rc = self.lib.ddp_mlp_forward(ctypes.byref(shape), _ptr(params, "params", torch.float32), _ptr(x, "x", torch.float32),
                                      _ptr(target, "target", torch.int64), batch, _ptr(workspace, "workspace"),
                                      _ptr(out_logits, "out_logits", torch.float32), _ptr(out_loss, "out_loss", torch.float32),
                                      _stream())
        self._check(rc, "ddp_mlp_forward")

    def mlp_backward(self, shape, params, dlogits, target, loss_scale, batch, workspace, grad_bucket):
        rc = self.lib.ddp_mlp_backward(ctypes.byref(shape), _ptr(params, "params", torch.float32),
                                       _ptr(dlogits, "dlogits", torch.float32), _ptr(target, "target", torch.int64),
                                       float(loss_scale), int(batch), _ptr(workspace, "workspace"),
                                       _ptr(grad_bucket, "grad_bucket", torch.float32), _stream())
        self._check(rc, "ddp_mlp_backward")

    def split_tf32(self, src, transposed=False):
        """(hi, lo[, hiT, loT]) tf32-exact planes of a 2-D fp32 tensor."""
        rows, cols = src.shape
        hi, lo = torch.empty_like(src), torch.empty_like(src)
        hiT = torch.empty(cols, rows, dtype=torch.float32, device=src.device) if transposed else None
        loT = torch.empty_like(hiT) if transposed else None
        rc = self.lib.ddp_split_tf32(_ptr(src, "src", torch.float32), rows, cols, _ptr(hi, "hi"), _ptr(lo, "lo"),
                                     _ptr(hiT, "hiT"), _ptr(loT, "loT"), _stream())
        self._check(rc, "ddp_split_tf32")
        return (hi, lo, hiT, loT) if transposed else (hi, lo)

    def gemm_3xtf32(self, a_hi, a_lo, b_hi, b_lo, out=None, a_mn_major=False, b_mn_major=False):
        """out[M][N] = A @ B^T from tf32 hi/lo planes.  A is [M][K] (or [K][M] with a_mn_major), B is [N][K]
        (or [K][N] with b_mn_major)."""
        (K, M) = a_hi.shape if a_mn_major else a_hi.shape[::-1]
        N = b_hi.shape[1] if b_mn_major else b_hi.shape[0]
        if (b_hi.shape[0] if b_mn_major else b_hi.shape[1]) != K:
            raise RuntimeError("gemm_3xtf32: K of A and B differ")
        if out is None:
            out = torch.empty(M, N, dtype=torch.float32, device=a_hi.device)
        if not out.is_cuda or out.dtype != torch.float32 or out.stride(1) != 1:
            raise RuntimeError("out must be a CUDA fp32 matrix with unit column stride (a row-strided view is fine)")
        rc = self.lib.ddp_gemm_3xtf32(_ptr(a_hi, "a_hi", torch.float32), _ptr(a_lo, "a_lo", torch.float32), int(a_mn_major),
                                      _ptr(b_hi, "b_hi", torch.float32), _ptr(b_lo, "b_lo", torch.float32), int(b_mn_major),
                                      M, N, K, ctypes.c_void_p(out.data_ptr()), out.stride(0), _stream())
        self._check(rc, "ddp_gemm_3xtf32")
        return out

    # -- elastic snapshot ---------------------------------------------------------------------------
    def state_snapshot(self, device_arena, pinned_host):
        nbytes = device_arena.numel() * device_arena.element_size()
        if pinned_host.is_cuda or pinned_host.numel() * pinned_host.element_size() < nbytes:
            raise RuntimeError("pinned_host must be a host tensor at least as large as the arena")
        self._check(self.lib.ddp_state_snapshot(_ptr(device_arena, "arena"), pinned_host.data_ptr(), nbytes,
                                                _stream()), "ddp_state_snapshot")

    def state_restore(self, device_arena, pinned_host):
        nbytes = device_arena.numel() * device_arena.element_size()
        self._check(self.lib.ddp_state_restore(_ptr(device_arena, "arena"), pinned_host.data_ptr(), nbytes,
                                               _stream()), "ddp_state_restore")


    # -- per-kernel timing aid ----------------------------------------------------------------------
    def prof_enable(self, on: bool):
        self._check(self.lib.ddp_prof_enable(int(bool(on))), "ddp_prof_enable")

    def prof_read(self):
        ms = (c_float * len(PROF_KERNELS))()
        self._check(self.lib.ddp_prof_read(ms), "ddp_prof_read")
        return dict(zip(PROF_KERNELS, [float(v) for v in ms]))


_INSTANCE = None


def get() -> Native:
    """The process-wide binding.  Raises if the CUDA library has not been built."""
    global _INSTANCE
    if _INSTANCE is None:
        _INSTANCE = Native()
    return _INSTANCE


def set_instance(obj):
    """Install a test double (host-logic unit tests only)."""
    global _INSTANCE
    _INSTANCE = obj
