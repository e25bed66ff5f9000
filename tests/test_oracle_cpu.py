"""CPU suite: the oracle against the committed golden vectors (which were produced by the REFERENCE's own
classes, oracle/gen_golden.py) and against torch itself for the integer paths."""
import hashlib

import numpy as np
import pytest
import torch

from oracle import net_oracle, sampler_oracle, synth_oracle
from helpers import load_golden


@pytest.mark.parametrize("name", ["fwdbwd_b1.pt", "fwdbwd_b8.pt", "fwdbwd_b48.pt", "fwdbwd_b64.pt"])
def test_oracle_forward_backward_matches_reference_golden(name):
    fx = load_golden(name)
    logp, loss, grads = net_oracle.forward_backward(fx["state_dict"], fx["x"], fx["y"], fx["mask2d"], fx["mask1d"])
    # generated on this image bit-for-bit; another host CPU may pick other oneDNN kernels -> tiny tolerance
    torch.testing.assert_close(logp, fx["logp"], rtol=1e-6, atol=1e-6)
    torch.testing.assert_close(loss, fx["loss"], rtol=1e-6, atol=1e-7)
    for k, g in fx["grads"].items():
        torch.testing.assert_close(grads[k], g, rtol=1e-5, atol=1e-7)
    o = net_oracle.OracleNet(); o.load_state_dict(fx["state_dict"]); o.eval()
    with torch.no_grad():
        le = o(fx["x"])
    torch.testing.assert_close(le, fx["logp_eval"], rtol=1e-6, atol=1e-6)
    assert torch.equal(le.argmax(1), fx["argmax_eval"])


@pytest.mark.parametrize("name", ["train_sgd.pt", "train_adamw.pt", "train_adam.pt", "train_sgdm.pt", "train_adadelta.pt"])
def test_oracle_training_loop_matches_reference_golden(name):
    fx = load_golden(name)
    losses, sd = net_oracle.train_steps(fx["state_dict0"], fx["xs"], fx["ys"], fx["masks2d"], fx["masks1d"],
                                        fx["opt_kind"], **fx["opt_kw"])
    torch.testing.assert_close(losses[:, 0], fx["losses"], rtol=1e-6, atol=1e-7)
    for k, v in fx["state_dict"].items():
        torch.testing.assert_close(sd[k], v, rtol=1e-5, atol=1e-7)


def test_oracle_world2_matches_real_gloo_ddp_golden():
    fx = load_golden("ddp_gloo_w2.pt")
    _, sd = net_oracle.train_steps(fx["state_dict0"], fx["xs"], fx["ys"], fx["masks2d"], fx["masks1d"], "sgd",
                                   world=fx["world"], **fx["opt_kw"])
    for k, v in fx["state_dict"].items():
        torch.testing.assert_close(sd[k], v, rtol=1e-5, atol=1e-7)


def test_sampler_oracle_matches_golden_and_torch():
    fx = load_golden("sampler.pt")
    for c in fx["cases"]:
        idx = sampler_oracle.shard_indices(c["n"], c["world"], c["rank"], True, c["seed"], c["epoch"], c["drop_last"])
        assert len(idx) == c["len"]
        assert hashlib.sha256(idx.astype(np.int64).tobytes()).hexdigest() == c["sha256"], c
        assert np.array_equal(idx[:16], c["head"].numpy()[:len(idx)])
    from torch.utils.data.distributed import DistributedSampler
    for n, w, r, ep, seed, dl in [(1000, 3, 1, 5, 9, False), (1000, 3, 2, 0, 0, True), (17, 5, 4, 1, 2, False)]:
        s = DistributedSampler(range(n), num_replicas=w, rank=r, seed=seed, drop_last=dl)
        s.set_epoch(ep)
        assert list(s) == sampler_oracle.shard_indices(n, w, r, True, seed, ep, dl).tolist()
    s = DistributedSampler(range(10), num_replicas=4, rank=3, shuffle=False)
    assert list(s) == sampler_oracle.shard_indices(10, 4, 3, shuffle=False).tolist()


def test_randperm_oracle_bit_exact():
    for n, seed in [(10, 0), (625, 1), (1249, 7), (10000, 3)]:
        g = torch.Generator(); g.manual_seed(seed)
        assert np.array_equal(sampler_oracle.randperm(n, seed), torch.randperm(n, generator=g).numpy())


def test_synth_oracle_properties():
    x, y = synth_oracle.synth_batch(1234, np.arange(64))
    assert x.shape == (64, 1, 28, 28) and x.dtype == np.float32 and y.dtype == np.int64
    assert x.min() >= np.float32(-0.42421296) - 1e-6 and x.max() <= np.float32(2.8214867) + 1e-6
    px, _ = synth_oracle.synth_pixels(1234, np.arange(64))
    assert px.dtype == np.uint8 and px.max() <= 239
    assert set(np.unique(y)) <= set(range(10))
    x2, y2 = synth_oracle.synth_batch(1234, np.arange(10, 20))
    assert np.array_equal(x2, x[10:20]) and np.array_equal(y2, y[10:20])      # pure function of (seed, index)
    # golden fixtures were generated from this same stream
    fx = load_golden("fwdbwd_b8.pt")
    xg, yg = synth_oracle.synth_batch(fx["synth_seed"], np.arange(fx["synth_start"], fx["synth_start"] + 8))
    assert np.array_equal(xg, fx["x"].numpy()) and np.array_equal(yg, fx["y"].numpy())
    # Philox4x32-10 known-answer vectors (Random123 kat_vectors)
    r = synth_oracle.philox4x32_10([0], [0], [0], [0], 0, 0)
    assert [int(v[0]) for v in r] == [0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8]
    r = synth_oracle.philox4x32_10([0xffffffff], [0xffffffff], [0xffffffff], [0xffffffff], 0xffffffff, 0xffffffff)
    assert [int(v[0]) for v in r] == [0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd]
    r = synth_oracle.philox4x32_10([0x243f6a88], [0x85a308d3], [0x13198a2e], [0x03707344], 0xa4093822, 0x299f31d0)
    assert [int(v[0]) for v in r] == [0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1]


@pytest.mark.parametrize("name", ["mlp_h1_f128.pt", "mlp_h2_f64.pt"])
def test_mlp_oracle_matches_reference_golden(name):
    """oracle/mlp_oracle.py against the fixtures written from the reference's own Model / Trainer._run_batch
    (ref: pytorch_elastic/mnist_ddp_elastic.py:70-75,133-159)."""
    from oracle import mlp_oracle
    fx = load_golden(name)
    H, Fe = fx["hidden_layers"], fx["features"]
    logits, loss, grads = mlp_oracle.forward_backward(fx["state_dict0"], fx["xs"][0], fx["ys"][0], H, Fe)
    torch.testing.assert_close(logits, fx["logits"], rtol=1e-6, atol=1e-6)
    torch.testing.assert_close(loss, fx["loss"], rtol=1e-6, atol=1e-7)
    for k, g in fx["grads"].items():
        torch.testing.assert_close(grads[k], g, rtol=1e-5, atol=1e-8)
    losses, sd = mlp_oracle.train_steps(fx["state_dict0"], fx["xs"], fx["ys"], H, Fe, "adam", lr=1e-3)
    torch.testing.assert_close(losses[:, 0], fx["losses"], rtol=1e-6, atol=1e-7)
    for k, v in fx["state_dict"].items():
        torch.testing.assert_close(sd[k], v, rtol=1e-5, atol=1e-6)
    # W = 2 semantics: averaging two half-batch gradients == the full-batch gradient of the mean loss
    l2, sd2 = mlp_oracle.train_steps(fx["state_dict0"], fx["xs"][:1], fx["ys"][:1], H, Fe, "adam", world=2, lr=1e-3)
    l1, sd1 = mlp_oracle.train_steps(fx["state_dict0"], fx["xs"][:1], fx["ys"][:1], H, Fe, "adam", world=1, lr=1e-3)
    for k in sd1:
        assert (sd2[k] - sd1[k]).abs().max().item() <= 2.5e-4     # one Adam step moves every weight by <= lr; sign(g) agrees
