"""snapshot.pt interchange — host logic only, no GPU: the file written here has the reference's
keys and CPU tensors, so the reference's ``_load_snapshot`` body (ref: pytorch_elastic/mnist_ddp_elastic.py:61-68)
reads it; the optional optimizer state rides along as an extra key the reference ignores."""
import os

import torch

from oracle import mlp_oracle
from pytorch_distributed_examples_b200.snapshot import load_snapshot, save_snapshot


def test_snapshot_round_trip_and_reference_reader(tmp_path):
    torch.manual_seed(0)
    model = mlp_oracle.OracleMLP(1, 32)
    opt = torch.optim.Adam(model.parameters(), lr=1e-3)
    x, y = torch.randn(8, 784), torch.randint(0, 10, (8,))
    torch.nn.functional.cross_entropy(model(x), y).backward()
    opt.step()
    path = os.path.join(tmp_path, "snapshot.pt")
    save_snapshot(model, 7, path, optimizer=opt)
    # the reference's reader, restated (ref :61-68): torch.load, MODEL_STATE -> load_state_dict, EPOCHS_RUN
    snap = torch.load(path)
    assert set(snap) >= {"MODEL_STATE", "EPOCHS_RUN"} and snap["EPOCHS_RUN"] == 7
    assert all(not v.is_cuda for v in snap["MODEL_STATE"].values())
    fresh = mlp_oracle.OracleMLP(1, 32)
    fresh.load_state_dict(snap["MODEL_STATE"])
    for a, b in zip(fresh.parameters(), model.parameters()):
        assert torch.equal(a, b)
    # our reader restores the optimizer state too
    m2 = mlp_oracle.OracleMLP(1, 32)
    o2 = torch.optim.Adam(m2.parameters(), lr=1e-3)
    assert load_snapshot(m2, path, optimizer=o2) == 7
    s1, s2 = opt.state_dict()["state"], o2.state_dict()["state"]
    assert s1.keys() == s2.keys()
    for k in s1:
        assert torch.equal(s1[k]["exp_avg"], s2[k]["exp_avg"]) and torch.equal(s1[k]["exp_avg_sq"], s2[k]["exp_avg_sq"])
    # a file written the reference's way (no optimizer key) loads as well
    torch.save({"MODEL_STATE": model.state_dict(), "EPOCHS_RUN": 3}, path)
    assert load_snapshot(m2, path, optimizer=o2) == 3


def test_async_writer_falls_back_to_a_synchronous_write_for_cpu_modules(tmp_path):
    from pytorch_distributed_examples_b200.snapshot import AsyncSnapshotWriter
    torch.manual_seed(1)
    model = mlp_oracle.OracleMLP(1, 16)
    path = os.path.join(tmp_path, "snapshot.pt")
    w = AsyncSnapshotWriter(path)
    w.save(model, 4)
    w.flush()
    snap = torch.load(path)
    assert snap["EPOCHS_RUN"] == 4 and all(torch.equal(snap["MODEL_STATE"][k], v) for k, v in model.state_dict().items())
    w.close()
