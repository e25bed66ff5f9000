"""Host-side pieces of the measurement contract that need no GPU: the W-rank gloo reference job (bench.py's
`cpu_baseline` leg and `--impl reference` arm, oracle/cpu_baseline.py), its rendezvous-port choice, and build()'s
content stamp.  The GPU arm of bench.py is exercised by the driver and by tools/evidence_n1.sh / evidence_n8.sh."""
import json
import os
import socket
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def test_rendezvous_port_is_free_and_below_the_ephemeral_range():
    from oracle import cpu_baseline
    seen = set()
    for _ in range(8):
        port = cpu_baseline._free_port()
        assert 15000 <= port < 29000          # outgoing connections of a GPU job draw from 32768+, never from here
        with socket.socket() as s:
            s.bind(("127.0.0.1", port))       # still free
        seen.add(port)
    assert len(seen) > 1                      # not a fixed port: two jobs on one box do not collide


def test_two_rank_gloo_reference_job_reports_global_throughput():
    from oracle import cpu_baseline
    r = cpu_baseline.time_reference_ddp(world=2, batch=32, steps=2, warmup=1, threads_per_rank=1, model="cnn",
                                        opt=dict(kind="sgd", lr=0.01, momentum=0.9), timeout=240.0)
    assert r["world"] == 2 and r["steps"] == 2 and r["batch"] == 32 and r["threads"] == 1
    assert r["samples_per_s"] > 0 and abs(r["samples_per_s"] - 2 * 32 * 2 / r["seconds"]) < 1e-6 * r["samples_per_s"]


def test_reference_arm_prints_one_json_line_with_the_contract_keys():
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="")
    for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK", "MASTER_ADDR", "MASTER_PORT"):
        env.pop(k, None)
    p = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "1", "--steps", "1",
                        "--warmup", "1", "--batch", "64"], cwd=ROOT, env=env, capture_output=True, text=True, timeout=600)
    assert p.returncode == 0, p.stderr[-2000:]
    lines = [l for l in p.stdout.splitlines() if l.strip()]
    assert len(lines) == 1, lines                       # stdout carries the JSON line only
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["unit"] == "samples/s" and d["higher_is_better"] is True and d["n_gpus"] == 1
    assert d["value"] > 0 and d["cpu_baseline"]["value"] == d["value"] and d["cpu_baseline"]["kind"] == "port"
    assert d["e2e"] == {"value": d["value"], "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert d["config"]["per_gpu_batch"] == 64 and "workload" in d["config"] and "model" not in d["config"]


def test_build_stamp_follows_content_not_modification_time(tmp_path):
    import __graft_entry__ as g
    src = tmp_path / "a.cu"
    src.write_text("// one\n")
    lib = str(tmp_path / "lib.so")
    d1 = g._digest([str(src)], ["-O3"])
    assert g._stale(lib, d1)                            # no library yet
    open(lib, "w").close()
    assert g._stale(lib, d1)                            # library without a stamp
    with open(lib + ".stamp", "w") as f:
        f.write(d1 + "\n")
    os.utime(src, (0, 0)); os.utime(lib, (1, 1))
    assert not g._stale(lib, g._digest([str(src)], ["-O3"]))      # any modification-time order: still current
    os.utime(src, (10 ** 9, 10 ** 9))
    assert not g._stale(lib, g._digest([str(src)], ["-O3"]))
    src.write_text("// two\n")
    assert g._stale(lib, g._digest([str(src)], ["-O3"]))          # content changed
    assert g._stale(lib, g._digest([str(src)], ["-O2"]))          # flags changed
