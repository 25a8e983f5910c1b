"""bench.py's JSON line contract, on the arm that runs without a GPU: `--impl reference` (the CPU arm: the
reference's own render from oracle/_ref or /root/reference, else the oracle port).  The GPU arm's line is
produced on the box by the driver."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(*extra, env=None):
    e = dict(os.environ)
    e.update(env or {})
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1",
                          "--warmup", "1", "--cpu-res", "8", *extra], capture_output=True, text=True, timeout=600,
                         cwd=ROOT, env=e)
    assert out.returncode == 0, out.stderr[-2000:]
    return out.stdout.strip().splitlines()


def test_reference_arm_prints_one_contract_line():
    lines = _run()
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "rays/sec (fwd+bwd)" and d["unit"] == "rays/s"
    assert d["higher_is_better"] is True and d["n_gpus"] == 1 and d["steps"] == 1 and d["warmup"] == 1
    assert d["value"] > 0 and d["ms_per_step"] > 0 and d["vs_baseline"] is None and d["data"] == "synthetic"
    assert "workload" in d["config"] and "model" not in d["config"]
    cb = d["cpu_baseline"]
    have_ref = os.path.exists("/root/reference/lib/nerf_utils.py") or \
        os.path.exists(os.path.join(ROOT, "oracle", "_ref", "lib", "nerf_utils.py"))
    assert cb["kind"] == ("reference" if have_ref else "port") and cb["cores"] >= 1 and cb["value"] == d["value"] and cb["sample"]
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}


def test_reference_arm_other_ranks_exit_quietly():
    """Under torchrun (N > 1) rank 0 alone runs and prints the CPU arm; the other ranks exit 0 without work."""
    assert _run("--gpus", "2", env={"RANK": "1", "WORLD_SIZE": "2", "LOCAL_RANK": "1"}) == []


def test_every_config_runs_on_the_reference_arm():
    for cid in ("1", "3s", "3c", "4", "5"):
        lines = _run("--config", cid)
        d = json.loads(lines[-1])
        assert d["config"]["id"] == cid and d["value"] > 0
        assert d["metric"] == ("rays/sec (fwd)" if cid == "1" else "rays/sec (fwd+bwd)")
        assert d["scaling"] == ("strong" if cid.startswith("3") else "weak")
