"""bench.py's contract on the CPU side: the reference arm prints exactly ONE JSON line on stdout (whatever a library writes
to fd 1 goes to stderr), with the keys the driver reads; under torchrun only rank 0 runs it."""
import json
import os
import subprocess
import sys

import common


def _run(extra_env=None):
    env = dict(os.environ)
    env.update(extra_env or {})
    return subprocess.run([sys.executable, os.path.join(common.ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0"],
                          capture_output=True, text=True, timeout=600, env=env, cwd=common.ROOT)


def test_reference_arm_prints_one_json_line():
    r = _run()
    assert r.returncode == 0, r.stderr
    lines = [ln for ln in r.stdout.splitlines() if ln.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "camera_paths_per_second" and d["unit"] == "paths/s"
    assert d["higher_is_better"] is True and d["value"] > 0 and d["gpu_launches"] == 0
    assert d["cpu_baseline"]["kind"] in ("reference", "port") and d["cpu_baseline"]["cores"] >= 1
    assert d["cpu_baseline"]["value"] == d["value"] == d["e2e"]["value"]
    assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["e2e"]["d2h_bytes_per_step"] == 0
    assert d["config"]["workload"].startswith("C4")


def test_reference_arm_other_ranks_exit_quietly():
    r = _run({"RANK": "1", "WORLD_SIZE": "2", "LOCAL_RANK": "1"})
    assert r.returncode == 0, r.stderr
    assert r.stdout.strip() == ""


def test_library_noise_on_fd1_does_not_reach_stdout():
    code = ("import os, sys; sys.path.insert(0, %r); import bench; bench.OUT = bench.claim_stdout(); "
            "os.write(1, b'NCCL version 0.0\\n'); print('python print'); bench.emit({'ok': 1})" % common.ROOT)
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stderr
    assert r.stdout == '{"ok": 1}\n'
    assert "NCCL version 0.0" in r.stderr and "python print" in r.stderr
