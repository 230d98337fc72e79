"""bench.py --impl reference on a CPU-runnable configuration (BASELINE configs[0]-like, Si2): the JSON line carries the
keys the contract names; under torchrun only rank 0 prints; OMP_NUM_THREADS=1 (what torchrun exports) does not pin the
GEMMs of the reference arm to one core."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(env_extra):
    env = dict(os.environ)
    env.update(env_extra)
    p = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--workload", "si2",
                        "--steps", "1", "--warmup", "1"], capture_output=True, text=True, env=env, cwd=ROOT, timeout=600)
    assert p.returncode == 0, p.stderr[-2000:]
    return p.stdout.strip().splitlines()


def test_reference_arm_line():
    lines = _run({"OMP_NUM_THREADS": "1"})
    assert len(lines) == 1
    j = json.loads(lines[0])
    assert j["impl"] == "reference"
    for k in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "config",
              "cpu_baseline", "e2e"):
        assert k in j
    assert j["metric"] == "h_psi band-applications/s" and j["value"] > 0
    assert j["cpu_baseline"]["kind"] in ("reference", "port") and j["cpu_baseline"]["cores"] >= 1
    assert j["e2e"]["h2d_bytes_per_step"] == 0 and j["e2e"]["d2h_bytes_per_step"] == 0
    assert j["gpu_launches"] == 0
    assert j["maps_libqeb200"] is False  # the reference arm never loads the product library
    # both arms print the same `config` keys for the same command line (the driver compares the dicts)
    assert sorted(j["config"]) == ["bands_per_step", "gamma_only", "grid", "l2", "nkb", "npw", "parallelism", "workload"]


def test_reference_arm_other_ranks_are_silent():
    assert _run({"RANK": "1", "WORLD_SIZE": "2", "LOCAL_RANK": "1"}) == []
