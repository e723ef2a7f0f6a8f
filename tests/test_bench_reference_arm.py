"""bench.py --impl reference (the CPU arm the driver runs next to the GPU arm): one JSON line with the contract's keys, the
same `config` shape as the GPU arm, parity of the CPU prover's five points against the trapdoor identity, no CUDA and no
libg16b200.so involved.  Small sizes so that the CPU suite stays short; both curves."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("curve,log_n", [("bn254", 11), ("bls12_377", 10)])
def test_reference_arm_prints_one_contract_line(curve, log_n):
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="")          # the arm must not need a GPU
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--curve", curve,
                          "--log-n", str(log_n), "--steps", "2", "--warmup", "1"], capture_output=True, text=True, env=env, timeout=600)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [l for l in out.stdout.splitlines() if l.strip()]
    assert len(lines) == 1, out.stdout
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == f"groth16_{curve}_prove_ms" and d["unit"] == "ms"
    assert d["higher_is_better"] is False and d["steps"] == 2 and d["value"] > 0 and d["ms_per_step"] == d["value"]
    assert d["config"]["constraints"] == 1 << log_n and d["config"]["workload"].startswith(f"groth16_{curve}_prove_2p{log_n}")
    assert d["parity"]["mismatches"] == 0 and d["parity"]["checked"] == ["ar", "bs1", "krs_k", "krs_hz", "bs2"]
    assert d["cpu_baseline"]["kind"] == "port" and d["cpu_baseline"]["cores"] >= 1 and d["cpu_baseline"]["value"] == d["value"]
    assert d["e2e"] == {"value": d["value"], "unit": "ms", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}


def test_reference_arm_idle_ranks_exit_quietly():
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="", RANK="1", WORLD_SIZE="2", LOCAL_RANK="1")
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--log-n", "10",
                          "--steps", "1", "--warmup", "0"], capture_output=True, text=True, env=env, timeout=120)
    assert out.returncode == 0 and out.stdout.strip() == ""
