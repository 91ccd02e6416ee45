"""CPU: bench.py's contract where no GPU is visible.  The product arm must refuse (there is no CPU path to fall back to); the
reference arm must still print its JSON line -- the CPU port of the reference algorithm on the host cores, a bounded sample -- with
the keys the driver reads.  The numbers themselves are the -m gpu runs' business (profiles/)."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _gpu_visible():
    return os.path.exists("/dev/nvidiactl")        # no torch import at collection time


pytestmark = pytest.mark.skipif(_gpu_visible(), reason="a GPU is visible: these are the no-GPU behaviours")


def run(*args):
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="")
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py")] + list(args), capture_output=True, text=True, timeout=600, env=env)


def test_product_arm_refuses_without_a_gpu():
    p = run("--steps", "1", "--warmup", "1", "--no-legs")
    assert p.returncode != 0
    assert "no CPU path" in p.stderr and p.stdout.strip() == ""


def test_reference_arm_prints_the_contract_line_from_the_host_cores():
    p = run("--impl", "reference", "--steps", "1", "--warmup", "1")
    assert p.returncode == 0, p.stderr[-500:]
    lines = [l for l in p.stdout.splitlines() if l.strip()]
    assert len(lines) == 1                                  # ONE JSON line on stdout, nothing else
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["unit"] == "rows/s" and d["higher_is_better"] is True and d["n_gpus"] == 1
    assert d["metric"].startswith("cudasp_scan rows/sec") and "workload" in d["config"]
    assert d["value"] > 0 and d["steps"] == 1 and d["warmup"] == 1
    cb = d["cpu_baseline"]
    assert cb["kind"] in ("port", "reference") and cb["cores"] >= 1 and cb["value"] == d["value"] and cb["sample"]
    assert d["e2e"] == {"value": d["value"], "unit": "rows/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
