"""`bench.py --impl reference`: the unmodified reference script (a byte copy under baseline/_ref) runs through its offline
stand-ins and prints the bench line.  CPU plumbing run with a tiny shape (FEDB200_REF_DEVICE=cpu); the timed GPU run is the
driver's."""
import hashlib
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, "baseline", "_ref")


@pytest.mark.skipif(not (os.path.exists(os.path.join(REF, "SHA256SUMS")) or os.path.isdir("/root/reference/src")),
                    reason="the reference tree is neither staged under baseline/_ref nor mounted at /root/reference")
def test_reference_arm_runs_the_unmodified_script_and_prints_one_json_line():
    env = dict(os.environ, FEDB200_REF_DEVICE="cpu")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--model", "bert-tiny", "--seq-len", "32",
                        "--train-per-client", "8", "--test-per-client", "8", "--batch-size", "4", "--steps", "1", "--warmup", "3"],
                       env=env, capture_output=True, text=True, timeout=900, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.strip()]
    assert len(lines) == 1, lines
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and "unavailable" not in d
    assert d["value"] > 0 and d["steps"] == 1 and d["warmup"] == 3 and d["dtype"] == "fp32"
    # the script that ran is byte-identical to the staged copy's manifest entry
    script = os.path.join(REF, d["config"]["script"])
    assert hashlib.sha256(open(script, "rb").read()).hexdigest() == d["config"]["sha256"]
    sums = open(os.path.join(REF, "SHA256SUMS")).read()
    assert d["config"]["sha256"] in sums
