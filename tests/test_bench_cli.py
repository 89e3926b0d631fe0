"""bench.py contract checks that do not need a GPU."""
import json
import os
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU behaviour")
def test_reference_arm_reports_unavailable_without_gpu():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference"],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-500:]
    line = json.loads(r.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and "unavailable" in line


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU behaviour")
def test_product_arm_refuses_to_run_without_gpu():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py")], capture_output=True,
                       text=True, timeout=300)
    assert r.returncode != 0 and "CUDA device" in (r.stderr + r.stdout)


def test_reference_install_is_byte_identical_when_present():
    ref = os.path.join(ROOT, "baseline", "_ref", "main.py")
    src = "/root/reference/main.py"
    if not (os.path.exists(ref) and os.path.exists(src)):
        pytest.skip("baseline/_ref or /root/reference not present on this box")
    assert open(ref, "rb").read() == open(src, "rb").read()


def test_graft_entry_exposes_build_and_smoke():
    sys.path.insert(0, ROOT)
    import __graft_entry__ as g

    assert callable(g.build) and callable(g.smoke)
