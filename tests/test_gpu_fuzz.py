"""Randomised differential checks of the kernel dispatch (tools/fuzz_paths.py, tools/fuzz_groups.py): the long-range kernels
against short calls (round-1 layouts) and oracle samples over random lengths, steps, shift ranges and group geometries."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("tool,cases,seed", [("fuzz_paths.py", 60, 101), ("fuzz_groups.py", 40, 102), ("fuzz_mt.py", 10, 103)])
def test_dispatch_fuzz(tool, cases, seed):
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", tool), str(cases), str(seed)], cwd=ROOT, capture_output=True,
                       text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-2000:]
    assert "0 bad" in r.stdout
