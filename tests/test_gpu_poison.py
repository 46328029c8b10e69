"""Uninitialised-memory check of the CUDA path: the parity cases of tests/poison_worker.py with every workspace and every
torch.empty buffer pre-filled with a byte pattern.  Stands in for compute-sanitizer's initcheck, which is not available
on the GPU pool.  One subprocess per pattern (the library reads B2S_POISON once per process)."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("pattern", [0x00, 0xff, 0x7f, 0xa5, "random"])
def test_results_do_not_depend_on_uninitialised_memory(pattern):
    out = subprocess.run([sys.executable, os.path.join(ROOT, "tests", "poison_worker.py")], capture_output=True, text=True,
                         timeout=900, env=dict(os.environ, B2S_POISON=str(pattern)))
    assert "POISON_OK" in out.stdout, (out.stdout[-2000:], out.stderr[-3000:])
