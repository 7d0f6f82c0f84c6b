"""The evaluation-mode forward passes with two tiles in flight per SM (csrc/acx_mlp_tc2.cu, off by default because it
measured slower -- profiles/r02_forward_experiments.md): the whole tensor-core forward parity file once more with
ACX_TC_TWO_TILES=1.  The switch is read when the library first launches such a pass, hence a process of its own."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_two_tile_forward_kernels_pass_the_forward_parity_file():
    env = dict(os.environ, ACX_TC_TWO_TILES="1")
    out = subprocess.run([sys.executable, "-m", "pytest", os.path.join(ROOT, "tests", "test_gpu_mlp_tc.py"), "-m", "gpu", "-q",
                          "-x", "-p", "no:cacheprovider"], cwd=ROOT, env=env, capture_output=True, text=True, timeout=900)
    assert out.returncode == 0, out.stdout[-3000:] + out.stderr[-2000:]
    assert " passed" in out.stdout
