"""The first half-sweeps after a random initialisation score the cost matrix by lane pairs (sweep_kernel<..., kPairMatrix>,
capi_run.cu.inc: cold_sweeps): a scheduling choice that must not change a single bit.  ACMMP_COLD_SWEEPS is read once per
process, so every setting runs in a process of its own.  pytest -m gpu."""
import json
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")
if not torch.cuda.is_available():
    pytest.skip("no CUDA device", allow_module_level=True)

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def one_view_crc(cold, tex):
    env = dict(os.environ, ACMMP_COLD_SWEEPS=str(cold))
    out = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "run_one_view.py"), "--width", "640", "--height", "480", "--nsrc", "4",
                          "--runs", "1", "--tex", tex], env=env, capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr[-2000:]
    line = [l for l in out.stdout.splitlines() if l.startswith("{")][-1]
    return json.loads(line)["crc"]


@pytest.mark.parametrize("tex", ["auto", "r32f"])
def test_results_do_not_depend_on_the_cold_sweep_count(tex):
    crcs = {k: one_view_crc(k, tex) for k in (0, 1, 6)}
    assert len(set(crcs.values())) == 1, crcs
