"""The real multi-rank exchange under pytest: needs at least two GPUs (skipped otherwise; the world-size-2 gloo test in
test_shard_gloo.py covers the shard arithmetic on CPU).  Launches tests/mp_exchange_check.py under torch.distributed.run."""
import os
import subprocess
import sys

import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("world", [2, 4, 8])
def test_exchange_variants_full_c_vs_oracle_on_every_rank(world):
    if not torch.cuda.is_available() or torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
           "--master-port", str(29540 + world), os.path.join(ROOT, "tests", "mp_exchange_check.py")]
    out = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=600)
    ok = out.stdout.count("MP_EXCHANGE_OK")
    assert out.returncode == 0 and ok == world, out.stdout[-3000:] + out.stderr[-3000:]
