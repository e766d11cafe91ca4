"""GPU (>= 2 devices): the partitioned solve -- halo exchange and all-reduce over peer memory (and over
NCCL with FK_NO_P2P=1) -- against a single-GPU solve of the same box.  Skipped on one-GPU boxes."""
import json
import os
import socket
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _ngpu():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:  # noqa: BLE001
        return 0


def _port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.mark.gpu
@pytest.mark.parametrize("no_p2p,rec,halo", [("", 0, "2"), ("", 0, "1"), ("", 1, ""), ("1", 0, "")])
def test_partitioned_solve_matches_single_gpu(no_p2p, rec, halo):
    n = _ngpu()
    if n < 2:
        pytest.skip("needs at least 2 GPUs")
    world = 2 if n < 4 else 4
    env = dict(os.environ, FK_NO_P2P=no_p2p, FK_RECURRENCE=str(rec), FK_HALO_MODE=halo)
    if not no_p2p:
        env.pop("FK_NO_P2P")
    if not halo:
        env.pop("FK_HALO_MODE")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
           "--master-port", str(_port()), os.path.join(ROOT, "scripts", "mgpu_check.py"), "16"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env, cwd=ROOT)
    lines = [ln for ln in r.stdout.splitlines() if ln.startswith("{")]
    assert r.returncode == 0 and lines, r.stdout[-2000:] + r.stderr[-4000:]
    out = json.loads(lines[-1])
    assert out["ok"] and out["comm_mode"] == (1 if no_p2p else 2), out
