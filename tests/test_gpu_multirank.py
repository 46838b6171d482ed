"""Two ranks on two GPUs of one box: the partitioned product path (ghost exchange over peer-memory windows, p2p = 1, or
NCCL send/recv, p2p = 0) against the one-rank run (skipped with < 2 GPUs)."""
import os
import socket
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.gpu
@pytest.mark.parametrize("p2p,tuning", [("1", ""), ("1", "exchange_groups=0"), ("0", "")])
def test_two_gpus_ghost_exchange(p2p, tuning):
    """p2p = 1: peer-memory windows; exchange_groups = 0: every exchange carries every point class instead of the classes written since
    the last one.  p2p = 0: NCCL send / recv."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    procs = []
    for r in range(2):
        env = dict(os.environ, RANK=str(r), LOCAL_RANK=str(r), WORLD_SIZE="2", MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), HB_P2P=p2p, HB_TUNING=tuning)
        procs.append(subprocess.Popen([sys.executable, os.path.join(ROOT, "tests", "multirank_gpu_worker.py")], env=env,
                                      stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True))
    outs = []
    for p in procs:
        try:
            out, _ = p.communicate(timeout=900)
        except subprocess.TimeoutExpired:
            for q in procs:
                q.kill()
            raise
        outs.append(out)
    for r, (p, out) in enumerate(zip(procs, outs)):
        assert p.returncode == 0, f"rank {r} failed:\n{out[-3000:]}"
        assert f"OK rank {r}/2" in out
