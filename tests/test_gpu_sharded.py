"""SURVEY.md section 8(e) on a GPU: the key-frame-sharded path with its one exchange step, two ranks.  On the
single-GPU test box both ranks share cuda:0 and the halo maps are staged over gloo; with >= 2 GPUs the
same worker runs one rank per GPU and the maps travel through the C ABI's NCCL exchange (sd_halo_exchange).  Each rank compares the key frames it owns with the
oracle run on the whole sequence (bit-exact)."""
import os
import socket
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _run(world, backend, extra=None):
    port = _free_port()
    procs = []
    for r in range(world):
        env = dict(os.environ, RANK=str(r), WORLD_SIZE=str(world), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port),
                   SHARD_BACKEND=backend, **(extra or {}))
        procs.append(subprocess.Popen([sys.executable, os.path.join(HERE, "shard_worker.py")], env=env,
                                      stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True))
    outs = [p.communicate(timeout=900)[0] for p in procs]
    for r, (p, o) in enumerate(zip(procs, outs)):
        assert p.returncode == 0, f"rank {r} failed:\n{o[-3000:]}"
        assert "mismatching planes 0" in o
    return outs


def test_two_ranks_one_gpu_gloo_staging(built):
    _run(2, "gloo")


def test_three_ranks_uneven_split(built):
    _run(3, "gloo", {"SHARD_NKF": "16", "SHARD_NBR": "5"})


def test_two_ranks_nccl(built):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs (run with gpurun --gpus 2)")
    _run(2, "nccl")
