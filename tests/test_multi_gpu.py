"""Multi-GPU tests (need >= 2 visible B200s; skipped on a single-GPU box): the peer-memory collectives against NCCL
(reference op being replaced: /root/reference/mpi_tools.py:30-37,47-53) and the data-parallel step check of bench.py."""
import json
import os
import socket
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _torchrun(n, args, timeout=300):
    with socket.socket(socket.AF_INET, socket.SOCK_STREAM) as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(n), "--master-addr",
           "127.0.0.1", "--master-port", str(port)] + args
    env = {k: v for k, v in os.environ.items() if k not in ("RANK", "WORLD_SIZE", "LOCAL_RANK", "MASTER_PORT")}
    p = subprocess.Popen(cmd, cwd=ROOT, env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True,
                         start_new_session=True)
    try:
        out, err = p.communicate(timeout=timeout)
    except subprocess.TimeoutExpired:
        os.killpg(p.pid, 9)                      # exactly the process group this test started
        raise
    return p.returncode, out, err


def _world():
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs at least 2 GPUs")
    return 2 if n < 4 else (4 if n < 8 else 8)


def test_peer_collectives_match_nccl():
    n = _world()
    rc, out, err = _torchrun(n, ["scripts/test_peer.py"])
    assert rc == 0, err[-2000:]
    res = json.loads([l for l in out.splitlines() if l.startswith("{")][-1])
    assert res["ok"] and res["world"] == n
    for c in res["checks"]:
        assert c["max_err"] < (2e-2 if c["op"] == "avg_bf16" else 1e-5), c


def test_data_parallel_step_equals_nccl_plus_torch_adam():
    n = _world()
    rc, out, err = _torchrun(n, ["bench.py", "--gpus", str(n), "--steps", "3", "--warmup", "3", "--no-e2e", "--model",
                                 "resnet18", "--batch", "64", "--image", "128"])
    assert rc == 0, err[-2000:]
    line = json.loads([l for l in out.splitlines() if l.startswith("{")][-1])
    chk = line["dp_check"]
    assert chk["param_digest_equal_across_ranks"] and chk["bf16_shadow_exact"]
    assert chk["vs_nccl_max_abs"] < 1e-5 and chk["update_max_abs"] > 1e-5, chk
    assert line["n_gpus"] == n and line["config"]["library_fallbacks"] == {}
