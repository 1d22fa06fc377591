"""Multi-GPU parity on hardware (SURVEY §4 tier 4, §8e): NCCL, one process per GPU.  Skipped on a one-GPU box;
`gpurun --gpus 2 -- python -m pytest tests/test_gpu_multi.py` runs it and the log is committed under profiles/."""
import json
import os
import socket
import subprocess
import sys

import pytest
import torch

from conftest import ROOT, report

pytestmark = pytest.mark.gpu


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("world", [2])
def test_nccl_allreduced_training_gradients_and_sharded_inference(world, tmp_path):
    if torch.cuda.device_count() < world:
        pytest.skip("needs %d GPUs" % world)
    out = tmp_path / "nccl.json"
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world),
           "--master-addr", "127.0.0.1", "--master-port", str(_free_port()),
           os.path.join(ROOT, "tests", "nccl_worker.py"), str(out)]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    res = json.load(open(out))
    report("nccl_world%d" % world, **res)
    assert res["world"] == world and res["tensors"] > 60
    assert res["allreduce_grad_max_rel_err"] <= 1e-5, res
    assert res["params_identical_after_step"] is True
    assert res["sharded_inference_bit_identical"] is True
