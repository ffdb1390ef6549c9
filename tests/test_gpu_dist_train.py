"""Launches tests/dist_train_check.py on 2 GPUs (NCCL) when the box has them; see that file."""

import subprocess
import sys
from pathlib import Path

import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs (run with gpurun --gpus 2)")
def test_data_parallel_training_two_gpus():
    script = Path(__file__).with_name("dist_train_check.py")
    res = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                          "--master-addr", "127.0.0.1", "--master-port", "29533", str(script)],
                         capture_output=True, text=True, timeout=600)
    assert res.returncode == 0 and "DIST_TRAIN_OK" in res.stdout, res.stdout[-2000:] + res.stderr[-2000:]
