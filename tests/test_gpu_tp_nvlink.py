"""The tensor-parallel decoder on REAL ranks (one process per GPU, NCCL + torch symmetric memory over NVLink): fused
exchanges against NCCL collectives against the unsharded decoder, eager and from a CUDA graph. Needs >= 2 GPUs (skipped on
the one-GPU test box; run with `gpurun --gpus 2`). The same kernels are covered on one GPU with emulated ranks in
tests/test_gpu_tp.py."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("world", [2, 4, 8])
def test_tp_decoder_on_real_ranks_is_bit_identical(world):
    if torch.cuda.device_count() < world:
        pytest.skip("needs %d GPUs" % world)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world), "--master-addr",
           "127.0.0.1", "--master-port", str(29540 + world), os.path.join(ROOT, "tests", "tp_nvlink_worker.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    if r.returncode == 0 and "TP_NVLINK_UNAVAILABLE" in r.stdout:
        pytest.skip(r.stdout.strip().splitlines()[-1])
    assert r.returncode == 0 and "TP_NVLINK_OK" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]
