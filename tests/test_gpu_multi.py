"""Hardware proof that the task-sharded N-GPU outer step equals the 1-GPU one (SURVEY.md section 4 item 4, 7.2 step 10:
"meta-grad equal to 1-GPU within 1e-6 rel"; the mean over tasks of src/nets/maml.py:231 becomes local sums + one NCCL
all-reduce).  Needs >= 2 GPUs: skipped on a 1-GPU box (run with `gpurun --gpus 2`); the CPU twin over gloo and the
emulator is tests/test_compat_host.py::test_task_sharding_world2_matches_single."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(world, out, shape):
    worker = os.path.join(ROOT, "tests", "multi_gpu_worker.py")
    if world == 1:
        cmd = [sys.executable, worker, out] + [str(x) for x in shape]
    else:
        port = 29600 + os.getpid() % 300
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
               "--master-addr", "127.0.0.1", "--master-port", str(port), worker, out] + [str(x) for x in shape]
    subprocess.run(cmd, check=True, timeout=600)
    return torch.load(out)


@pytest.mark.parametrize("shape", [(64, 3, 256, 4, 2), (128, 2, 200, 4, 2)])     # FP32-FMA path and tensor-core path
def test_two_gpus_equal_one_gpu(tmp_path, shape):
    if not torch.cuda.is_available() or torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    a = _run(1, str(tmp_path / "n1.pt"), shape)
    b = _run(2, str(tmp_path / "n2.pt"), shape)
    rel = lambda x, y: float((x.double() - y.double()).norm() / y.double().norm())
    # step 0: the same theta on both sides -> the all-reduced mean gradient agrees to the summation order
    assert rel(b["packed"][0], a["packed"][0]) <= 1e-6, rel(b["packed"][0], a["packed"][0])
    # step 1 starts from the Adam-updated theta: lr * g / (|g| + eps) is ill-conditioned where the task contributions cancel
    # (sign flips of entries at the rounding level move those parameters by 2 lr), so the bound is the propagated one
    assert rel(b["packed"][1], a["packed"][1]) <= 1e-3, rel(b["packed"][1], a["packed"][1])
    assert rel(b["gnorm"], a["gnorm"]) <= 1e-3
    assert float((b["theta"] - a["theta"]).abs().max()) <= 4.1e-3      # two steps of at most lr = 1e-3 each way
