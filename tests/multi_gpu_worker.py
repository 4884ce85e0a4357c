"""Worker of tests/test_gpu_multi.py (launched by torchrun, one rank per GPU): one MAML outer step on this rank's share
of a fixed global meta-batch; rank 0 saves the all-reduced packed gradient and the updated parameters."""
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT]
from meta_pde_b200 import synthetic, train  # noqa: E402


def main(out_path, W, L, n, T, K):
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    theta0 = synthetic.siren_theta([W] * L, 1, 30.0, 30.0, True, seed=0, bias_scale=0.05)
    tr = train.MamlTrainer("poisson", L, W, theta0, bsize=T, inner_steps=K, inner_lr=1e-2, outer_lr=1e-3, lr_inner_lr=0.5,
                           grad_clip=1e-3, device=dev)
    T_local = T // world
    res = []
    for step in range(2):
        batch = train.synthetic_maml_batch("poisson", T_local, K, n, n, seed=5 + step, task_offset=rank * T_local, T_global=T)
        losses, meta, aux, gnorm = tr.train_step(step, batch.to(dev))
        res.append(tr._packed.clone())
    torch.cuda.synchronize()
    if rank == 0:
        torch.save({"packed": [r.cpu() for r in res], "theta": tr.theta.cpu(), "lrs": tr.inner_lrs.cpu(), "gnorm": gnorm.cpu()}, out_path)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main(sys.argv[1], *(int(x) for x in sys.argv[2:7]))
