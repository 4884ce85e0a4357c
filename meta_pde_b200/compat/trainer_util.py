"""The parts of src/util/trainer_util.py the path needs: prepare_logging (text log only: TensorBoard / matplotlib are
out of scope and absent), get_optimizer, vmap_validation_error / vmap_validation_rollout (:501-590, ground-truth values
supplied by the caller: FEniCS is out of scope); plus src/util/timer.py's Timer with a device synchronise."""
import os
import shutil
import time

import torch

from .optim import get_optimizer  # noqa: F401


def vmap_validation_error(*args, **kwargs):
    from .deploy import vmap_validation_error as f
    return f(*args, **kwargs)


def vmap_validation_rollout(*args, **kwargs):
    from .deploy import vmap_validation_rollout as f
    return f(*args, **kwargs)


def prepare_logging(out_dir, expt_name):
    """src/util/trainer_util.py:473-498 (deletes an existing expt dir, as the reference does); tflogger is None."""
    if expt_name is not None:
        os.makedirs(out_dir, exist_ok=True)
        path = os.path.join(out_dir, expt_name)
        if os.path.exists(path):
            shutil.rmtree(path)
        os.mkdir(path)
        outfile = open(os.path.join(path, "log.txt"), "w")

        def log(*args, **kwargs):
            print(*args, **kwargs, flush=True)
            print(*args, **kwargs, file=outfile, flush=True)
    else:
        path = out_dir

        def log(*args, **kwargs):
            print(*args, **kwargs, flush=True)
    return path, log, None


class Timer:
    """src/util/timer.py:5-27; synchronises the device on exit so that the interval covers the kernels."""

    def __enter__(self):
        self.start = time.perf_counter()
        return self

    def __exit__(self, *args):
        if torch.cuda.is_available():
            torch.cuda.synchronize()
        self.end = time.perf_counter()
        self.interval = self.end - self.start
