"""The reference's ad-hoc checkpoint format (src/maml_pde.py:126-159,435-439; src/leap_pde.py:112-128,378-381;
src/nn_pde_maml.py:258-275,588-591): a pickle of ``flax.serialization.to_state_dict(Model)`` =
``{'params': {'Dense_i': {'kernel', 'bias'}, 'log_input_scale', 'log_output_scale'}}`` with numpy leaves;
MAML saves ``[model_state, inner_lrs_state]`` (lr leaves ``[K, *shape]``), LEAP / nn_pde save the model state alone.
Optimizer moments are not saved (neither does the reference)."""
import os
import pickle
import re

import numpy as np

from .field import Model


def to_state_dict(model):
    def conv(t):
        return {k: conv(v) for k, v in t.items()} if isinstance(t, dict) else v_np(t)

    def v_np(x):
        return x.detach().cpu().numpy().copy()
    return {"params": conv(model.params)}


def from_state_dict(module, state, device=None):
    return Model(module, state["params"], device=device)


def save(path, step, model, inner_lrs=None):
    obj = to_state_dict(model) if inner_lrs is None else [to_state_dict(model), to_state_dict(inner_lrs)]
    fn = os.path.join(path, "model_step_{}.pickle".format(step))
    with open(fn, "wb") as f:
        pickle.dump(obj, f, protocol=pickle.HIGHEST_PROTOCOL)
    return fn


def latest(model_path):
    """The file with the highest step number in ``--load_model_from_expt`` (src/maml_pde.py:126-136)."""
    files = [f for f in os.listdir(model_path) if "model_step" in f]
    if not files:
        raise FileNotFoundError(f"no model_step_*.pickle in {model_path}")
    steps = [int(re.findall("[0-9]+", f)[-1]) for f in files]
    return os.path.join(model_path, files[int(np.argsort(steps)[-1])])


def load(model_path, module, device=None):
    """-> (model, inner_lrs or None)"""
    with open(latest(model_path), "rb") as f:
        obj = pickle.load(f)
    if isinstance(obj, (list, tuple)):
        return from_state_dict(module, obj[0], device), from_state_dict(module, obj[1], device)
    return from_state_dict(module, obj, device), None
