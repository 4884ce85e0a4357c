"""The optimizers the reference takes from Flax 0.3.3 / jax.experimental (third-party, restated; see oracle/meta.py):
``Momentum(learning_rate, beta).create(model)`` / ``Adam(learning_rate, beta1, beta2).create(model)`` return an
``Optimizer`` with ``.target`` and ``.apply_gradient(grad_model)``; ``get_optimizer`` mirrors
src/util/trainer_util.py:593-612.  Updates run in the parameter-sized kernels of libmetapde_b200.so."""
import torch

from .. import ops
from .field import Model
from .flags import FLAGS


class Optimizer:
    def __init__(self, optimizer_def, target, state):
        self.optimizer_def, self.target, self.state = optimizer_def, target, state

    def apply_gradient(self, grads):
        return self.optimizer_def.apply_gradient(self, grads)


class Momentum:
    def __init__(self, learning_rate=None, beta=0.9):
        if beta != 0.0:
            raise NotImplementedError("metapde_b200: Momentum is implemented for beta = 0 (the reference's only use)")
        self.learning_rate, self.beta = float(learning_rate), beta

    def create(self, target):
        return Optimizer(self, target, {"step": 0})

    def apply_gradient(self, opt, grads):
        if grads.theta.requires_grad or opt.target.theta.requires_grad:     # differentiable inner step (generic MAML path)
            return Optimizer(self, opt.target.tree_map(lambda p, g: p - self.learning_rate * g, grads), {"step": opt.state["step"] + 1})
        K = ops.kernels(opt.target.theta.device)
        x = opt.target.theta.detach().clone().contiguous()
        K.sgd_step(x.view(-1), grads.theta.detach().contiguous().view(-1), self.learning_rate)
        return Optimizer(self, Model(opt.target.module, x), {"step": opt.state["step"] + 1})


class Adam:
    def __init__(self, learning_rate=None, beta1=0.9, beta2=0.999, eps=1e-8):
        self.learning_rate, self.beta1, self.beta2, self.eps = float(learning_rate), beta1, beta2, eps

    def create(self, target):
        z = torch.zeros_like(target.theta)
        return Optimizer(self, target, {"step": 0, "m": z, "v": z.clone()})

    def apply_gradient(self, opt, grads):
        K = ops.kernels(opt.target.theta.device)
        x = opt.target.theta.detach().clone().contiguous()
        m, v = opt.state["m"].clone(), opt.state["v"].clone()
        t = opt.state["step"] + 1
        K.adam_step(x.view(-1), grads.theta.detach().contiguous().view(-1), m.view(-1), v.view(-1), t, self.learning_rate,
                    self.beta1, self.beta2, self.eps)
        return Optimizer(self, Model(opt.target.module, x), {"step": t, "m": m, "v": v})


def get_optimizer(model_class, init_params):
    """src/util/trainer_util.py:593-612"""
    model = init_params if isinstance(init_params, Model) else Model(model_class, init_params)
    if FLAGS.optimizer == "adam":
        return Adam(learning_rate=FLAGS.outer_lr, beta1=0.9, beta2=0.99).create(model)
    elif FLAGS.optimizer == "rmsprop":
        return Adam(learning_rate=FLAGS.outer_lr, beta1=0.0, beta2=0.8).create(model)
    elif FLAGS.optimizer == "ranger":
        raise Exception("optimizer 'ranger' needs the external, un-pinned flaxOptimizers package (not restated); "
                        "every published command passes --optimizer adam")
    elif FLAGS.optimizer == "sgd":
        return Momentum(learning_rate=FLAGS.outer_lr, beta=0.0).create(model)
    else:
        raise Exception("unknown optimizer: ", FLAGS.optimizer)
