"""absl flags with the reference's names and defaults (src/util/common_flags.py:11-95 and the driver-local
flags of src/maml_pde.py:50-58, src/leap_pde.py:42-48, src/nn_pde_maml.py:43-45).  ``define_*`` are idempotent
so that several compat drivers can be imported in one process (the reference defines them at import time of
each driver, which makes its drivers mutually exclusive per process)."""
from absl import flags

FLAGS = flags.FLAGS


def _def(kind, name, default, help_=""):
    if name in FLAGS:
        return
    getattr(flags, "DEFINE_" + kind)(name, default, help_)


def define_common():
    _def("string", "out_dir", None, "out directory")
    _def("string", "expt_name", "default", "expt name")
    _def("float", "bc_scale", 1.0, "scale on random uniform bc")
    _def("float", "bc_weight", 100.0, "weight on bc loss")
    _def("float", "relaxation_parameter", 0.2, "Newton solver relaxation parameter (FEniCS side channel; unused)")
    _def("integer", "max_newton_steps", 500, "Newton solver max steps (FEniCS side channel; unused)")
    _def("string", "pde", "poisson", "which PDE")
    _def("integer", "ground_truth_resolution", 16, "mesh resolution for fenics ground truth (unused: no FEniCS)")
    _def("float", "boundary_resolution_factor", 3.0, "ratio of resolution to points around boundary of shape")
    _def("integer", "fixed_num_pdes", None, "fixed number of pdes")
    _def("integer", "n_eval", 16, "num eval PDEs")
    _def("integer", "validation_points", 1024, "num points in domain for validation")
    for n in ("vary_source", "vary_bc", "vary_geometry", "vary_ic"):
        _def("boolean", n, True, "")
    _def("string", "domain_loss", None, "")
    _def("float", "tmin", 0.0, "PDE initial time")
    _def("float", "tmax", 1.0, "PDE final time")
    _def("integer", "num_tsteps", 101, "number of time steps for td_burgers")
    _def("integer", "sample_tsteps", 64, "number of time steps for td_burgers")
    _def("boolean", "sample_time_random", True, "random time sample for NN")
    _def("boolean", "gpu_samplers", False, "metapde_b200 only: draw the collocation points on the GPU (same keys, same points; poisson and td_burgers)")
    _def("float", "max_reynolds", 100, "Reynolds number scale")
    _def("string", "burgers_pde", "default", "types of burgers equation")
    _def("float", "xmin", 0.0, "")
    _def("float", "xmax", 1.0, "")
    _def("float", "ymin", -1.0, "")
    _def("float", "ymax", 1.0, "")
    _def("integer", "max_holes", 12, "")
    _def("float", "max_hole_size", 0.4, "")
    _def("integer", "seed", 0, "set random seed")
    _def("integer", "outer_points", 256, "num support points on the boundary and in domain")
    _def("integer", "inner_points", 256, "num support points on the boundary and in domain")
    _def("integer", "outer_steps", int(1e8), "num outer steps")
    _def("integer", "num_layers", 3, "num fcnn layers")
    _def("integer", "layer_size", 64, "fcnn layer size")
    _def("boolean", "siren", True, "use siren")
    _def("float", "grad_clip", 1e14, "max grad for clipping")
    _def("float", "siren_omega", 1.0, "siren_omega")
    _def("float", "siren_omega0", 3.0, "siren_omega0")
    _def("boolean", "log_scale", True, "io_scale")
    _def("float", "io_scale_lr_factor", 1e1, "scale io lr by this factor")
    _def("string", "optimizer", "ranger", "adam or ranger, currently no adahess")
    _def("integer", "viz_every", int(1e4), "plot every N steps (plots are out of scope: checkpoints only)")
    _def("integer", "val_every", int(1e2), "validate every N steps")
    _def("integer", "log_every", int(1e3), "log every N steps")
    _def("integer", "measure_grad_norm_every", int(1e3), "")
    _def("string", "load_model_from_expt", None, "load pre-trained model")


def define_maml():
    define_common()
    _def("integer", "bsize", 16, "batch size (in tasks)")
    _def("float", "outer_lr", 1e-3, "outer learning rate")
    _def("float", "inner_lr", 3e-5, "inner learning rate")
    _def("float", "lr_inner_lr", 1.0 / 2, "lr for inner learning rate")
    _def("integer", "inner_steps", 5, "num_inner_steps")
    _def("float", "inner_grad_clip", 1e14, "inner grad clipping")
    _def("float", "outer_loss_decay", 0.1, "0. = just take final loss. 1.=sum all")


def define_leap():
    define_common()
    _def("integer", "bsize", 16, "batch size (in tasks)")
    _def("float", "outer_lr", 1e-3, "outer learning rate")
    _def("float", "inner_lr", 3e-5, "inner learning rate")
    _def("float", "inner_grad_clip", 1e14, "inner grad clipping")
    _def("integer", "inner_steps", 10, "num_inner_steps")


def define_nn():
    define_common()
    _def("integer", "bsize", 1, "batch size (in tasks)")
    _def("float", "outer_lr", 1e-5, "outer learning rate")
    _def("float", "inner_grad_clip", 100.0, "inner grad clipping")


def ensure_parsed(argv=("compat",)):
    if not FLAGS.is_parsed():
        FLAGS(list(argv))
