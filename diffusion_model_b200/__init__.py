"""B200-native (sm_100a) implementation of the EGNN-diffusion hot path of Ren-Okubo/diffusion_model.

The sub-modules keep the reference's module and symbol names so that ``main.py`` / ``train.py`` /
``generate_amorphous.py`` can import from here unchanged:
    EquivariantGraphNeuralNetwork.{EGCL, EquivariantGNN}
    diffusion_x_h.{E3DiffusionProcess, remove_mean}      E3diffusion_new.{E3DiffusionProcess, remove_mean}
    train_per_iretation.{diffuse_as_batch, train_epoch, eval_epoch, generate, EarlyStopping}
All compute goes through libegnn_b200.so (include/egnn_b200.h); there is no CPU or PyTorch fallback.
"""
from ._lib import EgnnError, LIB_PATH, load  # noqa: F401

__version__ = "0.1.0"
