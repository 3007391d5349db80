"""B200-native NIC-ES / NIC-NES population-evaluation hot path (drop-in for the reference's
Policy / PolicyNet / NESMaster / Optimizer API on that path).  Hand-written sm_100a CUDA behind a
C ABI (include/nes_b200.h); no CPU fallback."""
from . import _lib, layout
from ._lib import NesError
from .context import NesContext, PREC_EXACT, PREC_FAST, PREC_HYBRID, OPT_ADAM, OPT_SGD, gemm_selftest
from .nets import FCModel, MnistNet, ModelOptions, Mutation, PolicyNet
from .policies import CaptPolicy, ClfPolicy, Net, Policy, PolicyFactory, PolicyOptions, SuppDataset
from .cider import CiderD
from . import schedule
from .master import (Adam, ESMaster, NESMaster, Optimizer, SGD, all_gather_fitness, all_reduce_gradient, sharded_master_step, ESPopulation, es_generation, es_generation_mnist,
                     nes_iteration, shard_bounds)

__all__ = ['NesContext', 'NesError', 'FCModel', 'MnistNet', 'ClfPolicy', 'ModelOptions', 'Mutation', 'PolicyNet', 'CaptPolicy', 'Net', 'Policy',
           'PolicyFactory', 'PolicyOptions', 'SuppDataset', 'Adam', 'SGD', 'Optimizer', 'NESMaster', 'nes_iteration',
           'shard_bounds', 'all_gather_fitness', 'all_reduce_gradient', 'sharded_master_step', 'ESMaster', 'ESPopulation', 'es_generation', 'es_generation_mnist', 'CiderD', 'PREC_EXACT', 'PREC_FAST', 'PREC_HYBRID', 'OPT_ADAM', 'OPT_SGD', 'gemm_selftest']
