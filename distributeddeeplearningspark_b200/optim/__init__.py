from .flat import FlatSpace
from .keras_optimizers import KerasAdam, KerasAdagrad, KerasSGD, get_optimizer

__all__ = ['FlatSpace', 'KerasAdam', 'KerasAdagrad', 'KerasSGD', 'get_optimizer']
