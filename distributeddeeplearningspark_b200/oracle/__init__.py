from . import keras_math
from .keras_math import *  # noqa: F401,F403
