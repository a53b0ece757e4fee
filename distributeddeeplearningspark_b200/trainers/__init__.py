from .engine import StepEngine, categorical_crossentropy, mean_squared_error, get_loss
from .parameter_server import CollectiveCenter, MulticastCenter, ParameterServerView
from .trainers import (Trainer, SingleTrainer, DistributedTrainer, ADAG, DynSGD, DOWNPOUR, AEASGD, EAMSGD, AveragingTrainer,
                       EnsembleTrainer, AllreduceTrainer,
                       stage_partitions, load_partition)
from .predictors import (Predictor, ModelPredictor, Evaluator, AccuracyEvaluator, get_MAPE, evaluate_accuracy)

__all__ = ['StepEngine', 'categorical_crossentropy', 'mean_squared_error', 'get_loss', 'CollectiveCenter', 'MulticastCenter',
           'ParameterServerView', 'Trainer', 'SingleTrainer', 'DistributedTrainer', 'ADAG', 'DynSGD', 'DOWNPOUR', 'AEASGD', 'EAMSGD', 'AveragingTrainer', 'EnsembleTrainer',
           'AllreduceTrainer', 'stage_partitions', 'load_partition', 'Predictor', 'ModelPredictor', 'Evaluator',
           'AccuracyEvaluator', 'get_MAPE', 'evaluate_accuracy']
