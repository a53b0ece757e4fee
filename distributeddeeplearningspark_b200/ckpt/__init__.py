from .keras_npz import save_weights, load_weights, save_checkpoint, load_checkpoint, resume

__all__ = ['save_weights', 'load_weights', 'save_checkpoint', 'load_checkpoint', 'resume']
