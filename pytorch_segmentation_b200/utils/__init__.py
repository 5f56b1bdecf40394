from .utils import compute_loss, compute_metrics

__all__ = ['compute_loss', 'compute_metrics']
