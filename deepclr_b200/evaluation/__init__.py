from .data import Motion, Sequence

__all__ = ['Motion', 'Sequence']
