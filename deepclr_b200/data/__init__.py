from .labels import LabelType

__all__ = ['LabelType']
