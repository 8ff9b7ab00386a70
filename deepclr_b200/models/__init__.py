from .base import BaseModel, ModelInferenceHelper
from .stream import GraphedForward, GraphedSequentialStep, HostPipeline, RawScanPipeline
from .build import build_model, load_trained_model, ModelType, store_models_code

__all__ = ['BaseModel', 'ModelInferenceHelper', 'GraphedForward', 'GraphedSequentialStep', 'HostPipeline', 'RawScanPipeline',
           'build_model', 'load_trained_model', 'ModelType', 'store_models_code']
