from .config import Config, ConfigEnum
from .utils import Mode, create_default_config, load_config, load_model_config

__all__ = ['Config', 'ConfigEnum', 'Mode', 'create_default_config', 'load_config', 'load_model_config']
