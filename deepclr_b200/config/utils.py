"""Model-side configuration loading with the reference's entry points (deepclr/config/utils.py:18-24, 232-256).
Only the `model` group is interpreted; the training/data/optimizer schema of the reference is not rebuilt and
unknown top-level keys are kept as dynamic parameters.  Unlike the reference, loading needs neither a git
checkout (utils.py:192-195) nor the data-loader / optimizer packages."""
from enum import auto
from typing import Optional

from .config import Config, ConfigEnum


class Mode(ConfigEnum):
    NEW = auto()
    CONTINUE = auto()
    INFERENCE = auto()
    TEST = auto()


def create_default_config(mode: Mode) -> Config:
    cfg = Config(allow_dynamic_params=True)
    cfg.add_internal_param('mode', value=mode)
    cfg.define_param('device', default='cuda')
    cfg.define_param('checkpoint')
    model = cfg.define_group('model')
    cfg.define_param('weights', parent=model, default=None)
    cfg.define_param('input_dim', parent=model, required=True)
    cfg.define_param('point_dim', parent=model, default=3)
    cfg.define_param('label_type', parent=model, required=True)
    cfg.define_param('model_type', parent=model, required=True)
    cfg.define_param('params', parent=model, default={})
    return cfg


def finish_config(cfg: Config) -> None:
    from ..data.labels import LabelType
    from ..models.build import ModelType
    cfg.model.label_type = LabelType.create(cfg.model.label_type)
    cfg.model.model_type = ModelType.create(cfg.model.model_type)
    if cfg.model.point_dim > cfg.model.input_dim:
        raise RuntimeError("Model input dimension must be equal or smaller than point dimension.")
    cfg.freeze()


def load_config(cfg_filename: str, mode: Mode, ckpt_filename: Optional[str] = None) -> Config:
    config = create_default_config(mode=mode)
    config.read_file(cfg_filename)
    if ckpt_filename is not None:
        config.checkpoint = ckpt_filename
    finish_config(config)
    if not config.is_valid():
        raise RuntimeError("Configuration is not valid, missing required parameters.")
    return config


def load_model_config(cfg_filename: str, weights_filename: str) -> Config:
    """Load configuration for the model only (a reference model_config.yaml)."""
    config = create_default_config(mode=Mode.INFERENCE)
    config.model.read_file(cfg_filename)
    config.model.weights = weights_filename
    finish_config(config)
    return config.model
