"""build_model / load_trained_model with the reference's signatures (deepclr/models/build.py:14-49)."""
from __future__ import annotations

import os
import os.path as osp
import shutil
from enum import auto
from typing import Type

import torch

from ..config.config import Config, ConfigEnum
from .base import BaseModel


class ModelType(ConfigEnum):
    DEEPCLR = auto()

    def get_class(self) -> Type[BaseModel]:
        if self == ModelType.DEEPCLR:
            from .deepclr import DeepCLR
            return DeepCLR
        raise NotImplementedError("ModelType not implemented")


def build_model(model_cfg: Config) -> BaseModel:
    model_cls = model_cfg.model_type.get_class()
    return model_cls(input_dim=model_cfg.input_dim, point_dim=model_cfg.point_dim, label_type=model_cfg.label_type,
                     **model_cfg.params)


def store_models_code(directory: str) -> None:
    here = osp.dirname(osp.realpath(__file__))
    os.mkdir(directory)
    for name in os.listdir(here):
        if osp.isfile(osp.join(here, name)):
            shutil.copyfile(osp.join(here, name), osp.join(directory, name))


def load_trained_model(model_cfg: Config) -> BaseModel:
    """Build the model and load a reference `weights.tar` (a plain state dict, deepclr/utils/checkpoint.py:97-99)."""
    model = build_model(model_cfg)
    model.load_state_dict(torch.load(model_cfg.weights, map_location='cpu'))
    return model
