"""Shared helpers for the parity tests."""
import os

import numpy as np
import torch
import yaml

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CONFIG_DIR = os.path.join(ROOT, 'configs')


def load_golden(name):
    return np.load(os.path.join(ROOT, 'tests', 'golden', name + '.npz'), allow_pickle=False)


def build_from_golden(g, device=None):
    """deepclr_b200 model + raw config dict for a golden case; weights re-created from the recorded seed and checked
    against the recorded per-tensor checksums (the reference's initialisers are reproduced draw for draw)."""
    from deepclr_b200.config.config import Config
    from deepclr_b200.data.labels import LabelType
    from deepclr_b200.models.build import ModelType, build_model
    cfg = yaml.safe_load(str(g['cfg_yaml']))
    c = Config._from_dict(cfg)
    c.label_type = LabelType.create(c.label_type)
    c.model_type = ModelType.create(c.model_type)
    torch.manual_seed(int(g['weight_seed']))
    model = build_model(c).eval()
    sums = np.array([[float(v.double().sum()), float(v.double().abs().sum())] for v in model.state_dict().values()])
    np.testing.assert_allclose(sums, g['weight_checksums'], rtol=1e-12, atol=1e-12)
    if device is not None:
        model = model.to(device)
    return model, cfg


def golden_input(g):
    from deepclr_b200 import synth
    x = synth.batch(str(g['kind']), int(g['pairs']), int(g['seed0']), n=int(g['n']))
    chk = np.array([x.astype(np.float64).sum(), np.abs(x).astype(np.float64).sum()])
    np.testing.assert_allclose(chk, g['x_checksum'], rtol=1e-12)
    return x
