"""CPU suite: host-side mirror of the reference API, the C-ABI library's exports, the pair sharding (gloo, 2 ranks)."""
import ctypes
import os
import re
import warnings

import numpy as np
import pytest
import torch
import torch.multiprocessing as mp

from tests.util import ROOT

REF_KEYS_KITTI = 34


def test_library_exports_every_declared_symbol():
    from deepclr_b200 import _lib
    header = open(os.path.join(ROOT, 'include', 'deepclr_b200.h')).read()
    declared = sorted(set(re.findall(r'\b(dclr_[a-z_0-9]+)\s*\(', header)))
    assert len(declared) >= 14
    lib = ctypes.CDLL(_lib.LIB_PATH)
    for name in declared:
        assert hasattr(lib, name), f'{name} declared in include/deepclr_b200.h but not exported'
    assert set(_lib.EXPORTS) == set(declared)
    assert b'sm_100a' in _lib.lib().dclr_version()


def test_argument_errors_are_reported_not_fatal():
    from deepclr_b200 import _lib
    rc = _lib.lib().dclr_fps(None, 3, 1, 10, 4, 0, None, None, None)
    assert rc == -1 and b'null' in _lib.lib().dclr_last_error()
    rc = _lib.lib().dclr_knn(None, None, 1, 4, 4, 2, None, None, None)
    assert rc == -1


def test_package_default_tie_rule_is_the_reference_kernels():
    """Out of the box FPS orders exact ties like the reference's furthest_point_sampling_kernel; the C ABI
    has no process-wide setting any more."""
    import subprocess
    import sys
    out = subprocess.run([sys.executable, '-c', 'import deepclr_b200; print(deepclr_b200.get_fps_tie_rule())'],
                         capture_output=True, text=True, cwd=ROOT).stdout.strip()
    assert out == 'reference'
    from deepclr_b200 import _lib
    lib = ctypes.CDLL(_lib.LIB_PATH)
    assert not hasattr(lib, 'dclr_set_fps_tie_rule') and hasattr(lib, 'dclr_model_set_fps_tie_rule')
    assert _lib.lib().dclr_fps(ctypes.c_void_p(16), 3, 1, 10, 4, 7, ctypes.c_void_p(16), None, None) == -1
    assert b'tie_rule' in _lib.lib().dclr_last_error()


def test_no_cpu_route():
    from deepclr_b200.pointnet2 import furthest_point_sample
    with pytest.raises(RuntimeError, match='CUDA'):
        furthest_point_sample(torch.rand(1, 16, 3), 4)


def test_model_config_and_state_dict_layout():
    from deepclr_b200.config import load_model_config
    from deepclr_b200.data.labels import LabelType
    from deepclr_b200.models import ModelType, build_model
    cfg = load_model_config(os.path.join(ROOT, 'configs', 'kitti_model.yaml'), '')
    assert cfg.label_type == LabelType.POSE3D_DUAL_QUAT and cfg.model_type == ModelType.DEEPCLR
    assert cfg.is_frozen()
    model = build_model(cfg)
    sd = model.state_dict()
    assert len(sd) == REF_KEYS_KITTI and sum(v.numel() for v in sd.values()) == 1778312
    assert sd['_cloud_layers.0._sa0.mlps.1.layer2.conv.weight'].shape == (32, 16, 1, 1)
    assert sd['_merge_layers.0._embedding._conv._sequential.0._sequential.0.weight'].shape == (128, 131, 1)
    assert sd['_merge_layers.1.conv._sequential.4._sequential.0.weight'].shape == (1024, 512, 1)
    assert sd['_merge_layers.1.linear._sequential.1._sequential.0.weight'].shape == (256, 512)
    assert sd['_merge_layers.1.output.bias'].tolist() == [1, 0, 0, 0, 0, 0, 0, 0]
    assert model.get_input_dim() == 4 and not model.has_loss() and model._fused_supported()
    m2 = build_model(load_model_config(os.path.join(ROOT, 'configs', 'modelnet40_model.yaml'), ''))
    assert sum(v.numel() for v in m2.state_dict().values()) == 1778280
    with pytest.raises(RuntimeError, match='no CPU path'):
        model(torch.rand(2, 64, 4))


def test_inference_helper_contract():
    from deepclr_b200.models import ModelInferenceHelper
    from deepclr_b200.models.base import BaseModel

    class Fake(BaseModel):
        def get_input_dim(self): return 4
        def has_loss(self): return False
        def get_loss_weights(self): return {}
        def cloud_features(self, x, m=None): return x.sum(dim=1, keepdim=True)
        def forward(self, x, is_feat=False, m=None, y=None, debug=False):
            return x.reshape(1, -1)[:, :8] * (2.0 if is_feat else 1.0), None, None

    h = ModelInferenceHelper(Fake(), is_sequential=False)
    with pytest.raises(RuntimeError):
        h.predict(torch.rand(10, 4))
    with pytest.raises(RuntimeError):
        h.predict(torch.rand(10, 3), torch.rand(10, 4))
    with warnings.catch_warnings(record=True) as w:
        warnings.simplefilter('always')
        out = h.predict(torch.rand(12, 5), torch.rand(10, 4))
        assert len(w) == 1 and 'Truncate' in str(w[0].message)
    assert out.shape == (8,)
    assert ModelInferenceHelper.stack(torch.rand(7, 4), torch.rand(9, 4)).shape == (2, 7, 4)
    s = ModelInferenceHelper(Fake(), is_sequential=True)
    assert s.predict(torch.rand(10, 4)) is None and s.has_state()
    assert s.predict(torch.rand(10, 4)) is not None
    with pytest.raises(RuntimeError):
        s.predict(torch.rand(10, 4), torch.rand(10, 4))
    s.reset_state()
    assert not s.has_state()


def test_label_roundtrip():
    from deepclr_b200.data.labels import LabelType
    from deepclr_b200 import synth
    from oracle import model_oracle
    rng = np.random.default_rng(0)
    for _ in range(5):
        m = synth.random_motion(rng, 2.0, 30.0)
        for lt in LabelType:
            np.testing.assert_allclose(lt.to_matrix(lt.from_matrix(m)), m, atol=1e-9)
        dq = LabelType.POSE3D_DUAL_QUAT.from_matrix(m)
        np.testing.assert_allclose(model_oracle.dual_quat_to_matrix(dq), m, atol=1e-7)
    assert LabelType.create('pose3d_dual_quat') is LabelType.POSE3D_DUAL_QUAT


def test_factory_and_config():
    from deepclr_b200.config import Config
    from deepclr_b200.models.deepclr import DeepCLRModule, OutputSimple
    from deepclr_b200.utils.factory import factory
    from deepclr_b200.data.labels import LabelType
    mod = factory(DeepCLRModule, 'OutputSimple', input_dim=19, label_type=LabelType.POSE3D_QUAT, mlp=[8], linear=[8, 4],
                  dropout=1.0)
    assert isinstance(mod, OutputSimple) and mod.output_dim() == 7
    with pytest.raises(NotImplementedError):
        factory(DeepCLRModule, 'Nope')
    c = Config()
    c.define_param('a', default=1, required=True)
    with pytest.raises(KeyError):
        c.b = 2
    c.freeze()
    with pytest.raises(RuntimeError):
        c.a = 3


def _worker(rank, world, port, q):
    import torch.distributed as dist
    from deepclr_b200 import sharding, synth
    dist.init_process_group('gloo', init_method=f'tcp://127.0.0.1:{port}', rank=rank, world_size=world)
    x = synth.batch('modelnet', 5, 0, n=32)
    local = sharding.shard_batch(x, world, rank)
    b, e = sharding.pair_range(5, world, rank)
    # stand-in for the per-pair result: checksum of (template, source) -- what the host gathers after the forward
    half = local.shape[0] // 2
    res = torch.tensor([[float(local[i].sum()), float(local[half + i].sum())] for i in range(half)])
    sizes = [torch.zeros(1, dtype=torch.long) for _ in range(world)]
    dist.all_gather(sizes, torch.tensor([half]))
    parts = [torch.zeros(int(s), 2) for s in sizes]
    if rank == 0:
        parts[0] = res
        for r in range(1, world):
            dist.recv(parts[r], src=r)
        q.put((torch.cat(parts).numpy(), x))
    else:
        dist.send(res, dst=0)
    dist.barrier()
    dist.destroy_process_group()
    assert e - b == half


def test_pair_sharding_two_ranks_gloo():
    from deepclr_b200 import sharding
    assert [sharding.pair_range(64, 8, r) for r in (0, 7)] == [(0, 8), (56, 64)]
    assert [sharding.pair_range(5, 2, r) for r in (0, 1)] == [(0, 3), (3, 5)]
    ctx = mp.get_context('spawn')
    q = ctx.SimpleQueue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got, x = q.get()
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    want = np.array([[x[i].sum(), x[5 + i].sum()] for i in range(5)], dtype=np.float32)
    np.testing.assert_allclose(got, want, rtol=1e-6)
