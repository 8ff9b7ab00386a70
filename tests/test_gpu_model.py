"""GPU parity of the fused forward path against the oracle / reference goldens, through the reference-facing API
(deepclr_b200.models) which calls the C ABI.  Tolerances:
  fp32 mode : features max-abs-err <= 1e-4 * max|ref|; pose <= 0.01 deg and 1e-3 m (KITTI) / 1e-4 (ModelNet)
  fp16 mode : features <= 5e-3 * max|ref|;             pose <= 0.2 deg and 1e-2 m / 1e-3   (tcgen05, fp16 operands)
  bf16 mode : features <= 2e-2 * max|ref|;             pose <= 0.6 deg and 5e-2 m / 5e-3   (tcgen05, bf16 operands)
The 16-bit modes accumulate in fp32.  What the pose tolerances mean (tools/precision_class.py ->
profiles/r2_precision_class.json, 16 full-size KITTI-shaped pairs, seeded random-init weights, all against the reference
composition in strict fp32): the REFERENCE ITSELF with torch's default TF32 convolutions (cudnn.allow_tf32 = True, the
default of its pinned torch 1.8.1) deviates by 0.069 deg max / 0.046 mean and 3.6e-3 m (0.102 deg with TF32 matmuls too);
fp16 operands carry the same 11-bit significand as TF32 and land in the same place: 0.093 deg max / 0.045 mean, 4.9e-3 m
(0.101 deg on one of another 16 pairs).  bf16 (8-bit significand): 0.41 deg max / 0.27 mean, 3.2e-2 m.  fp32 mode: 6e-5
deg.  So the fp16 bound is 2x the reference's own TF32 self-deviation, not a tighter claim than the arithmetic supports;
SURVEY 8c's proposal of 0.1 deg for the 16-bit mode is met on average, not in the maximum.
Indices (FPS, ball-query counts, kNN sets) are bit-exact in every mode."""
import os
import warnings

import numpy as np
import pytest
import torch

from oracle import model_oracle
from tests.util import ROOT, build_from_golden, golden_input, load_golden

pytestmark = pytest.mark.gpu
DEV = 'cuda:0'

TOL = {'fp32': dict(feat=1e-4, rot=0.01, trans={'kitti': 1e-3, 'modelnet': 1e-4}),
       'fp16': dict(feat=5e-3, rot=0.2, trans={'kitti': 1e-2, 'modelnet': 1e-3}),
       'bf16': dict(feat=2e-2, rot=0.6, trans={'kitti': 5e-2, 'modelnet': 5e-3})}


def _modes():
    return ['fp32', 'fp16', 'bf16'] if os.environ.get('DCLR_TEST_TC', '1') == '1' else ['fp32']


def _check_against(g, model, x_np, mode, idx_exact=True):
    kind = str(g['kind'])
    model.numerics = mode
    x = torch.from_numpy(x_np).to(DEV)
    with torch.no_grad():
        feat = model.cloud_features(x)
        y, _, _ = model(x)
        native = model._native
        _, _, motion = native.forward_pairs(x, want_motion=True)
        y2, _, _ = model(feat, is_feat=True)
    B, M = x.shape[0] // 2, native.npoint
    fps = native.debug_tensor(0, (2 * B, M)).cpu().numpy()
    cnt0 = native.debug_tensor(1, (2 * B, M)).cpu().numpy()
    cnt1 = native.debug_tensor(2, (2 * B, M)).cpu().numpy()
    knn = native.debug_tensor(3, (B, M, native.knn_k)).cpu().numpy()
    assert np.array_equal(fps, g['fps_idx'])
    assert np.array_equal(cnt0, g['ball_cnt0']) and np.array_equal(cnt1, g['ball_cnt1'])
    want_knn = g['knn_idx'][1].reshape(B, M, -1) - (np.arange(B) * M)[:, None, None]
    assert np.array_equal(knn, want_knn)
    tol = TOL[mode]
    fe = np.abs(feat.cpu().numpy() - g['cloud_feat']).max() / np.abs(g['cloud_feat'][:, 3:]).max()
    me = np.abs(motion.cpu().numpy() - g['motion_feat']).max() / np.abs(g['motion_feat'][:, 3:]).max()
    assert np.array_equal(feat.cpu().numpy()[:, :3], g['cloud_feat'][:, :3])     # sampled xyz rows are copies
    assert fe <= 1e-4, f'SA features rel err {fe}'                               # SA is fp32 in both modes
    assert me <= tol['feat'], f'motion features rel err {me} ({mode})'
    yc = y.cpu().numpy()
    np.testing.assert_allclose(y2.cpu().numpy(), yc, atol=1e-6)
    for b in range(B):
        rot, trans = model_oracle.pose_errors(yc[b], g['y'][b])
        assert rot <= tol['rot'] and trans <= tol['trans'][kind], f'pose err {rot} deg {trans} m ({mode})'
    return fe, me


@pytest.mark.parametrize('case', ['modelnet_b2', 'kitti_small_b2', 'kitti_b1'])
def test_forward_matches_reference_goldens(case):
    g = load_golden(case)
    model, _ = build_from_golden(g, DEV)
    x = golden_input(g)
    for mode in _modes():
        _check_against(g, model, x, mode)


def test_forward_matches_live_oracle_ragged_sizes():
    """Fresh seeds, odd sizes (N not a multiple of anything, N < npoint for one case) vs the oracle run on the spot."""
    from deepclr_b200 import synth
    g = load_golden('modelnet_b2')
    model, cfg = build_from_golden(g, DEV)
    sd = {k: v.cpu() for k, v in model.state_dict().items()}
    for n, seed in ((700, 3), (333, 4)):
        x_np = synth.batch('modelnet', 2, seed, n=n)
        dbg = {}
        y_ref = model_oracle.forward(sd, cfg, torch.from_numpy(x_np), debug=dbg).numpy()
        live = dict(kind=np.array('modelnet'), fps_idx=dbg['fps_idx'], ball_cnt0=dbg['ball_cnt0'],
                    ball_cnt1=dbg['ball_cnt1'], knn_idx=dbg['knn_idx'], cloud_feat=dbg['cloud_feat'],
                    motion_feat=dbg['motion_feat'], y=y_ref)
        for mode in _modes():
            _check_against(live, model, x_np, mode)


def test_reference_shape_tests():
    """The reference's own test cases (tests/model/test_deepclr.py:17-57): 96-point clouds, npoint 1024 > N,
    module-by-module shapes and whole-model forward incl. the in-model loss and is_feat=True."""
    from deepclr_b200.config import load_model_config
    from deepclr_b200.models import build_model
    from deepclr_b200.models.deepclr import SetAbstraction, MotionEmbedding, OutputSimple
    cfg = load_model_config(os.path.join(ROOT, 'tests', 'model_test.yaml'), '')
    clouds_internal = torch.rand(10, cfg.input_dim, 96).to(DEV)
    sa = SetAbstraction(input_dim=cfg.input_dim, point_dim=cfg.point_dim, **cfg.params.cloud_features.params).to(DEV)
    f = sa(clouds_internal)
    assert f.shape == (10, 67, 1024)
    me = MotionEmbedding(input_dim=sa.output_dim(), point_dim=cfg.point_dim, **cfg.params.merge.params).to(DEV)
    e = me(f)
    assert e.shape == (5, 259, 1024)
    out = OutputSimple(input_dim=me.output_dim(), label_type=cfg.label_type, **cfg.params.output.params).to(DEV)
    assert out(e).shape == (5, cfg.label_type.dim)

    clouds = torch.rand(10, 96, cfg.input_dim).to(DEV)
    y = torch.rand(5, cfg.label_type.dim).to(DEV)
    model = build_model(cfg).to(DEV)
    model.numerics = 'fp32'                  # compared with the composed (torch fp32) route below
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    with torch.no_grad():
        y1, loss1, _ = model(clouds, y=y)
        assert y1.shape == (5, 8) and loss1.shape == ()
        feat = model.cloud_features(clouds)
        y2, loss2, _ = model(feat, y=y, is_feat=True)
        assert y2.shape == (5, 8) and loss2.shape == ()
        np.testing.assert_allclose(y1.cpu().numpy(), y2.cpu().numpy(), atol=1e-6)
    # composed route (grad enabled) == fused route
    yc, lossc, _ = model(clouds, y=y)
    assert yc.requires_grad
    np.testing.assert_allclose(yc.detach().cpu().numpy(), y1.cpu().numpy(), atol=2e-4)
    lossc.backward()


def test_inference_helper_sequential_equals_pairwise():
    from deepclr_b200 import synth
    from deepclr_b200.models import ModelInferenceHelper
    g = load_golden('kitti_small_b2')
    model, _ = build_from_golden(g, DEV)
    t, s, _ = synth.kitti_pair(0, 2048)
    t, s = torch.from_numpy(t).to(DEV), torch.from_numpy(s).to(DEV)
    pair = ModelInferenceHelper(model, is_sequential=False)
    y_pair = pair.predict(s, t)
    seq = ModelInferenceHelper(model, is_sequential=True)
    assert seq.predict(t) is None
    y_seq = seq.predict(s)
    np.testing.assert_allclose(y_seq.cpu().numpy(), y_pair.cpu().numpy(), atol=1e-6)
    with warnings.catch_warnings(record=True) as w:
        warnings.simplefilter('always')
        wide = torch.cat([s, s[:, :1]], dim=1)
        y_wide = pair.predict(wide, t)
        assert any('Truncate' in str(m.message) for m in w)
    np.testing.assert_allclose(y_wide.cpu().numpy(), y_pair.cpu().numpy(), atol=1e-6)


def test_batch_independence_full_size():
    """KITTI-shaped 16384-point pairs at batch 4: each pair's pose equals its batch-1 result (pairs do not interact),
    which is what makes the multi-GPU sharding exact."""
    from deepclr_b200 import sharding, synth
    g = load_golden('kitti_b1')
    model, _ = build_from_golden(g, DEV)
    x = synth.batch('kitti', 4, 200)
    with torch.no_grad():
        y_all, _, _ = model(torch.from_numpy(x).to(DEV))
        parts = [model(torch.from_numpy(sharding.shard_batch(x, 4, r)).to(DEV))[0] for r in range(4)]
    np.testing.assert_array_equal(torch.cat(parts).cpu().numpy(), y_all.cpu().numpy())
    assert torch.isfinite(y_all).all()


def test_host_pipeline_equals_direct_forward():
    """HostPipeline (double-buffered H2D on a copy stream, poses into pinned buffers) returns, batch by batch, exactly
    what a direct forward of the same batch returns -- 7 batches through 2 device slots, ragged last batch."""
    from deepclr_b200 import synth
    from deepclr_b200.models import HostPipeline
    g = load_golden('kitti_small_b2')
    model, _ = build_from_golden(g, DEV)
    hosts = [torch.from_numpy(synth.batch('kitti', 3 if i < 6 else 1, 300 + 10 * i, n=2048)).pin_memory()
             for i in range(7)]
    pipe = HostPipeline(model, depth=2)
    got = pipe.run_all(hosts)
    assert len(got) == 7 and pipe.h2d_bytes == sum(h.numel() * 4 for h in hosts)
    for h, y in zip(hosts, got):
        with torch.no_grad():
            want = model(h.to(DEV))[0].cpu()
        np.testing.assert_array_equal(y.numpy(), want.numpy())


def test_graphed_forward_replays_the_same_poses():
    """The fused forward (incl. the side-stream fork/join of the grid build) is CUDA-graph capturable: replays give
    bit-identical poses to direct calls, for new inputs copied into the static buffer."""
    from deepclr_b200 import synth
    from deepclr_b200.models import GraphedForward
    g = load_golden('kitti_small_b2')
    model, _ = build_from_golden(g, DEV)
    xs = [torch.from_numpy(synth.batch('kitti', 1, 300 + i, n=int(g['n']))).to(DEV) for i in range(3)]
    gf = GraphedForward(model, xs[0])
    for x in xs + xs[::-1]:
        with torch.no_grad():
            want = model(x)[0].clone()
        got = gf(x).clone()
        assert torch.equal(got, want)
    with pytest.raises(ValueError):
        gf(torch.cat([xs[0], xs[1]]))


def test_inference_helper_with_cuda_graph():
    from deepclr_b200 import synth
    from deepclr_b200.models import ModelInferenceHelper
    g = load_golden('kitti_small_b2')
    model, _ = build_from_golden(g, DEV)
    plain, graphed = ModelInferenceHelper(model), ModelInferenceHelper(model, use_cuda_graph=True)
    for i in range(3):
        x = torch.from_numpy(synth.batch('kitti', 1, 400 + i, n=int(g['n']))).to(DEV)
        assert torch.equal(plain.predict(x[1], x[0]), graphed.predict(x[1], x[0]))
    assert len(graphed._graphs) == 1


def test_host_pipeline_with_cuda_graph_equals_plain():
    from deepclr_b200 import synth
    from deepclr_b200.models import HostPipeline
    g = load_golden('kitti_small_b2')
    model, _ = build_from_golden(g, DEV)
    batches = [torch.from_numpy(synth.batch('kitti', 2, 500 + i, n=int(g['n']))).pin_memory() for i in range(7)]
    plain = HostPipeline(model).run_all(batches)
    graphed = HostPipeline(model, use_cuda_graph=True).run_all(batches)
    assert len(plain) == len(graphed) == 7
    for a, b in zip(plain, graphed):
        assert torch.equal(a, b)
