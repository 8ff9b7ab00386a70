"""GPU parity at the configurations BASELINE.json names, beyond the committed goldens ("weak" 3):
  * a B = 16 batch of full-size KITTI-shaped pairs against the oracle run on the spot -- the persistent tcgen05 kernels
    walk several waves of tiles per CTA there and pool with atomicMax, which the B <= 2 goldens never exercise;
  * a ModelNet40-shaped pair with partial overlap (configs[3]): a visible share of the flow-embedding rows is
    radius-masked (deepclr/models/deepclr.py:220-223);
  * cloud_features(x, m) with an augmentation matrix (deepclr.py:512-514, tgm.transform_points);
  * the stage-level C entries (dclr_sa_msg_fused / dclr_flow_embed_fused / dclr_pose_head_fused) against the whole path;
  * CUDA graphs that outlive a workspace re-allocation or a weight update.
Tolerances are those of tests/test_gpu_model.py."""
import os

import numpy as np
import pytest
import torch

from oracle import model_oracle
from tests.test_gpu_model import TOL, _check_against, _modes
from tests.util import build_from_golden, load_golden

pytestmark = pytest.mark.gpu
DEV = 'cuda:0'


def _live(kind, dbg, y_ref):
    return dict(kind=np.array(kind), fps_idx=dbg['fps_idx'], ball_cnt0=dbg['ball_cnt0'], ball_cnt1=dbg['ball_cnt1'],
                knn_idx=dbg['knn_idx'], cloud_feat=dbg['cloud_feat'], motion_feat=dbg['motion_feat'], y=y_ref)


def test_forward_matches_oracle_at_batch_16_full_size():
    from deepclr_b200 import synth
    g = load_golden('kitti_b1')
    model, cfg = build_from_golden(g, DEV)
    sd = {k: v.cpu() for k, v in model.state_dict().items()}
    x_np = synth.batch('kitti', 16, 700)
    dbg = {}
    y_ref = model_oracle.forward(sd, cfg, torch.from_numpy(x_np), debug=dbg).numpy()
    for mode in _modes():
        _check_against(_live('kitti', dbg, y_ref), model, x_np, mode)


@pytest.mark.parametrize('noise', [0.0, 0.02, 0.04])
def test_partial_overlap_modelnet_pairs(noise):
    from deepclr_b200 import synth
    g = load_golden('modelnet_b2')
    model, cfg = build_from_golden(g, DEV)
    sd = {k: v.cpu() for k, v in model.state_dict().items()}
    x_np = synth.batch('modelnet', 3, 40, partial=True, noise=noise)
    dbg = {}
    y_ref = model_oracle.forward(sd, cfg, torch.from_numpy(x_np), debug=dbg).numpy()
    # how many (template sample, neighbour) rows lie outside the merge radius: the crop must make the mask matter
    B, M = 3, dbg['cloud_feat'].shape[2]
    xyz = dbg['cloud_feat'][:, :3].transpose(0, 2, 1).reshape(-1, 3)            # [2B*M, 3] rows, templates first
    row, col = dbg['knn_idx'][0], dbg['knn_idx'][1]
    dist = np.linalg.norm(xyz[B * M + col] - xyz[row], axis=1)
    radius = float(cfg['params']['merge']['params']['radius'])
    masked = float((dist >= radius).mean())
    assert masked > 0.02, f'only {masked:.3%} of the rows are radius-masked'
    for mode in _modes():
        _check_against(_live('modelnet', dbg, y_ref), model, x_np, mode)


def test_cloud_features_with_augmentation_matrix():
    """`m` [2B,4,4] moves every cloud before the set abstraction, in place, as tgm.transform_points(m, xyz) does
    (homogeneous 4x4; rigid here, so no perspective divide); everything after it must equal the oracle on the moved cloud."""
    from deepclr_b200 import synth
    g = load_golden('kitti_small_b2')
    model, cfg = build_from_golden(g, DEV)
    sd = {k: v.cpu() for k, v in model.state_dict().items()}
    x_np = synth.batch('kitti', 2, 77, n=int(g['n']))
    rng = np.random.default_rng(5)
    m_np = np.stack([synth.random_motion(rng, 2.0, 10.0) for _ in range(4)])
    x = torch.from_numpy(x_np.copy()).to(DEV)
    m = torch.from_numpy(m_np).float().to(DEV)
    torch.backends.cudnn.allow_tf32 = False        # composed route: torch's 1x1 convolutions, strict fp32 for this comparison
    torch.backends.cuda.matmul.allow_tf32 = False
    with torch.no_grad():
        feat = model.cloud_features(x, m=m)
    moved = x.cpu().numpy()                                        # the reference transforms x in place (deepclr.py:514)
    want_xyz = np.einsum('bij,bnj->bni', m_np[:, :3, :3], x_np[:, :, :3].astype(np.float64)) + m_np[:, None, :3, 3]
    np.testing.assert_allclose(moved[:, :, :3], want_xyz, rtol=1e-5, atol=1e-4)      # fp32 matmul of ~80 m coordinates
    np.testing.assert_array_equal(moved[:, :, 3], x_np[:, :, 3])
    dbg = {}
    ref = model_oracle.cloud_features(sd, cfg['params'], torch.from_numpy(moved), debug=dbg).numpy()
    got = feat.cpu().numpy()
    assert np.array_equal(got[:, :3], ref[:, :3])
    assert np.abs(got - ref).max() <= 1e-4 * np.abs(ref[:, 3:]).max()


def test_stage_level_entries_equal_the_whole_path():
    from deepclr_b200 import synth
    g = load_golden('kitti_small_b2')
    model, _ = build_from_golden(g, DEV)
    x = torch.from_numpy(synth.batch('kitti', 3, 90, n=int(g['n']))).to(DEV)
    for mode in _modes():
        model.numerics = mode
        with torch.no_grad():
            model(x)
            native = model._native
            y, feat, motion = native.forward_pairs(x, want_feat=True, want_motion=True)
            assert native.workspace_bytes(6, int(g['n'])) > 0
            new_xyz = feat[:, :3].transpose(1, 2).contiguous()
            sa = native.sa_msg(x, new_xyz)
            assert torch.equal(sa, feat[:, 3:])
            emb = native.flow_embed(feat)
            assert torch.equal(emb[:, :3], motion[:, :3])
            tol = 1e-6 if mode == 'fp32' else TOL[mode]['feat'] * 0.1
            assert float((emb - motion).abs().max()) <= tol * float(motion[:, 3:].abs().max())
            y2 = native.pose_head(motion)
            np.testing.assert_allclose(y2.cpu().numpy(), y.cpu().numpy(), atol=2e-3 if mode != 'fp32' else 1e-6)


def test_graphs_survive_workspace_growth_and_weight_updates():
    """A graph captured for a small shape, then a larger shape grows the native workspace (the old one is freed), then
    the small graph is used again: it must notice (dclr_model_generation) and re-capture instead of replaying kernels
    that point into freed memory.  Same after the parameters change (the native handle is rebuilt)."""
    from deepclr_b200 import synth
    from deepclr_b200.models import GraphedForward, ModelInferenceHelper
    g = load_golden('kitti_small_b2')
    model, _ = build_from_golden(g, DEV)
    small = [torch.from_numpy(synth.batch('kitti', 1, 800 + i, n=1024)).to(DEV) for i in range(2)]
    big = [torch.from_numpy(synth.batch('kitti', 2, 810 + i, n=4096)).to(DEV) for i in range(2)]
    helper = ModelInferenceHelper(model, use_cuda_graph=True)
    plain = ModelInferenceHelper(model)
    for rounds in range(2):
        for xs in (small, big, small):
            for x in xs:
                assert torch.equal(helper.predict(x[-1], x[0]), plain.predict(x[-1], x[0]))
    gf = GraphedForward(model, small[0])
    first = gf.captures
    with torch.no_grad():
        model(torch.from_numpy(synth.batch('kitti', 4, 820, n=8192)).to(DEV))       # grows the workspace
        want = model(small[1])[0].clone()
    assert torch.equal(gf(small[1]), want) and gf.captures == first + 1
    with torch.no_grad():
        for p in model.parameters():
            p.mul_(1.01)
        want = model(small[0])[0].clone()
    assert torch.equal(gf(small[0]), want) and gf.captures == first + 2


def test_sequential_mode_with_cuda_graph_equals_plain():
    """Odometry mode (deepclr/models/base.py:97-112): the per-frame step replayed from one CUDA graph gives the poses of
    the plain per-frame calls, over a sequence that changes its cloud size midway and is reset once."""
    from deepclr_b200 import synth
    from deepclr_b200.models import ModelInferenceHelper
    g = load_golden('kitti_small_b2')
    model, _ = build_from_golden(g, DEV)
    frames = [torch.from_numpy(synth.kitti_pair(900 + i, 2048 if i < 5 else 3000)[i % 2]).to(DEV) for i in range(8)]
    plain = ModelInferenceHelper(model, is_sequential=True)
    graphed = ModelInferenceHelper(model, is_sequential=True, use_cuda_graph=True)
    for rep in range(2):
        for f in frames:
            a, b = plain.predict(f), graphed.predict(f)
            assert (a is None) == (b is None)
            if a is not None:
                assert torch.equal(a, b)
            assert plain.has_state() and graphed.has_state()
        plain.reset_state()
        graphed.reset_state()
        assert not graphed.has_state()


@pytest.mark.parametrize('use_graph', [False, True])
def test_host_pipeline_with_replicas_equals_plain(use_graph):
    """Several model replicas on several streams (DeepCLR.replica(), HostPipeline(replicas=3)): consecutive batches
    overlap on the device, the poses are those of the single-stream pipeline, in order."""
    from deepclr_b200 import synth
    from deepclr_b200.models import HostPipeline
    g = load_golden('kitti_small_b2')
    model, _ = build_from_golden(g, DEV)
    batches = [torch.from_numpy(synth.batch('kitti', 2, 600 + i, n=int(g['n']))).pin_memory() for i in range(11)]
    plain = HostPipeline(model).run_all(batches)
    multi = HostPipeline(model, depth=6, replicas=3, use_cuda_graph=use_graph).run_all(batches)
    assert len(plain) == len(multi) == 11
    for a, b in zip(plain, multi):
        assert torch.equal(a, b)


@pytest.mark.parametrize('golden,kind,n', [('modelnet_b2', 'modelnet', None), ('kitti_small_b2', 'kitti', None),
                                           ('kitti_b1', 'kitti', 16384)])
def test_sa_tensor_core_kernel_equals_simt_kernel(golden, kind, n):
    """sa_tc_kernel (tcgen05, fp16 hi/lo split operands) against sa_mlp_kernel (fp32 FFMA) on the same centres and
    neighbour lists: fp32-class agreement (the split carries 22 significant bits), dense ModelNet-shaped and sparse
    KITTI-shaped neighbourhoods, input_dim 3 and 4, partial last block of centres."""
    from deepclr_b200 import synth
    g = load_golden(golden)
    model, _ = build_from_golden(g, DEV)
    model.numerics = 'fp16'
    n = int(g['n']) if n is None else n
    x = torch.from_numpy(synth.batch(kind, 3, 33, n=n)).to(DEV)
    with torch.no_grad():
        model(x)
        native = model._native
        _, feat, _ = native.forward_pairs(x, want_feat=True, want_motion=True)
        new_xyz = feat[:, :3].transpose(1, 2).contiguous()
        old = os.environ.get('DCLR_SA_ROUTE')
        try:
            os.environ['DCLR_SA_ROUTE'] = 'simt'
            ref = native.sa_msg(x, new_xyz)
            os.environ['DCLR_SA_ROUTE'] = 'tc'
            got = native.sa_msg(x, new_xyz)
            again = native.sa_msg(x, new_xyz)
        finally:
            if old is None:
                os.environ.pop('DCLR_SA_ROUTE', None)
            else:
                os.environ['DCLR_SA_ROUTE'] = old
    assert torch.equal(got, again)                                   # atomicMax on integer images: order independent
    scale = float(ref.abs().max())
    assert scale > 0
    assert float((got - ref).abs().max()) <= 2e-6 * scale + 1e-7


@pytest.mark.parametrize('golden,kind,pairs,n', [('kitti_b1', 'kitti', 5, 16384), ('modelnet_b2', 'modelnet', 24, None)])
def test_forward_is_bit_reproducible_over_repeated_calls(golden, kind, pairs, n):
    """40 forwards of one batch, alone and with a second replica running other batches on another stream: every result
    equals the first bit for bit.  Nothing in the path may depend on timing -- tiles drawn by ticket (head), atomicMax
    pooling (head, tensor-core SA), samples per FPS round, which warp issues which weight tile."""
    from deepclr_b200 import synth
    g = load_golden(golden)
    model, _ = build_from_golden(g, DEV)
    model.numerics = 'fp16'
    n = int(g['n']) if n is None else n
    x = torch.from_numpy(synth.batch(kind, pairs, 71, n=n)).to(DEV)
    other = [torch.from_numpy(synth.batch(kind, pairs, 72 + i, n=n)).to(DEV) for i in range(2)]
    rep = model.replica()
    side = torch.cuda.Stream(device=DEV)
    with torch.no_grad():
        first = model(x)[0].clone()
        for i in range(40):
            if i >= 20:                                   # the second half shares the GPU with another replica
                side.wait_stream(torch.cuda.current_stream())
                with torch.cuda.stream(side):
                    rep(other[i % 2])
            assert torch.equal(model(x)[0], first), f'call {i} differs'
        torch.cuda.synchronize()


@pytest.mark.parametrize('npoint,k', [(100, 20), (77, 9)])
def test_tensor_core_sa_with_a_partial_block_of_centres(npoint, k):
    """npoint not a multiple of the 32 centres a CTA of sa_tc_kernel owns (last block partial) and smaller than a tile of
    the other kernels: both SA kernels agree, and the whole fp16 path agrees with the fp32 path within the fp16 tolerance."""
    import yaml
    from deepclr_b200 import synth
    from deepclr_b200.config.config import Config
    from deepclr_b200.data.labels import LabelType
    from deepclr_b200.models.build import ModelType, build_model
    g = load_golden('kitti_small_b2')
    cfg = yaml.safe_load(str(g['cfg_yaml']))
    cfg['params']['cloud_features']['params']['npoint'] = [npoint]
    cfg['params']['merge']['params']['k'] = k
    c = Config._from_dict(cfg)
    c.label_type = LabelType.create(c.label_type)
    c.model_type = ModelType.create(c.model_type)
    torch.manual_seed(3)
    model = build_model(c).eval().to(DEV)
    x = torch.from_numpy(synth.batch('kitti', 2, 55, n=3000)).to(DEV)
    with torch.no_grad():
        model.numerics = 'fp32'
        y32 = model(x)[0].clone()
        model.numerics = 'fp16'
        y16 = model(x)[0].clone()
        native = model._native
        _, feat, _ = native.forward_pairs(x, want_feat=True, want_motion=True)
        new_xyz = feat[:, :3].transpose(1, 2).contiguous()
        old = os.environ.get('DCLR_SA_ROUTE')
        try:
            os.environ['DCLR_SA_ROUTE'] = 'simt'
            ref = native.sa_msg(x, new_xyz)
            os.environ['DCLR_SA_ROUTE'] = 'tc'
            got = native.sa_msg(x, new_xyz)
        finally:
            if old is None:
                os.environ.pop('DCLR_SA_ROUTE', None)
            else:
                os.environ['DCLR_SA_ROUTE'] = old
    assert float((got - ref).abs().max()) <= 2e-6 * float(ref.abs().max()) + 1e-7
    np.testing.assert_allclose(y16.cpu().numpy(), y32.cpu().numpy(), atol=2e-2)
