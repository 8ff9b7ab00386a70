"""Generate tests/golden/*.npz from the reference's own Python model code.  TEST INFRASTRUCTURE ONLY.

Run in the build container (needs /root/reference):   python -m oracle.gen_golden

For each case the unmodified reference modules (deepclr.models.deepclr.DeepCLR, pointnet2.PointnetSAModuleMSG, ...)
are imported on CPU with the five native ops bound to the C restatements (oracle/ref_import.py), fed seeded synthetic
pairs and seeded reference-initialised weights, and their outputs are stored.  The standalone restatement
oracle/model_oracle.py -- the checker that travels to the GPU box -- must reproduce them, otherwise this script fails.
Weights are not stored (7 MB each): both sides rebuild them from torch.manual_seed(seed) with the reference's
initialisers, and the fixture keeps per-tensor checksums so a mismatch is detected.
"""
import os
import sys

import numpy as np
import torch
import yaml

from . import model_oracle, ops, ref_import

HERE = os.path.dirname(os.path.abspath(__file__))
GOLDEN = os.path.join(os.path.dirname(HERE), 'tests', 'golden')
sys.path.insert(0, os.path.dirname(HERE))
from deepclr_b200 import synth  # noqa: E402

CASES = [
    # name, model yaml (reference tree), generator, pairs, points, seed0
    ('modelnet_b2', 'models/modelnet40/model_config.yaml', 'modelnet', 2, 1024, 0),
    ('kitti_b1', 'models/kitti_00-10/model_config.yaml', 'kitti', 1, 16384, 0),
    ('kitti_small_b2', 'models/kitti_00-10/model_config.yaml', 'kitti', 2, 2048, 10),
]


def weight_checksums(sd):
    return np.array([[float(v.double().sum()), float(v.double().abs().sum())] for v in sd.values()])


def run_case(name, cfg_rel, kind, pairs, n, seed0):
    cfg = yaml.safe_load(open(os.path.join(ref_import.REF_ROOT, cfg_rel)))
    model = ref_import.build_reference_model(cfg, seed=0)
    sd = model.state_dict()
    x = torch.from_numpy(synth.batch(kind, pairs, seed0, n=n))
    out = {}
    for mode in ('reference', 'lowidx'):
        ref_import._Fn.fps_mode = mode
        with torch.no_grad():
            feat = model.cloud_features(x.clone())
            y, _, _ = model(feat, is_feat=True)
            motion = model._merge_layers[0](feat)
        out[mode] = (y.numpy(), feat.numpy(), motion.numpy())
    same_fps = all(np.array_equal(a, b) for a, b in zip(out['reference'], out['lowidx']))
    dbg = {}
    y_o = model_oracle.forward(sd, cfg, x, debug=dbg).numpy()
    y_r, feat_r, motion_r = out['lowidx']
    err = dict(y=np.abs(y_o - y_r).max(), feat=np.abs(dbg['cloud_feat'] - feat_r).max(),
               motion=np.abs(dbg['motion_feat'] - motion_r).max())
    print(f'{name}: reference-vs-lowidx FPS identical={same_fps}; restatement vs reference modules: {err}')
    assert max(err.values()) <= 1e-5 * max(1.0, float(np.abs(motion_r).max())), 'model_oracle diverges from the reference'
    _, ties = ops.fps_lowidx(x.numpy(), cfg['params']['cloud_features']['params']['npoint'][0], return_ties=True)
    np.savez_compressed(
        os.path.join(GOLDEN, f'{name}.npz'),
        cfg_yaml=np.array(yaml.safe_dump(cfg)), kind=np.array(kind), pairs=pairs, n=n, seed0=seed0, weight_seed=0,
        weight_checksums=weight_checksums(sd), x_checksum=np.array([x.double().sum().item(), x.abs().double().sum().item()]),
        y=y_r, cloud_feat=feat_r.astype(np.float32), motion_feat=motion_r.astype(np.float32),
        y_reference_fps=out['reference'][0], fps_rules_agree=same_fps, fps_ties=ties,
        fps_idx=dbg['fps_idx'], ball_cnt0=dbg['ball_cnt0'], ball_cnt1=dbg['ball_cnt1'],
        knn_idx=dbg['knn_idx'].astype(np.int64))


def ops_case():
    """Small op-level fixtures incl. the reference test's degenerate npoint > N case (tests/model/test_deepclr.py:10-25)."""
    rng = np.random.Generator(np.random.PCG64(42))
    xyz = rng.uniform(-1, 1, (3, 300, 3)).astype(np.float32)
    xyz[1, 100] = xyz[1, 7]                    # duplicated point -> an exact distance tie
    small = rng.uniform(0, 1, (2, 96, 3)).astype(np.float32)
    fps_ref, fps_low = ops.fps(xyz, 64), ops.fps_lowidx(xyz, 64)
    new_xyz = np.stack([xyz[b, fps_low[b]] for b in range(3)])
    idx, cnt = ops.ball_query(0.35, 16, xyz, new_xyz, return_cnt=True)
    feats = rng.normal(size=(3, 5, 300)).astype(np.float32)
    np.savez_compressed(
        os.path.join(GOLDEN, 'ops_small.npz'), xyz=xyz, small=small, feats=feats,
        fps_ref=fps_ref, fps_low=fps_low, small_fps_ref=ops.fps(small, 128), small_fps_low=ops.fps_lowidx(small, 128),
        new_xyz=new_xyz, ball_idx=idx, ball_cnt=cnt, gathered=ops.gather(feats, fps_low), grouped=ops.group(feats, idx),
        knn=ops.knn(xyz[:2].reshape(-1, 3), new_xyz[:2].reshape(-1, 3), 9, 2))
    print('ops_small: fps rules agree =', np.array_equal(fps_ref, fps_low),
          '| degenerate agree =', np.array_equal(ops.fps(small, 128), ops.fps_lowidx(small, 128)))


def preproc_case():
    """Ragged pre-processing fixtures made by the reference's OWN transform classes (SystematicErasing, RangeSelection,
    TruncateDimension: deepclr/data/transforms/transforms.py:90-113,244-283) in the order of transforms/build.py:37-38;
    the numpy restatement in oracle/preproc.py is checked against them here."""
    import importlib
    from oracle import preproc
    ref_import.load()
    T = importlib.import_module('deepclr.data.transforms.transforms')
    rng = np.random.Generator(np.random.PCG64(7))
    sizes = [5000, 0, 2048, 1, 4097, 333]
    clouds = []
    for n in sizes:
        c = rng.uniform(-60, 60, (n, 5)).astype(np.float32)
        c[:, 2] = rng.uniform(-3, 3, n)
        clouds.append(c)
    clouds[0][10, 0] = np.nan                       # NaN coordinate: dropped by the range test
    clouds[0][20, :2] = [2.0, -1.0]                 # max(|x|,|y|) exactly on the lower bound: kept
    clouds[0][24, :2] = [-50.0, 3.0]                # exactly on the upper bound: kept
    clouds[0][28, :2] = [np.nextafter(np.float32(2.0), np.float32(0)), 0.5]     # just inside the hole: dropped
    out = {}
    cfgs = [dict(nth=1, start=0, rmin=0.0, rmax=np.inf, dim=4), dict(nth=2, start=0, rmin=2.0, rmax=50.0, dim=4),
            dict(nth=3, start=2, rmin=0.0, rmax=30.0, dim=3), dict(nth=1, start=0, rmin=5.0, rmax=np.inf, dim=5)]
    for ci, c in enumerate(cfgs):
        chain = [T.SystematicErasing(c['nth'], c['start']), T.RangeSelection(c['rmin'], c['rmax'], dim=3),
                 T.TruncateDimension(c['dim'])]
        sample = {'clouds': [x.copy() for x in clouds]}
        for t in chain:
            sample = t(sample)
        for b, (ref_out, raw) in enumerate(zip(sample['clouds'], clouds)):
            mine = preproc.filter_cloud(raw, c['nth'], c['start'], c['rmin'], c['rmax'], c['dim'])
            assert np.array_equal(ref_out, mine, equal_nan=True), (ci, b)
            out[f'out_{ci}_{b}'] = np.ascontiguousarray(ref_out)
        out[f'cfg_{ci}'] = np.array([c['nth'], c['start'], c['rmin'], c['rmax'], c['dim']], np.float64)
    np.savez_compressed(os.path.join(GOLDEN, 'preproc_small.npz'), num_cfgs=len(cfgs), num_clouds=len(clouds),
                        **{f'cloud_{b}': c for b, c in enumerate(clouds)}, **out)
    print('preproc_small: restatement == reference transforms on', len(cfgs), 'configs x', len(clouds), 'clouds')


def main():
    if not ref_import.available():
        raise SystemExit('reference tree not mounted; goldens can only be generated in the build container')
    os.makedirs(GOLDEN, exist_ok=True)
    ops_case()
    preproc_case()
    for case in CASES:
        run_case(*case)


if __name__ == '__main__':
    main()
