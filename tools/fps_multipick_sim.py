"""How many samples can one block-wide round of the bucketed FPS kernel emit, exactly?  (CPU simulation, no GPU.)

State at the barrier of fps_bucket.cu: every warp w (NW = 8, buckets g = i * NW + w of the Morton-sorted cloud) knows its
best point (t, tie key) and an upper bound `w2` of every other point's running minimum in the warp.  Sort the NW warp
winners c1 >= c2 >= ... (value, then tie order).  c_i is PROVABLY the i-th next sample when, for every l < i,
    d(c_i, c_l) >= t[c_i]          (its minimum is untouched by the earlier picks of the round) and
    w2[warp(c_l)] <  t[c_i]        (nothing left in an earlier winner's warp can overtake it) and t[c_i] > 0.
The simulation runs plain FPS (numpy fp32, lowest index on ties) and, at each round start, evaluates how long that
accepted prefix would be for a cap of m picks; it then advances by exactly that many samples, so the histogram is the one
the kernel would see.  Output: JSON lines {kind, cloud, m, rounds, mean_picks, hist}.
"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from deepclr_b200 import synth  # noqa: E402


def morton_order(xyz, G=64):
    lo, hi = xyz.min(0), xyz.max(0)
    ext = hi - lo
    drop = int(np.argmin(ext))
    a0, a1 = [a for a in range(3) if a != drop]
    c0 = np.clip(np.floor((xyz[:, a0] - lo[a0]) / max(ext[a0] / G, 1e-30)), 0, G - 1).astype(np.int64)
    c1 = np.clip(np.floor((xyz[:, a1] - lo[a1]) / max(ext[a1] / G, 1e-30)), 0, G - 1).astype(np.int64)

    def spread(v):
        v = (v | (v << 4)) & 0x30f
        v = (v | (v << 2)) & 0x333
        v = (v | (v << 1)) & 0x555
        return v
    code = spread(c0) | (spread(c1) << 1)
    return np.argsort(code, kind='stable')


def simulate(xyz, npoint, m_cap, NW=8, bucket=64, group_of=None):
    """-> (picks per round, chosen indices).  group_of[k] = reduction group (warp) of point k in sorted order."""
    order = morton_order(xyz)
    p = xyz[order].astype(np.float32)
    orig = order.astype(np.int64)
    N = p.shape[0]
    if group_of is None:
        group_of = (np.arange(N) // bucket) % NW
    ngroups = int(group_of.max()) + 1
    members = [np.nonzero(group_of == g)[0] for g in range(ngroups)]
    t = np.full(N, 1e10, np.float32)
    start = int(np.nonzero(orig == 0)[0][0])

    def upd(s):
        d = p - p[s]
        d2 = (d[:, 2] * d[:, 2] + (d[:, 0] * d[:, 0] + d[:, 1] * d[:, 1])).astype(np.float32)
        np.minimum(t, d2, out=t)

    upd(start)
    chosen = [int(orig[start])]
    picks = []
    while len(chosen) < npoint:
        win, wval, w2 = [], [], []
        for g in range(ngroups):
            mem = members[g]
            tv = t[mem]
            mx = tv.max()
            cand = mem[tv == mx]
            w = cand[np.argmin(orig[cand])]
            win.append(int(w))
            wval.append(float(mx))
            if len(cand) > 1:
                w2.append(float(mx))
            else:
                rest = tv[mem != w]
                w2.append(float(rest.max()) if rest.size else -1.0)
        idx = sorted(range(ngroups), key=lambda g: (-wval[g], orig[win[g]]))
        acc = [idx[0]]
        for g in idx[1:m_cap]:
            c = win[g]
            tc = wval[g]
            if not tc > 0:
                break
            ok = True
            for a in acc:
                d = p[c] - p[win[a]]
                d2 = np.float32(d[2] * d[2] + (d[0] * d[0] + d[1] * d[1]))
                if not (d2 >= np.float32(tc) and w2[a] < tc):
                    ok = False
                    break
            if not ok:
                break
            acc.append(g)
        acc = acc[: npoint - len(chosen)]
        for g in acc:
            chosen.append(int(orig[win[g]]))
        for g in acc:
            upd(win[g])
        picks.append(len(acc))
    return picks, chosen


def plain_fps(xyz, npoint):
    p = xyz.astype(np.float32)
    t = np.full(p.shape[0], 1e10, np.float32)
    out = [0]
    for _ in range(npoint - 1):
        d = p - p[out[-1]]
        d2 = (d[:, 2] * d[:, 2] + (d[:, 0] * d[:, 0] + d[:, 1] * d[:, 1])).astype(np.float32)
        np.minimum(t, d2, out=t)
        out.append(int(np.argmax(t)))
    return out


def main():
    nclouds = int(sys.argv[1]) if len(sys.argv) > 1 else 16
    out = []
    for kind, n, npoint in (('kitti', 16384, 1024), ('modelnet', 1024, 512)):
        for c in range(nclouds):
            pair = synth.kitti_pair(100 + c // 2, n) if kind == 'kitti' else synth.modelnet_pair(100 + c // 2, n)
            xyz = pair[c % 2][:, :3]
            ref = plain_fps(xyz, npoint) if c < 2 else None
            for m in (2, 3, 4, 8):
                if kind == 'kitti':
                    picks, chosen = simulate(xyz, npoint, m)
                else:           # small clouds: the resident kernel's groups = warps of consecutive 32-point runs
                    N = xyz.shape[0]
                    picks, chosen = simulate(xyz, npoint, m, group_of=(np.arange(N) // 32) % 8)
                if ref is not None:
                    assert chosen == ref, (kind, c, m)
                hist = np.bincount(picks, minlength=m + 1)[1:].tolist()
                rec = {'kind': kind, 'cloud': c, 'm': m, 'rounds': len(picks), 'serial_rounds': npoint - 1,
                       'mean_picks': round((npoint - 1) / len(picks), 3), 'hist': hist}
                print(json.dumps(rec), flush=True)
                out.append(rec)
    for kind in ('kitti', 'modelnet'):
        for m in (2, 3, 4, 8):
            r = [x for x in out if x['kind'] == kind and x['m'] == m]
            print(json.dumps({'kind': kind, 'm': m, 'clouds': len(r),
                              'mean_rounds': float(np.mean([x['rounds'] for x in r])),
                              'mean_picks': float(np.mean([x['mean_picks'] for x in r]))}))
    # sensitivity to the number of candidate groups (one KITTI-shaped cloud): finer candidates accept more per round
    xyz = synth.kitti_pair(100, 16384)[0][:, :3]
    for NW in (8, 16, 32, 256):
        for m in (2, 4, 8):
            picks, _ = simulate(xyz, 1024, m, NW=NW)
            print(json.dumps({'kind': 'kitti', 'groups': NW, 'm': m, 'rounds': len(picks),
                              'hist': np.bincount(picks, minlength=m + 1)[1:].tolist()}), flush=True)


if __name__ == '__main__':
    main()
