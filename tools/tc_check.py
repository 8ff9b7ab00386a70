"""Error report of the tensor-core numerics modes against the golden fixtures (GPU box)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from oracle import model_oracle
from tests.util import build_from_golden, golden_input, load_golden

for case in sys.argv[1:] or ['modelnet_b2', 'kitti_small_b2', 'kitti_b1']:
    g = load_golden(case)
    model, _ = build_from_golden(g, 'cuda:0')
    x = torch.from_numpy(golden_input(g)).cuda()
    for mode in ('fp32', 'bf16', 'fp16'):
        model.numerics = mode
        with torch.no_grad():
            y = model(x)[0]
            _, _, motion = model._native.forward_pairs(x, want_motion=True)
        torch.cuda.synchronize()
        me = np.abs(motion.cpu().numpy() - g['motion_feat']).max() / np.abs(g['motion_feat'][:, 3:]).max()
        errs = [model_oracle.pose_errors(y[b].cpu().numpy(), g['y'][b]) for b in range(y.shape[0])]
        print(case, mode, 'motion rel err %.3e' % me, 'pose (deg, m):', ['%.2e %.2e' % e for e in errs],
              'ymax %.3e' % np.abs(y.cpu().numpy() - g['y']).max(), flush=True)
