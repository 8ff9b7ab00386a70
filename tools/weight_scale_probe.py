import sys, torch, numpy as np
sys.path.insert(0, '.')
from tests.util import build_from_golden, load_golden
from deepclr_b200 import synth
g = load_golden('kitti_b1')
x = torch.from_numpy(synth.batch('kitti', 4, 123)).cuda()
for scale, bstd in ((1.0, 0.0), (1.1, 0.05), (1.2, 0.1), (1.35, 0.2), (1.5, 0.3)):
    model, _ = build_from_golden(g, 'cuda:0')
    torch.manual_seed(1)
    with torch.no_grad():
        for n_, p in model.named_parameters():
            if p.dim() > 1: p.mul_(scale)
            else: p.add_(bstd * torch.randn_like(p))
        out = {}
        for mode in ('fp32', 'fp16', 'bf16'):
            model.numerics = mode
            out[mode] = model(x)[0].double().cpu().numpy()
    def rot_deg(a, b):
        qa, qb = a[:, :4] / np.linalg.norm(a[:, :4], axis=1, keepdims=True), b[:, :4] / np.linalg.norm(b[:, :4], axis=1, keepdims=True)
        d = np.abs((qa * qb).sum(1)).clip(0, 1)
        return np.degrees(2 * np.arccos(d)).max()
    print(scale, bstd, 'max|y|', np.abs(out['fp32']).max().round(3), 'fp16 rot', rot_deg(out['fp32'], out['fp16']).round(4), 'abs', np.abs(out['fp32'] - out['fp16']).max().round(5),
          '| bf16 rot', rot_deg(out['fp32'], out['bf16']).round(4), 'finite', np.isfinite(out['fp16']).all())
