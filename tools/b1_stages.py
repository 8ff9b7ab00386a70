import sys, torch, json
sys.path.insert(0, '.')
import bench
model, _ = bench.build_model_and_cfg('kitti'); model = model.cuda(); model.numerics = 'fp16'
x = torch.from_numpy(bench.make_batches('kitti', 1, 16384, 1, 1)[0]).cuda()
with torch.no_grad():
    for _ in range(5): model(x)
    n = model._native; n.set_timing(True)
    for _ in range(30): model(x)
    torch.cuda.synchronize(); st, calls = n.collect_timing(); n.set_timing(False)
print({k: round(v / calls, 4) for k, v in st.items()}, round(sum(st.values()) / calls, 4))
