"""bench.py -- registered pairs/s of the DeepCLR forward path (KITTI-shaped 16384-point pairs) on N B200s.

  python bench.py [--gpus N --steps K --warmup W]            this repo's sm_100a path (C ABI, libdeepclr_b200.so)
  python bench.py --impl reference [...]                     the reference algorithm on the host cores (CPU oracle)

One "step" = one forward over `--pairs` synthetic KITTI-shaped pairs.  Default = the configuration BASELINE.json quotes
(configs[2]): 64 pairs IN TOTAL per step, sharded by pairs over the GPUs (strong scaling, 64/32/16/8 per GPU); the
weak-scaling number (64 pairs per GPU) of the same run is reported beside it (`weak_scaling`).  Pairs are independent:
no collective on the data path.  Prints ONE JSON line.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = 'registered pairs/sec (KITTI 16k pts)'
FLOP_PER_PAIR = {'kitti': 10.0849e9, 'modelnet': 4.3844e9}     # dense-equivalent MLP work, SURVEY.md section 8
ROTATE = 4                                                     # distinct resident input batches cycled per step


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=20)
    ap.add_argument('--warmup', type=int, default=5)
    ap.add_argument('--impl', default='native', choices=['native', 'reference'])
    ap.add_argument('--pairs', type=int, default=64, help='pairs per step: in total (strong scaling) / per GPU (weak)')
    ap.add_argument('--scaling', default='strong', choices=['strong', 'weak'],
                    help='strong (default, BASELINE.json configs[2]): --pairs pairs in total sharded over the GPUs')
    ap.add_argument('--streams', type=int, default=0, help='model replicas / streams per GPU (default 4)')
    ap.add_argument('--partial', action='store_true', help='ModelNet-shaped pairs with partial overlap (configs[3])')
    ap.add_argument('--noise', type=float, default=0.02, help='ModelNet point noise sigma (0, 0.02, 0.04: the reference ladder)')
    ap.add_argument('--points', type=int, default=16384)
    ap.add_argument('--kind', default='kitti', choices=['kitti', 'modelnet'])
    ap.add_argument('--numerics', default=os.environ.get('DCLR_NUMERICS', 'fp16'), choices=['fp32', 'fp16', 'bf16'])
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--no-latency', action='store_true', help='skip the batch-1 latency loop (ncu launch lists)')
    ap.add_argument('--no-raw-scans', action='store_true', help='skip the raw-scan front-end arm (e2e_raw_scans)')
    ap.add_argument('--raw-points', type=int, default=120000, help='mean points of a raw scan (KITTI HDL-64: ~120 k)')
    ap.add_argument('--cpu-pairs', type=int, default=40, help='pairs of the bounded CPU sample')
    return ap.parse_args()


def workload_name(args, world):
    if args.scaling == 'strong':
        return f'{args.kind}_{args.points}pts_batch{args.pairs}_total_sharded_over_{world}_gpus'
    return f'{args.kind}_{args.points}pts_batch{args.pairs}_per_gpu'


def load_peaks():
    p = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(hbm=float(d['hbm_gbs']), tf=float(d['bf16_tflops_sustained']), src='measured')
    return dict(hbm=6650.0, tf=1400.0, src='fallback')


def build_model_and_cfg(kind):
    """Random-init DeepCLR (torch.manual_seed(0)) from the reference-shaped model config."""
    import torch
    import yaml
    from deepclr_b200.config import load_model_config
    from deepclr_b200.models import build_model
    path = os.path.join(ROOT, 'configs', f'{"kitti" if kind == "kitti" else "modelnet40"}_model.yaml')
    cfg = load_model_config(path, '')
    torch.manual_seed(0)
    model = build_model(cfg).eval()
    return model, yaml.safe_load(open(path))


def make_batches(kind, pairs, points, count, seed0, **gen_kw):
    """`count` distinct [2*pairs, points, C] batches.  Generating 16k-point scans is slow in numpy, so a batch is
    assembled from a pool of seeded pairs; every batch is a different selection (no two identical inputs in a row)."""
    from deepclr_b200 import synth
    pool_n = min(pairs, 16)
    gen = synth.kitti_pair if kind == 'kitti' else synth.modelnet_pair
    pool = [gen(seed0 + i, points, **gen_kw)[:2] for i in range(pool_n)]
    rng = np.random.default_rng(seed0)
    out = []
    for _ in range(count):
        sel = rng.integers(0, pool_n, pairs)
        # rigidly shift every copy a little so the duplicates are not bit-identical clouds
        t = np.stack([pool[i][0] for i in sel])
        s = np.stack([pool[i][1] for i in sel])
        off = rng.uniform(-0.05, 0.05, (pairs, 1, 3)).astype(np.float32)
        t[:, :, :3] += off
        s[:, :, :3] += off
        out.append(np.ascontiguousarray(np.concatenate([t, s], 0)))
    return out


STAGE_KERNEL = {'fps': 'fps_tmem_kernel', 'ball_query': 'bq_grid_query_kernel', 'knn': 'knn_fast_kernel',
                'flow_embed': 'flow_tc_kernel', 'pose_head_conv': 'head_tc_kernel', 'sa_mlp_maxpool': 'sa_mlp_kernel'}


def ncu_traffic(stage, args):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the stage's main kernel, from the newest committed
    `ncu --set full` capture of this same workload (profiles/*_ncu_traffic.json, tools/ncu_summary.py); None when the
    workload differs from the captured one or no capture exists."""
    import glob
    if (args.kind, args.pairs, args.points) != ('kitti', 64, 16384):
        return None
    files = sorted(glob.glob(os.path.join(ROOT, 'profiles', '*_ncu_traffic.json')), key=os.path.basename)
    for f in reversed(files):
        d = json.load(open(f)).get(STAGE_KERNEL.get(stage, ''))
        if d:
            return d['dram_read_bytes'] + d['dram_write_bytes']
    return None


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons sampled during the timed region (B200_PROFILING.md)."""
    Q = ('clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,'
         'clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap')

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self._halt = index, [], threading.Event()
        # NVML is initialised HERE, before the timed region starts (nvmlInit alone can take longer than the region);
        # the thread then only polls: ~0.1 ms per sample
        self._nv = self._h = None
        try:
            import pynvml as nv
            nv.nvmlInit()
            self._h = nv.nvmlDeviceGetHandleByIndex(index)
            self._max = nv.nvmlDeviceGetMaxClockInfo(self._h, nv.NVML_CLOCK_SM)
            self._reasons = (getattr(nv, 'nvmlDeviceGetCurrentClocksEventReasons', None)
                             or nv.nvmlDeviceGetCurrentClocksThrottleReasons)
            self._nv = nv
            self._sample_nvml()
            self.rows.clear()                        # that one only proved the calls work
        except Exception as e:
            self._nv = None
            print(f'[bench] NVML unavailable ({type(e).__name__}: {e}); clocks come from nvidia-smi', file=sys.stderr)

    def _sample_nvml(self):
        nv, h = self._nv, self._h
        r = self._reasons(h)
        bits = [0x8, 0x40, 0x20, 0x4]       # hw_slowdown, hw_thermal_slowdown, sw_thermal_slowdown, sw_power_cap
        self.rows.append([str(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)), str(self._max),
                          str(nv.nvmlDeviceGetPowerUsage(h) / 1000.0)] +
                         ['Active' if r & b else 'Not Active' for b in bits])

    def _run_nvml(self):
        while not self._halt.is_set():
            self._sample_nvml()
            self._halt.wait(0.002)

    def run(self):
        try:
            if self._nv is not None:
                self._run_nvml()
                return
        except Exception as e:                      # reported, then the nvidia-smi route takes over
            print(f'[bench] NVML clock sampling failed ({type(e).__name__}: {e}); falling back to nvidia-smi',
                  file=sys.stderr, flush=True)
        while not self._halt.is_set():
            try:
                o = subprocess.run(['nvidia-smi', '-i', str(self.index), f'--query-gpu={self.Q}',
                                    '--format=csv,noheader,nounits'], capture_output=True, text=True, timeout=5).stdout
                f = [v.strip() for v in o.strip().split(',')]
                if len(f) >= 7:
                    self.rows.append(f)
            except Exception as e:
                print(f'[bench] nvidia-smi clock sample failed ({type(e).__name__}: {e})', file=sys.stderr, flush=True)
            self._halt.wait(0.1)

    def finish(self):
        self._halt.set()
        self.join(6)
        if not self.rows:                            # region shorter than one nvidia-smi call: one sample right after it
            try:
                o = subprocess.run(['nvidia-smi', '-i', str(self.index), f'--query-gpu={self.Q}',
                                    '--format=csv,noheader,nounits'], capture_output=True, text=True, timeout=5).stdout
                f = [v.strip() for v in o.strip().split(',')]
                if len(f) >= 7:
                    self.rows.append(f)
            except Exception:
                pass
        if not self.rows:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': ['unavailable']}
        sm = sorted(float(r[0]) for r in self.rows)
        names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
        reasons = [n for j, n in enumerate(names) if any(r[3 + j].lower().startswith('active') for r in self.rows)]
        return {'sm_mhz': sm[len(sm) // 2], 'sm_max_mhz': float(self.rows[0][1]), 'reasons': reasons,
                'samples': len(self.rows)}


def cpu_oracle_pairs_per_s(kind, points, n_pairs, warm=1):
    """The reference algorithm on the host cores: oracle/model_oracle.py (torch CPU fp32 dense layers with all host
    threads + OpenMP C restatements of the reference point kernels).  Bounded sample of the bench workload."""
    import torch
    from oracle import model_oracle, ops
    model, cfg = build_model_and_cfg(kind)
    sd = {k: v.detach().cpu() for k, v in model.state_dict().items()}
    xs = make_batches(kind, 1, points, min(n_pairs, 4), 1234)
    for i in range(warm):
        model_oracle.forward(sd, cfg, torch.from_numpy(xs[0]))
    t0 = time.perf_counter()
    for i in range(n_pairs):
        model_oracle.forward(sd, cfg, torch.from_numpy(xs[i % len(xs)]))
    dt = time.perf_counter() - t0
    cores = max(torch.get_num_threads(), ops.num_threads())
    return n_pairs / dt, cores, dt


def run_reference(args, rank, world):
    """--impl reference: the reference's CPU implementation of the path (the oracle port; the reference's own Python
    needs /root/reference and third-party wheels that do not exist on the GPU box).  Rank 0 only."""
    if rank != 0:
        return
    per_step = max(1, args.cpu_pairs // 10)
    import torch
    torch.set_num_threads(os.cpu_count() or 1)     # torchrun exports OMP_NUM_THREADS=1 per rank; this arm owns the host
    from oracle import model_oracle, ops
    model, cfg = build_model_and_cfg(args.kind)
    sd = {k: v.detach().cpu() for k, v in model.state_dict().items()}
    xs = make_batches(args.kind, 1, args.points, 4, 1234)
    step = lambda i: [model_oracle.forward(sd, cfg, torch.from_numpy(xs[(i + j) % 4])) for j in range(per_step)]
    for i in range(args.warmup):
        step(i)
    t0 = time.perf_counter()
    for i in range(args.steps):
        step(i)
    dt = time.perf_counter() - t0
    value = per_step * args.steps / dt
    cores = max(torch.get_num_threads(), ops.num_threads())
    sample = f'{per_step} pair(s) of {args.points} points per step, batch 1 forwards'
    line = {'impl': 'reference', 'metric': METRIC, 'value': value, 'unit': 'pairs/s', 'n_gpus': args.gpus,
            'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': 1e3 * dt / args.steps,
            'higher_is_better': True, 'scaling': args.scaling, 'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic',
            'config': {'workload': workload_name(args, world), 'sample': sample},
            'cpu_baseline': {'value': value, 'unit': 'pairs/s', 'cores': cores, 'kind': 'port', 'sample': sample},
            'e2e': {'value': value, 'unit': 'pairs/s', 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
            'gpu_launches': 0}
    print(json.dumps(line), file=RESULT, flush=True)


def ref_cuda_pairs_per_s(kind, points, pairs, tf32, steps=5, warmup=2):
    """Baseline B1: the reference's OWN CUDA kernels (oracle/_ref/libpointnet2_ref.so, compiled unmodified
    from the reference tree for sm_100a) + torch fp32 dense layers (cuDNN / cuBLAS) = the reference forward as it runs
    on this B200; kNN is a cdist + top-k STAND-IN (torch_cluster is not vendored).  Protocol of the reference's
    scripts/timing.py:32-47: CUDA events around each forward, mean over the timed calls."""
    import torch
    from oracle import ref_cuda
    torch.backends.cuda.matmul.allow_tf32 = bool(tf32)
    torch.backends.cudnn.allow_tf32 = bool(tf32)
    model, cfg = build_model_and_cfg(kind)
    sd = {k: v.cuda() for k, v in model.state_dict().items()}
    xs = [torch.from_numpy(b).cuda() for b in make_batches(kind, pairs, points, 2, 1234)]
    for i in range(warmup):
        ref_cuda.forward(sd, cfg, xs[i % 2])
    torch.cuda.synchronize()
    times = []
    for i in range(steps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        ref_cuda.forward(sd, cfg, xs[i % 2])
        b.record()
        b.synchronize()
        times.append(a.elapsed_time(b))
    del sd, xs
    torch.cuda.empty_cache()
    return pairs / (sum(times) / len(times) * 1e-3)


def run_native(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist
    from deepclr_b200 import sharding
    from deepclr_b200.models import GraphedForward, HostPipeline
    assert torch.cuda.is_available(), 'bench.py (native arm) needs a CUDA device: there is no CPU path'
    torch.cuda.set_device(local_rank)
    dev = torch.device('cuda', local_rank)
    if world > 1:
        dist.init_process_group('nccl', device_id=dev)
    peaks = load_peaks()
    model, _ = build_model_and_cfg(args.kind)
    model = model.to(dev)
    model.numerics = args.numerics
    replicas = [model] + [model.replica() for _ in range((args.streams if args.streams > 0 else 8) - 1)]
    gen_kw = dict(partial=True, noise=args.noise) if args.kind == 'modelnet' and args.partial else \
        (dict(noise=args.noise) if args.kind == 'modelnet' else {})
    # strong scaling (BASELINE.json configs[2]): `--pairs` pairs IN TOTAL, contiguous shards of pairs
    # (deepclr_b200.sharding.pair_range), 64/32/16/8 per GPU at 1/2/4/8 GPUs.  weak: `--pairs` per GPU.
    b0, b1 = sharding.pair_range(args.pairs, world, rank)
    share = {'strong': b1 - b0, 'weak': args.pairs}
    assert min(share.values()) >= 1, 'more ranks than pairs'
    full = [torch.from_numpy(b).pin_memory() for b in make_batches(args.kind, args.pairs, args.points, ROTATE,
                                                                   1234 + 1000 * rank, **gen_kw)]

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(ms):
        t = torch.tensor([ms], device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def default_lanes(B):
        """Streams / model replicas per GPU: --streams, else by the size of a step.  The sampling kernel holds one SM per
        cloud (2 B of 148) for ~0.4 ms whatever the batch, so small steps need more of them in flight.  Measured on one
        B200 (end to end, pairs/s): 64 pairs per step 2 | 3 | 4 | 6 streams 65.6 k | 66.2 k | 65.8 k | 65.6 k; 8 pairs per
        step 4 | 8 | 12 | 16 streams 37.4 k | 51.5 k | 50.8 k | 50.0 k.  Eight ranks on one box (8 pairs per GPU and step,
        resident | end to end): 4 streams 320 k | 267 k, 6 streams 308 k | 316 k, 8 streams 391 k | 235 k -- with eight
        processes the host side of the end-to-end loop (Python: ~15 stream / event calls per step) competes for the
        cores, so small steps default to 6."""
        if args.streams > 0:
            return args.streams
        return 4 if B >= 32 else 6

    def measure(B, steps, warmup):
        """Resident and end-to-end throughput of this rank's B pairs per step.  Consecutive steps rotate over `--streams`
        model replicas on as many streams (default 4; 1 = no overlap): the sampling kernel occupies one SM per cloud and is
        latency-bound, so the other steps in flight fill the rest of the GPU; every step still does all of its work.
        Measured on one B200, resident pairs/s with 1 | 2 | 4 streams: 64 pairs per step 55.0 k | 65.3 k | 66.3 k,
        32 pairs 37.0 k | 59.8 k | 63.5 k, 16 pairs 22.7 k | 40.0 k | 56.6 k, 8 pairs 12.8 k | 24.4 k | 40.3 k."""
        host = [h[torch.cat([torch.arange(B), args.pairs + torch.arange(B)])].contiguous().pin_memory() for h in full]
        resident = [h.to(dev) for h in host]
        lanes = default_lanes(B)
        models = replicas[:lanes]
        streams = [torch.cuda.Stream(device=dev) for _ in range(lanes)] if lanes > 1 else [None]
        cur = torch.cuda.current_stream(dev)
        use_graph = os.environ.get('DCLR_GRAPH', '1') == '1'
        # one CUDA graph per (replica, resident batch), captured over the batch's own storage: a step is one graph launch,
        # as in the end-to-end arm -- at 8 pairs per step the ~110 plain launches of a forward cost more CPU time than the
        # step takes on the GPU
        graphs = {}

        def forward(i):
            key = (i % lanes, i % ROTATE)
            if not use_graph:
                return models[key[0]](resident[key[1]])
            if key not in graphs:
                graphs[key] = GraphedForward(models[key[0]], resident[key[1]], alias_input=True)
            return graphs[key].replay()

        def run_resident(n):
            for s in streams:
                if s is not None:
                    s.wait_stream(cur)
            for i in range(n):
                s = streams[i % lanes]
                with torch.no_grad():
                    if s is None:
                        forward(i)
                    else:
                        with torch.cuda.stream(s):
                            forward(i)
            for s in streams:
                if s is not None:
                    cur.wait_stream(s)

        if use_graph:                      # capture every (replica, batch) combination outside the timed region
            with torch.no_grad():
                for i in range(lanes * ROTATE):
                    forward(i)
            torch.cuda.synchronize(dev)
        run_resident(warmup)
        # short steps (a few pairs per GPU): W steps are ~1 ms of GPU work, not enough to reach steady clocks and a warm
        # allocator on every rank -- keep going, untimed, until 40 ms have passed (the timed K steps are unchanged)
        t_warm = time.perf_counter()
        while time.perf_counter() - t_warm < 0.04:
            run_resident(max(warmup, lanes))
            torch.cuda.synchronize(dev)
        import gc
        gc.collect()
        gc.disable()                    # no collector pause inside a timed region of a few milliseconds
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        sampler = ClockSampler(local_rank) if rank == 0 else None
        barrier()
        if sampler:
            sampler.start()
        e0.record()
        run_resident(steps)            # no per-stage events here: stage times come from a separate pass
        e1.record()
        barrier()
        clocks = sampler.finish() if sampler else None
        ms = max_over_ranks(e0.elapsed_time(e1))
        launches = sum(m_._native.last_launches() for m_ in models) * steps // lanes

        # end to end through the public host-buffer API: H2D of every step's inputs from pinned memory and D2H of its poses
        # inside the timed region (deepclr_b200.models.HostPipeline, forward replayed from CUDA graphs)
        pipe = HostPipeline(model, depth=2 * lanes, use_cuda_graph=use_graph, replicas=lanes)
        pipe.run_all(host[i % ROTATE] for i in range(max(2 * lanes + 1, warmup // 2)))
        t_warm = time.perf_counter()
        while time.perf_counter() - t_warm < 0.04:                       # as above: untimed, until 40 ms have passed
            pipe.run_all(host[i % ROTATE] for i in range(2 * lanes + 1))
        pipe.h2d_bytes = pipe.d2h_bytes = 0
        barrier()
        e0.record()
        outs = pipe.run_all(host[i % ROTATE] for i in range(steps))
        e1.record()
        barrier()
        ms_e2e = max_over_ranks(e0.elapsed_time(e1))
        gc.enable()
        assert len(outs) == steps and all(torch.isfinite(o).all() for o in outs)
        e2e = {'ms_per_step': ms_e2e / steps, 'h2d_bytes_per_step': pipe.h2d_bytes // steps,
               'd2h_bytes_per_step': pipe.d2h_bytes // steps,
               'api': 'deepclr_b200.models.HostPipeline.run_all (pinned host batches%s%s)' % (
                   ', forward replayed from CUDA graphs' if use_graph else '',
                   f', {lanes} model replicas on {lanes} streams' if lanes > 1 else '')}
        del pipe
        return dict(B=B, ms=ms, ms_e2e=ms_e2e, e2e=e2e, clocks=clocks, launches=launches, resident=resident, host=host,
                    lanes=lanes)

    main = measure(share[args.scaling], args.steps, args.warmup)
    B = main['B']
    pairs_per_step = args.pairs if args.scaling == 'strong' else world * args.pairs
    value = pairs_per_step * args.steps / (main['ms'] * 1e-3)
    e2e = dict(main['e2e'], value=pairs_per_step * args.steps / (main['ms_e2e'] * 1e-3), unit='pairs/s')
    other = None
    if world > 1:                       # the other scaling mode, same run, reported beside the headline
        o_mode = 'weak' if args.scaling == 'strong' else 'strong'
        o = measure(share[o_mode], args.steps, args.warmup)
        o_pairs = args.pairs if o_mode == 'strong' else world * args.pairs
        other = {'scaling': o_mode, 'pairs_per_gpu_per_step': o['B'], 'value': o_pairs * args.steps / (o['ms'] * 1e-3),
                 'ms_per_step': o['ms'] / args.steps, 'e2e_value': o_pairs * args.steps / (o['ms_e2e'] * 1e-3),
                 'streams': o['lanes']}
        del o
    resident, host = main['resident'], main['host']

    # ---- per-stage device times: a SECOND pass with CUDA events between the stages (dclr_model_set_timing), one stream
    native = model._native
    native.set_timing(True)
    with torch.no_grad():
        for i in range(min(args.steps, 16)):
            model(resident[i % ROTATE])
    stage_ms, calls = native.collect_timing()
    native.set_timing(False)
    per_call = {k: v / max(calls, 1) for k, v in stage_ms.items()}
    # genuine (centre, neighbour) rows of the last forward: what the grouped-MLP kernel really computes
    M = native.npoint
    sa_rows = int(sum(int(native.debug_tensor(1 + s, (2 * B, M)).clamp(min=1).sum().item())
                      for s in range(len(native.nsamples))))

    # ---- plain pinned host -> device copies, all ranks at once: the box's H2D ceiling next to the e2e line
    nbytes = host[0].numel() * 4
    dst = torch.empty_like(resident[0])
    for _ in range(3):
        dst.copy_(host[0], non_blocking=True)
    h0, h1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    h0.record()
    for i in range(20):
        dst.copy_(host[i % ROTATE], non_blocking=True)
    h1.record()
    barrier()
    ms_h2d = max_over_ranks(h0.elapsed_time(h1))
    h2d = {'gbps_per_rank': nbytes * 20 / (ms_h2d * 1e-3) / 1e9, 'gbps_all_ranks': world * nbytes * 20 / (ms_h2d * 1e-3) / 1e9,
           'bytes_per_copy': nbytes, 'e2e_needs_gbps_all_ranks': world * e2e['h2d_bytes_per_step'] / (e2e['ms_per_step'] * 1e-3) / 1e9}
    del dst

    # ---- raw-scan arm: what the reference's inference loop does before its batch copy -- every
    # 2nd point, range filter, random subsample to N points, per scan in numpy on the host (transforms/build.py:37-39)
    # -- moved behind the H2D copy: raw ~120 k-point scans (ragged) in pinned memory -> H2D -> dclr_filter_clouds ->
    # dclr_subsample_clouds -> forward -> D2H poses, double-buffered (deepclr_b200.models.RawScanPipeline).
    e2e_raw = None
    if args.kind == 'kitti' and not args.no_raw_scans and not args.no_latency:
        from deepclr_b200 import synth
        from deepclr_b200.models import RawScanPipeline
        rng = np.random.Generator(np.random.PCG64(4321 + rank))
        raw = []
        for _ in range(2):
            lengths = rng.integers(int(args.raw_points * 0.95), int(args.raw_points * 1.05), 2 * B).tolist()
            stage = torch.empty((sum(lengths), 4), dtype=torch.float32).pin_memory()
            pos = 0
            for n in lengths:
                xyz, inten = synth._lidar_scan(rng, n)
                stage[pos:pos + n, :3] = torch.from_numpy(xyz.astype(np.float32))
                stage[pos:pos + n, 3] = torch.from_numpy(inten.astype(np.float32))
                pos += n
            raw.append((stage, lengths))
        rpipe = RawScanPipeline(model, args.points, nth=2, start=0, min_range=2.0, max_range=80.0, seed=7)
        rpipe.run_all(raw[i % 2] for i in range(3))
        rpipe.h2d_bytes = rpipe.d2h_bytes = 0
        raw_steps = max(4, args.steps // 2)
        r0, r1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        r0.record()
        raw_out = rpipe.run_all(raw[i % 2] for i in range(raw_steps))
        r1.record()
        barrier()
        ms_raw = max_over_ranks(r0.elapsed_time(r1))
        assert len(raw_out) == raw_steps and all(torch.isfinite(o).all() for o in raw_out)
        e2e_raw = {'value': world * B * raw_steps / (ms_raw * 1e-3), 'unit': 'pairs/s', 'steps': raw_steps,
                   'ms_per_step': ms_raw / raw_steps, 'h2d_bytes_per_step': rpipe.h2d_bytes // raw_steps,
                   'd2h_bytes_per_step': rpipe.d2h_bytes // raw_steps,
                   'h2d_gbps': rpipe.h2d_bytes / raw_steps / (ms_raw / raw_steps) / 1e6,
                   'front_end': 'every 2nd point -> range [2, 80] m -> random subsample to %d points, on the device' % args.points,
                   'api': 'deepclr_b200.models.RawScanPipeline.run_all (pinned raw scans, ~%d points each)' % args.raw_points}
        del raw, rpipe

    # ---- batch-1 latency (BASELINE.json configs[1]): p50 of single-pair forwards, CUDA events per call
    lat = []
    one = [r[[0, B]].contiguous() for r in resident]
    with torch.no_grad():
        for i in range(0 if args.no_latency else 10):
            model(one[i % ROTATE])
        for i in range(1 if args.no_latency else 50):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            model(one[i % ROTATE])
            b.record()
            b.synchronize()
            lat.append(a.elapsed_time(b))
    lat.sort()
    # the same single-pair forward replayed from a CUDA graph (deepclr_b200.models.GraphedForward): one launch per call
    lat_graph = []
    if not args.no_latency:
        from deepclr_b200.models import GraphedForward
        gf = GraphedForward(model, one[0])
        for i in range(10):
            gf(one[i % ROTATE])
        for i in range(50):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            gf(one[i % ROTATE])
            b.record()
            b.synchronize()
            lat_graph.append(a.elapsed_time(b))
        lat_graph.sort()
        del gf
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- rooflines.  Per-stage times are those of the second (instrumented, single-stream) pass.
    clocks = main['clocks']
    step_ms = main['ms'] / args.steps
    kitti = args.kind == 'kitti'
    flop_stage = {'flow_embed': 2700.1e6 if kitti else 2025.1e6, 'pose_head_conv': 2149.1e6 if kitti else 1074.5e6,
                  'sa_mlp_maxpool': 5234.4e6 if kitti else 1283.4e6}
    C = 4 if kitti else 3
    ns = (512, 1024) if kitti else (256, 512)
    G, N = 2 * B, args.points
    # algorithmic bytes per launch, SURVEY.md 8(d): FPS B(12 N + 4 npoint); ball query B(12 N + 12 M + 4 M ns) summed over
    # the radii (one pass serves both); kNN 12 (Nx + Ny) + 16 Ny k per pair
    bytes_stage = {'fps': G * (12 * N + 4 * M),
                   'ball_query': G * (12 * N + 12 * M + 4 * M * sum(ns)),
                   'knn': B * (12 * 2 * M + 16 * M * native.knn_k)}
    sm_mhz = (clocks or {}).get('sm_mhz') or (clocks or {}).get('sm_max_mhz') or 1965.0
    sm_max = (clocks or {}).get('sm_max_mhz') or 1965.0
    ffma_peak = 148 * 128 * 2 * sm_max * 1e6 / 1e12                    # fp32 FFMA, TFLOP/s
    sa_flop_row = 2.0 * (C * 16 + 16 * 16 + 16 * 32)
    roof_stages = {}
    for k, t_ms in per_call.items():
        if t_ms <= 0:
            continue
        if k in ('flow_embed', 'pose_head_conv'):
            v = B * flop_stage[k] / (t_ms * 1e-3) / 1e12
            roof_stages[k] = {'bound': 'tensor', 'achieved': v, 'unit': 'TFLOP/s', 'peak': peaks['tf'], 'frac': v / peaks['tf'],
                              'ms': t_ms}
        elif k == 'sa_mlp_maxpool':
            actual = sa_rows * sa_flop_row / (t_ms * 1e-3) / 1e12
            roof_stages[k] = {'bound': 'fp32 issue (SIMT)', 'achieved': actual, 'unit': 'TFLOP/s', 'peak': ffma_peak,
                              'frac': actual / ffma_peak, 'ms': t_ms, 'rows_computed': sa_rows,
                              'rows_dense_equivalent': G * M * sum(ns),
                              'dense_equivalent_tflops': B * flop_stage[k] / (t_ms * 1e-3) / 1e12,
                              'note': 'padded duplicate rows of the reference are skipped exactly; peak = 148 SMs x 128 FFMA '
                                      'lanes x 2 x max SM clock'}
        elif k in bytes_stage:
            v = bytes_stage[k] / (t_ms * 1e-3) / 1e9
            roof_stages[k] = {'bound': 'hbm', 'achieved': v, 'unit': 'GB/s', 'peak': peaks['hbm'], 'frac': v / peaks['hbm'],
                              'ms': t_ms}
    if 'fps' in roof_stages:
        # serial rounds, one CTA per cloud: the meaningful figures are cycles per emitted sample and the SMs it occupies
        roof_stages['fps'].update(bound='latency', cycles_per_sample=per_call['fps'] * 1e-3 * sm_mhz * 1e6 / max(M - 1, 1),
                                  sms_occupied=min(G, 148), sm_count=148, hbm_frac=roof_stages['fps']['frac'],
                                  note='dependent sampling rounds (up to 3 exact samples per round); HBM bytes are read once')
    if 'ball_query' in roof_stages:
        # the fused path writes only genuine hits: no fraction of the HBM peak is claimed for bytes it does not move
        roof_stages['ball_query'].update(dense_equiv_frac=roof_stages['ball_query']['frac'], frac=None, bound='issue / L2 latency',
                                         note='dense-equivalent bytes (the reference writes fully padded index tensors); the '
                                              'fused path writes only genuine hits')
    top = max((k for k in per_call if k in roof_stages and per_call[k] > 0), key=per_call.get)
    roof = dict(roof_stages[top])
    roof.update({'kernel': STAGE_KERNEL.get(top, top), 'stage': top, 'traffic': ncu_traffic(top, args), 'peak_source': peaks['src'],
                 'ms_per_launch': per_call[top], 'share_of_step': per_call[top] / sum(per_call.values())})
    tensor_flop = flop_stage['flow_embed'] + flop_stage['pose_head_conv']
    per_gpu = value / world
    roof_e2e = {'dense_equiv_frac': per_gpu * FLOP_PER_PAIR[args.kind] / 1e12 / peaks['tf'],
                'actual_tensor_frac': per_gpu * tensor_flop / 1e12 / peaks['tf'], 'peak': peaks['tf'], 'unit': 'TFLOP/s',
                'note': 'dense_equiv = the reference\'s 10.085 / 4.384 GFLOP per pair incl. padded SA rows; actual_tensor = '
                        'flow + head GEMMs only (what the tensor pipe really computes)'}

    line = {'metric': METRIC, 'value': value, 'unit': 'pairs/s', 'n_gpus': world, 'steps': args.steps,
            'warmup': args.warmup, 'ms_per_step': step_ms, 'higher_is_better': True, 'scaling': args.scaling,
            'vs_baseline': None, 'dtype': {'bf16': 'bf16', 'fp16': 'f16', 'fp32': 'f32'}[args.numerics], 'data': 'synthetic',
            'config': {'workload': workload_name(args, world),
                       'pairs_per_step_all_gpus': pairs_per_step, 'pairs_per_gpu_per_step': B,
                       'points_per_cloud': N, 'numerics': args.numerics, 'fps_tie_rule': native.fps_tie_rule,
                       'parallelism': f'dp{world} (independent pairs, no collective on the data path)',
                       'streams_per_gpu': main['lanes'],
                       'partial_overlap': bool(args.partial), 'noise': args.noise if args.kind == 'modelnet' else None,
                       'l2': f'{ROTATE} distinct resident input batches cycled + >126 MB of per-step workspace traffic'},
            'clocks': clocks, 'e2e': e2e, 'gpu_launches': main['launches'], 'roofline': roof,
            'stage_ms': per_call, 'roofline_stages': roof_stages, 'roofline_e2e': roof_e2e, 'h2d_copy': h2d,
            'latency_b1': {'p50_ms': lat[len(lat) // 2], 'p90_ms': lat[int(len(lat) * 0.9)], 'n': len(lat),
                           'p50_graph_ms': lat_graph[len(lat_graph) // 2] if lat_graph else None,
                           'p90_graph_ms': lat_graph[int(len(lat_graph) * 0.9)] if lat_graph else None},
            'model_tflops': value * FLOP_PER_PAIR[args.kind] / 1e12}
    if other is not None:
        line[other['scaling'] + '_scaling'] = other
    if e2e_raw is not None:
        line['e2e_raw_scans'] = e2e_raw
    if world > 1:
        dist.destroy_process_group()
    if not args.no_cpu_baseline and world == 1:      # reported at N=1 only (rank 0 owns the host cores there)
        del resident, host, full
        torch.cuda.empty_cache()
        # B1: the reference's own CUDA kernels + torch fp32 layers on this GPU (north_star target: >= 10x this)
        try:
            from oracle import ref_cuda
            if ref_cuda.available():
                b1 = {'what': 'reference CUDA ops (oracle/_ref, sm_100a rebuild of the unmodified reference sources) + torch '
                              'fp32 dense layers; kNN = cdist + top-k stand-in; CUDA events per forward (scripts/timing.py:32-47)',
                      'unit': 'pairs/s', 'runs': []}
                for pairs_b1 in (1, 8, 64):
                    for tf32 in (0, 1):
                        try:
                            v = ref_cuda_pairs_per_s(args.kind, args.points, pairs_b1, tf32)
                            b1['runs'].append({'pairs_per_step': pairs_b1, 'tf32': bool(tf32), 'value': v})
                        except RuntimeError as e:             # e.g. out of memory at batch 64 (grouped tensors are materialised)
                            b1['runs'].append({'pairs_per_step': pairs_b1, 'tf32': bool(tf32), 'value': None,
                                               'error': str(e).split('\n')[0][:120]})
                            torch.cuda.empty_cache()
                best = max((r['value'] for r in b1['runs'] if r['value']), default=None)
                b1['best'] = best
                b1['speedup_vs_best'] = value / best if best else None
                b1['e2e_speedup_vs_best'] = e2e['value'] / best if best else None
                line['b1_baseline'] = b1
        except Exception as e:
            line['b1_baseline'] = {'unavailable': f'{type(e).__name__}: {e}'[:200]}
        v, cores, dt = cpu_oracle_pairs_per_s(args.kind, args.points, args.cpu_pairs)
        line['cpu_baseline'] = {'value': v, 'unit': 'pairs/s', 'cores': cores, 'kind': 'port',
                                'sample': f'{args.cpu_pairs} batch-1 forwards of {N}-point pairs ({dt:.1f} s)'}
    print(json.dumps(line), file=RESULT, flush=True)


RESULT = sys.stdout


def claim_stdout():
    """fd 1 carries the ONE result line and nothing else: whatever a library writes to stdout from here on (NCCL prints
    its version banner there when NCCL_DEBUG is set) is sent to stderr; the JSON line goes to the saved descriptor."""
    global RESULT
    sys.stdout.flush()
    real = os.dup(1)
    os.dup2(2, 1)
    RESULT = os.fdopen(real, 'w')


def main():
    args = parse_args()
    claim_stdout()
    if args.impl == 'reference':
        # rank 0 alone works in this arm and may use every host core: undo torchrun's OMP_NUM_THREADS=1 before libgomp
        # (torch, the C oracle) is loaded
        os.environ['OMP_NUM_THREADS'] = str(os.cpu_count() or 1)
    rank = int(os.environ.get('RANK', 0))
    world = int(os.environ.get('WORLD_SIZE', 1))
    local_rank = int(os.environ.get('LOCAL_RANK', 0))
    if args.impl == 'reference':
        run_reference(args, rank, world)
    else:
        run_native(args, rank, world, local_rank)


if __name__ == '__main__':
    main()
