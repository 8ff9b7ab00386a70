"""Host-buffer pipeline around a model: the call a user with HOST batches makes.

The reference's inference loop (`scripts/inference.py:83-119`) copies a pair to the device, runs `predict`, synchronises
and copies the pose back, one pair at a time.  With the forward at ~2 ms per 64 KITTI pairs the 33 MB input copy
(~0.6 ms over PCIe 5) is a third of a step if it is serialised, so this pipeline keeps `depth` device input buffers and
issues the host->device copy of batch i+1 on a copy stream while batch i computes; poses come back into pinned buffers
on the compute stream.  Every byte still moves every step -- only the waiting is overlapped.
"""
from typing import Iterable, Iterator, List, Optional

import torch


def _graph_token(model, device):
    """What a graph captured from `model`'s forward depends on besides its input buffer (DeepCLR.graph_token: native
    handle identity + workspace generation); models without a native handle have none."""
    fn = getattr(model, 'graph_token', None)
    return fn(device) if fn is not None else None


class HostPipeline:
    """`for y in HostPipeline(model).run(batches): ...` -- `batches` yields pinned (or pageable) CPU tensors
    [2B, N, C] (templates first, deepclr/data/build.py:82); yields CPU tensors [B, label_dim] in order.

    A yielded pose tensor is valid once the pipeline has moved `depth` batches on (or after `synchronize()`), the usual
    contract of a non-blocking copy; `run_all` returns fully synchronised results."""

    def __init__(self, model, depth: int = 2, device: Optional[torch.device] = None, use_cuda_graph: bool = False,
                 replicas: int = 1):
        """`use_cuda_graph`: the forward over each of the `depth` device input buffers is captured once per batch shape
        and replayed (one graph launch per batch instead of ~11 kernel launches + the side-stream fork/join).
        `replicas` > 1: batch i runs on model replica i % replicas on its own compute stream (DeepCLR.replica()), so
        consecutive batches overlap on the device -- the sampling kernel of a small batch occupies one SM per cloud for
        ~0.4 ms and leaves the rest of the GPU to the dense stages of the batch before it.  Results are unchanged."""
        self.model = model
        self._use_graph = bool(use_cuda_graph)
        self._graphs: dict = {}
        self.replicas = max(1, int(replicas))
        self.depth = max(2, int(depth))
        if self.depth % self.replicas:
            self.depth = self.replicas * ((self.depth + self.replicas - 1) // self.replicas)   # slot <-> replica is fixed
        self.device = device if device is not None else next(model.parameters()).device
        if self.device.type != 'cuda':
            raise RuntimeError('HostPipeline needs the model on a CUDA device: there is no CPU path')
        self._models = [model] + [model.replica() for _ in range(self.replicas - 1)]
        self._compute = [None] + [torch.cuda.Stream(device=self.device) for _ in range(self.replicas - 1)]
        self._copy = torch.cuda.Stream(device=self.device)
        self._in: List[Optional[torch.Tensor]] = [None] * self.depth
        self._out: List[Optional[torch.Tensor]] = [None] * (self.depth + 1)
        self._loaded = [torch.cuda.Event() for _ in range(self.depth)]
        self._freed = [torch.cuda.Event() for _ in range(self.depth)]
        self._used = [False] * self.depth
        self._done = [torch.cuda.Event() for _ in range(self.depth + 1)]
        self.h2d_bytes = 0
        self.d2h_bytes = 0
        self._main_stream = None                  # the stream run() computes on (set there)

    def _stage(self, slot: int, host: torch.Tensor) -> None:
        buf = self._in[slot]
        with torch.cuda.stream(self._copy):
            if self._used[slot]:
                self._copy.wait_event(self._freed[slot])       # the forward that read this buffer has finished
            if buf is None or buf.shape != host.shape:
                # (re)allocated ON the copy stream, after the wait above: the caching allocator then never hands out a
                # block that kernels queued on the main stream still use, and the old buffer is released only once the
                # forward that read it is done; the main stream is recorded as a user of the new block
                buf = self._in[slot] = torch.empty(host.shape, dtype=torch.float32, device=self.device)
                buf.record_stream(self._compute[slot % self.replicas] or self._main())
            buf.copy_(host, non_blocking=True)
            self._loaded[slot].record(self._copy)
        self._used[slot] = True
        self.h2d_bytes += host.numel() * host.element_size()

    def _main(self):
        return torch.cuda.default_stream(self.device) if self._main_stream is None else self._main_stream

    def _forward(self, slot: int) -> torch.Tensor:
        x = self._in[slot]
        model = self._models[slot % self.replicas]
        if not self._use_graph:
            return model(x)[0]
        key = (slot, x.data_ptr(), tuple(x.shape), _graph_token(model, self.device))
        hit = self._graphs.get(slot)
        if hit is None or hit[0] != key:
            # first batch of this shape in this slot: run it directly (this also sizes the native workspace), then
            # capture the same call over the slot's own buffer -- the graph reads whatever the copy stream puts there
            y = model(x)[0]
            torch.cuda.synchronize(self.device)
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, capture_error_mode='thread_local'):   # other threads (NCCL watchdog) may call CUDA
                y_static = model(x)[0]
            # the capture itself may have grown the workspace of ANOTHER slot's shape: key on the token after it
            key = key[:3] + (_graph_token(model, self.device),)
            self._graphs[slot] = (key, g, y_static)
            return y
        hit[1].replay()
        return hit[2]

    def run(self, batches: Iterable[torch.Tensor]) -> Iterator[torch.Tensor]:
        it = iter(batches)
        nxt = next(it, None)
        i = 0
        main = self._main_stream = torch.cuda.current_stream(self.device)
        if nxt is not None:
            self._stage(0, nxt)
        while nxt is not None:
            slot = i % self.depth
            nxt = next(it, None)
            if nxt is not None:
                self._stage((i + 1) % self.depth, nxt)
            cs = self._compute[slot % self.replicas] or main          # this batch's compute stream
            if cs is not main and i < self.replicas:
                cs.wait_stream(main)                                  # work queued before run() comes first
            cs.wait_event(self._loaded[slot])
            o = i % (self.depth + 1)
            with torch.cuda.stream(cs), torch.no_grad():
                y = self._forward(slot)
                self._freed[slot].record(cs)
                if self._out[o] is None or self._out[o].shape != y.shape:
                    self._out[o] = torch.empty(y.shape, dtype=y.dtype).pin_memory()
                self._out[o].copy_(y, non_blocking=True)
                self._done[o].record(cs)
            self.d2h_bytes += y.numel() * y.element_size()
            i += 1
            yield self._out[o]

    def run_all(self, batches: Iterable[torch.Tensor]) -> List[torch.Tensor]:
        """All poses, synchronised.  The host collects batch i - (depth - 1) while the batches after it are already queued
        on the device (with `replicas` compute streams several forwards are in flight at once: waiting for batch i-1 before
        queueing batch i+1 would serialise them); a pose buffer is reused only depth + 1 batches later."""
        outs: List[torch.Tensor] = []
        pending: List = []
        lag = self.depth - 1
        for i, y in enumerate(self.run(batches)):
            pending.append((i, y))
            if len(pending) > lag:
                j, yj = pending.pop(0)
                self._done[j % (self.depth + 1)].synchronize()
                outs.append(yj.clone())
        self.synchronize()
        outs.extend(yj.clone() for _, yj in pending)
        return outs

    def synchronize(self) -> None:
        self._copy.synchronize()
        for cs in self._compute:
            (cs or torch.cuda.current_stream(self.device)).synchronize()


class RawScanPipeline(HostPipeline):
    """The same double-buffered pipeline fed with RAW scans of different lengths: `batches` yields
    `(stage, lengths)` -- stage = [total, C] float32 host tensor (pinned: deepclr_b200.data.datasets.stage_scans) with
    the 2B scans of B pairs back to back, templates first (deepclr/data/build.py:82), lengths = rows per scan.
    Per batch: one H2D copy of the raw bytes, then on the device SystematicErasing -> RangeSelection ->
    TruncateDimension (dclr_filter_clouds), the size-equalising random subsample to `n_points` rows per cloud
    (dclr_subsample_clouds) and the forward -- what the reference does per cloud in numpy on the host before its
    batch copy (deepclr/data/transforms/build.py:37-39, deepclr/models/base.py:130-135)."""

    def __init__(self, model, n_points: int, nth: int = 1, start: int = 0, min_range: float = 0.0,
                 max_range: float = float('inf'), seed: int = 0, depth: int = 2,
                 device: Optional[torch.device] = None):
        super().__init__(model, depth, device)
        self.n_points, self.seed = int(n_points), int(seed)
        self.filter_args = dict(nth=int(nth), start=int(start), min_range=float(min_range), max_range=float(max_range))
        self._meta: List[Optional[tuple]] = [None] * self.depth
        self._batch_no = 0

    def _stage(self, slot: int, item) -> None:
        stage, lengths = item
        total, C = int(stage.shape[0]), int(stage.shape[1])
        if total != sum(lengths):
            raise ValueError('stage rows != sum(lengths)')
        buf = self._in[slot]
        with torch.cuda.stream(self._copy):
            if self._used[slot]:
                self._copy.wait_event(self._freed[slot])
            if buf is None or buf[0].shape[0] < total or buf[0].shape[1] != C or buf[1].shape[0] < len(lengths) + 1:
                cap = int(total * 1.25) + 1               # allocated on the copy stream, after the wait (see HostPipeline)
                buf = self._in[slot] = (torch.empty((cap, C), dtype=torch.float32, device=self.device),
                                        torch.empty(len(lengths) + 1, dtype=torch.int32, device=self.device))
                for t in buf:
                    t.record_stream(self._main())
            # a fresh pinned block per batch: torch's caching host allocator hands it out again only after this copy
            off = torch.empty(len(lengths) + 1, dtype=torch.int32, pin_memory=True)
            off[0] = 0
            off[1:] = torch.cumsum(torch.tensor(lengths, dtype=torch.int64), 0).to(torch.int32)
            buf[0][:total].copy_(stage, non_blocking=True)
            buf[1][:len(lengths) + 1].copy_(off[:len(lengths) + 1], non_blocking=True)
            self._loaded[slot].record(self._copy)
        self._used[slot] = True
        self._meta[slot] = (total, list(lengths))
        self.h2d_bytes += total * C * 4 + (len(lengths) + 1) * 4

    def _forward(self, slot: int) -> torch.Tensor:
        from ..data import device as dev_ops
        total, lengths = self._meta[slot]
        pts, off = self._in[slot]
        rc = dev_ops.RaggedClouds(pts[:total], off[:len(lengths) + 1], lengths)
        rc = dev_ops.filter_clouds(rc, out_dim=self.model.get_input_dim(), **self.filter_args)
        x = dev_ops.subsample_clouds(rc, self.n_points, seed=self.seed + self._batch_no)
        self._batch_no += 1
        return self.model(x)[0]


class GraphedForward:
    """The whole fused forward of ONE input shape captured in a CUDA graph: `y = GraphedForward(model, x_example)(x)`.
    For the latency case (one pair per call, `scripts/inference.py:104`): the ~11 kernels of a forward, the fork / join
    of the ball-query grid build and the output allocation become one graph launch, so the CPU-side launch cost and
    the gaps between small kernels disappear.  `x` is copied into the graph's static input buffer (device -> device, or
    host -> device when it is a pinned CPU tensor); the returned poses are a static buffer that the next call overwrites
    (clone to keep)."""

    def __init__(self, model, example: torch.Tensor, warmup: int = 3, alias_input: bool = False):
        """`alias_input`: capture over `example`'s own storage (a contiguous fp32 CUDA tensor the caller refills in
        place) instead of a private copy; `replay()` then runs the forward on whatever it holds."""
        dev = next(model.parameters()).device
        if dev.type != 'cuda':
            raise RuntimeError('GraphedForward needs the model on a CUDA device: there is no CPU path')
        self.model = model
        if alias_input:
            if not (example.is_cuda and example.is_contiguous() and example.dtype == torch.float32):
                raise ValueError('alias_input needs a contiguous float32 CUDA tensor')
            self.static_x = example
        else:
            self.static_x = torch.empty(tuple(example.shape), dtype=torch.float32, device=dev)
            self.static_x.copy_(example)
        self._dev, self._warmup = dev, max(1, warmup)
        self.captures = 0
        self._capture()

    def _capture(self) -> None:
        dev, model = self._dev, self.model
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side), torch.no_grad():
            for _ in range(self._warmup):        # sizes the native workspace, creates the side stream: no cudaMalloc later
                model(self.static_x)
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph, capture_error_mode='thread_local'), torch.no_grad():
            self.static_y = model(self.static_x)[0]
        self._token = _graph_token(model, dev)
        self.captures += 1

    def replay(self) -> torch.Tensor:
        """Forward over the current content of the static input buffer.  The graph holds pointers into the native
        handle's workspace and weights: when either was replaced since the capture (another shape grew the workspace,
        `load_state_dict` / a numerics change rebuilt the handle) the forward is captured again first."""
        if _graph_token(self.model, self._dev) != self._token:
            self._capture()
        self.graph.replay()
        return self.static_y

    def __call__(self, x: torch.Tensor) -> torch.Tensor:
        if tuple(x.shape) != tuple(self.static_x.shape):
            raise ValueError(f'graph captured for {tuple(self.static_x.shape)}, got {tuple(x.shape)}')
        self.static_x.copy_(x, non_blocking=True)
        return self.replay()


class GraphedSequentialStep:
    """One frame of sequential (odometry) prediction as a single CUDA-graph launch: set abstraction of the new scan,
    merge + output against the cached features of the previous frame, and the cache update -- the per-frame step of
    ModelInferenceHelper(is_sequential=True) (deepclr/models/base.py:97-112, driver loop scripts/inference.py:83-119).
    `step(cloud)` returns the pose of `cloud` against the previous one (a static buffer: clone to keep), None for the
    first frame.  One instance per cloud shape; a different shape or `reset()` starts a new sequence."""

    def __init__(self, model, example: torch.Tensor):
        dev = next(model.parameters()).device
        if dev.type != 'cuda':
            raise RuntimeError('GraphedSequentialStep needs the model on a CUDA device: there is no CPU path')
        self.model, self._dev = model, dev
        self.static_x = torch.empty((1,) + tuple(example.shape), dtype=torch.float32, device=dev)
        self.static_x[0].copy_(example)
        self._prev: Optional[torch.Tensor] = None
        self._has_prev = False
        self.graph = None
        self._token = None

    def reset(self) -> None:
        self._has_prev = False

    def _body(self):
        feat = self.model.cloud_features(self.static_x)                  # [1, 3+F, npoint]
        y = self.model.forward(torch.cat((self._prev, feat), 0), is_feat=True)[0]
        self._prev.copy_(feat)
        return y

    def _capture(self) -> None:
        keep = self._prev.clone()
        side = torch.cuda.Stream(device=self._dev)
        side.wait_stream(torch.cuda.current_stream(self._dev))
        with torch.cuda.stream(side), torch.no_grad():
            for _ in range(2):                   # sizes the native workspace; the cache is restored below
                self._body()
        torch.cuda.current_stream(self._dev).wait_stream(side)
        torch.cuda.synchronize(self._dev)
        x_keep = self.static_x.clone()
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph, capture_error_mode='thread_local'), torch.no_grad():
            self.static_y = self._body()
        self._prev.copy_(keep)
        self.static_x.copy_(x_keep)
        self._token = _graph_token(self.model, self._dev)

    def step(self, cloud: torch.Tensor) -> Optional[torch.Tensor]:
        if tuple(cloud.shape) != tuple(self.static_x.shape[1:]):
            raise ValueError(f'graph captured for {tuple(self.static_x.shape[1:])}, got {tuple(cloud.shape)}')
        self.static_x[0].copy_(cloud, non_blocking=True)
        if not self._has_prev:
            with torch.no_grad():
                feat = self.model.cloud_features(self.static_x)
            if self._prev is None:
                self._prev = torch.empty_like(feat)
            self._prev.copy_(feat)
            self._has_prev = True
            return None
        if self.graph is None or _graph_token(self.model, self._dev) != self._token:
            self._capture()
        self.graph.replay()
        return self.static_y
