"""Host-buffer pipeline around a model: the call a user with HOST batches makes.

The reference's inference loop (`scripts/inference.py:83-119`) copies a pair to the device, runs `predict`, synchronises
and copies the pose back, one pair at a time.  With the forward at ~2 ms per 64 KITTI pairs the 33 MB input copy
(~0.6 ms over PCIe 5) is a third of a step if it is serialised, so this pipeline keeps `depth` device input buffers and
issues the host->device copy of batch i+1 on a copy stream while batch i computes; poses come back into pinned buffers
on the compute stream.  Every byte still moves every step -- only the waiting is overlapped.
"""
from typing import Iterable, Iterator, List, Optional

import torch


class HostPipeline:
    """`for y in HostPipeline(model).run(batches): ...` -- `batches` yields pinned (or pageable) CPU tensors
    [2B, N, C] (templates first, deepclr/data/build.py:82); yields CPU tensors [B, label_dim] in order.

    A yielded pose tensor is valid once the pipeline has moved `depth` batches on (or after `synchronize()`), the usual
    contract of a non-blocking copy; `run_all` returns fully synchronised results."""

    def __init__(self, model, depth: int = 2, device: Optional[torch.device] = None):
        self.model = model
        self.depth = max(2, int(depth))
        self.device = device if device is not None else next(model.parameters()).device
        if self.device.type != 'cuda':
            raise RuntimeError('HostPipeline needs the model on a CUDA device: there is no CPU path')
        self._copy = torch.cuda.Stream(device=self.device)
        self._in: List[Optional[torch.Tensor]] = [None] * self.depth
        self._out: List[Optional[torch.Tensor]] = [None] * (self.depth + 1)
        self._loaded = [torch.cuda.Event() for _ in range(self.depth)]
        self._freed = [torch.cuda.Event() for _ in range(self.depth)]
        self._used = [False] * self.depth
        self._done = [torch.cuda.Event() for _ in range(self.depth + 1)]
        self.h2d_bytes = 0
        self.d2h_bytes = 0

    def _stage(self, slot: int, host: torch.Tensor) -> None:
        buf = self._in[slot]
        if buf is None or buf.shape != host.shape:
            buf = self._in[slot] = torch.empty(host.shape, dtype=torch.float32, device=self.device)
        with torch.cuda.stream(self._copy):
            if self._used[slot]:
                self._copy.wait_event(self._freed[slot])       # the forward that read this buffer has finished
            buf.copy_(host, non_blocking=True)
            self._loaded[slot].record(self._copy)
        self._used[slot] = True
        self.h2d_bytes += host.numel() * host.element_size()

    def run(self, batches: Iterable[torch.Tensor]) -> Iterator[torch.Tensor]:
        it = iter(batches)
        nxt = next(it, None)
        i = 0
        main = torch.cuda.current_stream(self.device)
        if nxt is not None:
            self._stage(0, nxt)
        while nxt is not None:
            slot = i % self.depth
            nxt = next(it, None)
            if nxt is not None:
                self._stage((i + 1) % self.depth, nxt)
            main.wait_event(self._loaded[slot])
            with torch.no_grad():
                y = self.model(self._in[slot])[0]
            self._freed[slot].record(main)
            o = i % (self.depth + 1)
            if self._out[o] is None or self._out[o].shape != y.shape:
                self._out[o] = torch.empty(y.shape, dtype=y.dtype).pin_memory()
            self._out[o].copy_(y, non_blocking=True)
            self._done[o].record(main)
            self.d2h_bytes += y.numel() * y.element_size()
            i += 1
            yield self._out[o]

    def run_all(self, batches: Iterable[torch.Tensor]) -> List[torch.Tensor]:
        """All poses, synchronised: the host collects batch i-1 while batch i is already queued on the device."""
        outs: List[torch.Tensor] = []
        prev = None
        for i, y in enumerate(self.run(batches)):
            if prev is not None:
                self._done[(i - 1) % (self.depth + 1)].synchronize()
                outs.append(prev.clone())
            prev = y
        if prev is not None:
            self.synchronize()
            outs.append(prev.clone())
        return outs

    def synchronize(self) -> None:
        self._copy.synchronize()
        torch.cuda.current_stream(self.device).synchronize()
