"""One pinned host arena + one device arena for a step's inputs, moved with ONE copy.

A caller that holds the step's inputs on the host (q/k/v of the new tokens, block
table, seq_lens, query_start_loc, slot_mapping) would otherwise issue one
host-to-device copy per tensor; at decode sizes each copy is a few hundred KB at
most and the per-copy launch + DMA latency (about 8 us each) costs more than the
bytes.  `StepStaging` lays all of them out in a single pinned buffer with the same
layout on the device: fill the host views, call `upload()`, pass the device views
to `do_kv_cache_update` / `forward`.  This mirrors how vLLM's model runner keeps
its per-step metadata in persistent pinned buffers
(site-packages/vllm/v1/worker/gpu_model_runner.py, `_make_buffer`).

Plain PyTorch memory plumbing; no arithmetic happens here.
"""
from __future__ import annotations

from typing import Dict, Iterable, Tuple

import torch

_ALIGN = 256  # keeps every view 16-byte aligned for the kernels' vector loads (and TMA-friendly)


class StepStaging:
    def __init__(self, spec: Iterable[Tuple[str, Tuple[int, ...], torch.dtype]], device) -> None:
        """spec: (name, shape, dtype) of every tensor of a step, in any order."""
        self.device = torch.device(device)
        self._layout = []
        off = 0
        for name, shape, dtype in spec:
            n = 1
            for s in shape:
                n *= int(s)
            nbytes = n * torch.empty((), dtype=dtype).element_size()
            self._layout.append((name, tuple(int(s) for s in shape), dtype, off, nbytes))
            off = (off + nbytes + _ALIGN - 1) // _ALIGN * _ALIGN
        self.nbytes = max(off, _ALIGN)
        pin = self.device.type == "cuda"
        self._host = torch.empty(self.nbytes, dtype=torch.uint8, pin_memory=pin)
        self._dev = torch.empty(self.nbytes, dtype=torch.uint8, device=self.device)
        self.host: Dict[str, torch.Tensor] = {}
        self.dev: Dict[str, torch.Tensor] = {}
        for name, shape, dtype, o, nb in self._layout:
            self.host[name] = self._host[o:o + nb].view(dtype).view(shape)
            self.dev[name] = self._dev[o:o + nb].view(dtype).view(shape)

    @property
    def payload_bytes(self) -> int:
        """Bytes of tensors (without alignment padding)."""
        return sum(nb for *_, nb in self._layout)

    def upload(self) -> None:
        """One asynchronous host-to-device copy of the whole arena on the current stream."""
        self._dev.copy_(self._host, non_blocking=True)


class _Slot:
    __slots__ = ("staging", "dev_out", "host_out", "uploaded", "computed", "downloaded", "busy")


class StepPipeline:
    """`depth` StepStaging arenas + output buffers cycled through three streams: the upload of
    step i+1 (copy stream) and the read-back of step i-1 (second copy stream) overlap the
    kernels of step i (the caller's current stream).  Every step still moves all of its inputs
    host->device and its result device->host; the host blocks only when it needs a slot back,
    at which point that slot's result is complete and readable in `slot.host_out`.

        pipe = StepPipeline(spec, out_shape, out_dtype, device)
        for step in steps:
            slot = pipe.acquire()            # host_out of the step that used this slot is valid now
            ... fill slot.staging.host[...] ...
            pipe.run(slot, lambda dev, out: impl.forward(..., dev["q"], ..., output=out))
        pipe.drain()
    """

    def __init__(self, spec, out_shape, out_dtype, device, depth: int = 2) -> None:
        self.device = torch.device(device)
        spec = list(spec)
        self.h2d = torch.cuda.Stream(self.device)
        self.d2h = torch.cuda.Stream(self.device)
        self.slots = []
        for _ in range(depth):
            s = _Slot()
            s.staging = StepStaging(spec, self.device)
            s.dev_out = torch.empty(out_shape, dtype=out_dtype, device=self.device)
            s.host_out = torch.empty(out_shape, dtype=out_dtype, pin_memory=True)
            s.uploaded, s.computed, s.downloaded = (torch.cuda.Event() for _ in range(3))
            s.busy = False
            self.slots.append(s)
        self._next = 0

    def acquire(self) -> _Slot:
        s = self.slots[self._next]
        self._next = (self._next + 1) % len(self.slots)
        if s.busy:
            s.downloaded.synchronize()  # result of the step that last used this slot is on the host
            s.busy = False
        return s

    def run(self, slot: _Slot, compute) -> None:
        main = torch.cuda.current_stream(self.device)
        with torch.cuda.stream(self.h2d):
            slot.staging.upload()  # the slot's previous kernels finished: acquire() waited for its read-back
            slot.uploaded.record(self.h2d)
        main.wait_event(slot.uploaded)
        compute(slot.staging.dev, slot.dev_out)
        slot.computed.record(main)
        self.d2h.wait_event(slot.computed)
        with torch.cuda.stream(self.d2h):
            slot.host_out.copy_(slot.dev_out, non_blocking=True)
            slot.downloaded.record(self.d2h)
        slot.busy = True

    def drain(self) -> None:
        for s in self.slots:
            if s.busy:
                s.downloaded.synchronize()
                s.busy = False
