"""Double-buffered host->device upload on a side stream.

The reference uploads the step's pixel tensor synchronously from pageable memory (image_modality.py:131-132).  Here a
pinned host tensor is copied on a dedicated copy stream into one of `depth` persistent device buffers; the compute
stream only waits for that copy's event.  Because the host enqueues step i+1 while the GPU still runs step i, the copy
of step i+1 overlaps step i's kernels and disappears from the critical path.  A buffer is written again only after the
step that read it has released it (event recorded on the compute stream)."""
from __future__ import annotations

import torch


class H2DUploader:
    def __init__(self, device, depth: int = 2):
        self.device = torch.device(device)
        self.stream = torch.cuda.Stream(device=self.device)
        self.slots = [dict(buf=None, free=None) for _ in range(depth)]
        self.i = 0
        self.in_use = []
        self.last_ready = None

    def upload(self, host: torch.Tensor) -> torch.Tensor:
        """-> device copy of `host`, valid on the CURRENT stream.  Non-pinned input degrades to a plain async copy."""
        if host.is_cuda:
            return host
        if not host.is_pinned():
            return host.to(self.device, non_blocking=True)
        slot = self.slots[self.i]
        self.i = (self.i + 1) % len(self.slots)
        cur = torch.cuda.current_stream(self.device)
        buf = slot["buf"]
        if buf is None or buf.shape != host.shape or buf.dtype != host.dtype:
            buf = slot["buf"] = torch.empty(host.shape, dtype=host.dtype, device=self.device)
            slot["free"] = None
            self.stream.wait_stream(cur)      # the allocator may hand out memory whose last user is still queued on `cur`
        if slot["free"] is not None:
            self.stream.wait_event(slot["free"])
        with torch.cuda.stream(self.stream):
            buf.copy_(host, non_blocking=True)
            ready = torch.cuda.Event()
            ready.record(self.stream)
        cur.wait_event(ready)
        self.last_ready = ready
        self.in_use.append(slot)
        return buf

    def release(self):
        """Call once the kernels that read the uploaded buffers have been enqueued on the current stream."""
        if not self.in_use:
            return
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(self.device))
        for slot in self.in_use:
            slot["free"] = ev
        self.in_use = []


class HostStager:
    """The collator hands a modality its inputs as a Python LIST of per-image CPU tensors (data_loader.py:126,131,
    144-152); the reference stacks them into pageable memory and copies synchronously (image_modality.py:131-132).
    Here the list is stacked straight into one of `depth` cached PINNED staging buffers (one host memcpy per image, no
    intermediate tensor) and uploaded on the copy stream; a staging buffer is reused only after its own H2D finished
    (event), so with depth 2 the host never waits for the copy it issued one step earlier."""

    def __init__(self, device, uploader: "H2DUploader | None" = None, depth: int = 2):
        self.device = torch.device(device)
        self.uploader = uploader or H2DUploader(device)
        self.slots = [dict(buf=None, done=None) for _ in range(depth)]
        self.i = 0

    def stack_upload(self, inputs) -> torch.Tensor:
        """list of equally-shaped CPU tensors -> [B, ...] on the device, valid on the current stream."""
        shape, dt = (len(inputs), *inputs[0].shape), inputs[0].dtype
        slot = self.slots[self.i]
        self.i = (self.i + 1) % len(self.slots)
        buf = slot["buf"]
        if buf is None or buf.shape != shape or buf.dtype != dt:
            buf = slot["buf"] = torch.empty(shape, dtype=dt, pin_memory=True)
            slot["done"] = None
        if slot["done"] is not None:
            slot["done"].synchronize()      # the H2D issued `depth` steps ago out of this buffer
        torch.stack(inputs, dim=0, out=buf)
        out = self.uploader.upload(buf)
        slot["done"] = self.uploader.last_ready
        return out


class HostTensorStager:
    """Small host tensors that arrive in PAGEABLE memory (the collator's `input_ids`, `batch_idx`, `token_range`): a
    pageable `cudaMemcpyAsync` only returns once the copy has been staged, i.e. once the STREAM has reached it — the host
    then cannot run ahead of the GPU and every per-step host cost (stacking, checks) lands on the critical path (measured:
    9 ms per step at cfg4, whose ids are 4 MB).  Staging through one of `depth` cached pinned buffers makes the upload
    truly asynchronous; a buffer is reused only after its own copy has completed (event)."""

    def __init__(self, device, depth: int = 3):
        self.device = torch.device(device)
        self.slots = {}
        self.depth = depth

    def upload(self, host: torch.Tensor) -> torch.Tensor:
        if host.is_cuda:
            return host
        if host.is_pinned():
            return host.to(self.device, non_blocking=True)
        key = (tuple(host.shape), host.dtype)
        ring = self.slots.setdefault(key, dict(i=0, bufs=[None] * self.depth, evs=[None] * self.depth))
        i = ring["i"]
        ring["i"] = (i + 1) % self.depth
        if ring["bufs"][i] is None:
            ring["bufs"][i] = torch.empty(host.shape, dtype=host.dtype, pin_memory=True)
        elif ring["evs"][i] is not None:
            ring["evs"][i].synchronize()
        ring["bufs"][i].copy_(host)
        out = ring["bufs"][i].to(self.device, non_blocking=True)
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(self.device))
        ring["evs"][i] = ev
        if len(self.slots) > 16:       # shapes change from batch to batch: keep the cache bounded
            for k in list(self.slots)[:-8]:
                del self.slots[k]
        return out
