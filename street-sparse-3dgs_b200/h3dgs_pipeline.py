"""Host pipeline of a training iteration: what the reference's loops do synchronously each step --
``viewpoint_cam.*.cuda()`` of four camera tensors, ``original_image.cuda()`` of a 25 MB image from pageable memory,
``loss.item()`` and ``.cpu()`` (train_post.py:77-80,109,134,148; train_single.py:116-139;
utils/camera_utils.py:126-133 keeps every image on the CPU) -- taken off the critical path:

* ``HostPrefetcher``   the inputs of step i+1 (ground-truth image, masks, camera block) cross PCIe from pinned memory
                       on a copy stream while step i computes; every step still performs its own copies.
* ``CutAhead``         the hierarchy cut of step i+1 (expand_to_size + get_interpolation_weights: functions of the
                       camera and the hierarchy's boxes, not of the parameters) runs on a side stream under step i's
                       issue-bound backward; double-buffered index tensors.
* ``LossReadback``     the scalar loss travels through pinned memory and is read one step late: the host never waits
                       for the step it has just queued (``loss.item()`` stalls the queue for a whole step).

All three only order work with CUDA events; none of them changes a value (tests/test_gpu_pipeline.py)."""
from types import SimpleNamespace

import torch


class CutBuffers:
    """The five pre-allocated cut tensors of train_post.py:59-63 and the two cut calls of :91-113."""

    def __init__(self, n_store, device):
        self.render_indices = torch.zeros(n_store, dtype=torch.int32, device=device)
        self.parent_indices = torch.zeros(n_store, dtype=torch.int32, device=device)
        self.nodes_for_render_indices = torch.zeros(n_store, dtype=torch.int32, device=device)
        self.interpolation_weights = torch.zeros(n_store, dtype=torch.float32, device=device)
        self.num_siblings = torch.zeros(n_store, dtype=torch.int32, device=device)
        self.count = 0

    def compute(self, gh, nodes, boxes, camera_center, limit, camera_center_host=None):
        """-> number of entries.  ``camera_center`` is a CUDA tensor (expand_to_size reads it on the device);
        get_interpolation_weights wants it on the host (GH/torch/torch_interface.cpp:116-117): pass the pinned host
        copy as ``camera_center_host`` to avoid a device->host sync."""
        zeros3 = torch.zeros(3)
        self.count = gh.expand_to_size(nodes, boxes, limit, camera_center, zeros3, self.render_indices,
                                       self.parent_indices, self.nodes_for_render_indices)
        host = camera_center.cpu() if camera_center_host is None else camera_center_host
        gh.get_interpolation_weights(self.nodes_for_render_indices[:self.count].contiguous(), limit, nodes, boxes, host,
                                     zeros3, self.interpolation_weights, self.num_siblings)
        return self.count


class HostPrefetcher:
    """Double-buffered host->device staging of per-step inputs on a copy stream.

    ``slots``: dict name -> (shape, dtype) of the device buffers.  ``fetch(step, host)`` queues the copies of
    ``host`` (dict name -> pinned CPU tensor) for ``step``; ``get(step)`` makes the current stream wait for them and
    returns the device tensors; ``done(step)`` (after the last kernel that reads them has been queued) frees the
    buffer for step + 2.  Typical loop: ``fetch(0)``, then per step ``fetch(i + 1)``, ``get(i)``, ..., ``done(i)``."""

    def __init__(self, slots, device, depth=2):
        self.depth = depth
        self.dev = [{k: torch.empty(shape, dtype=dtype, device=device) for k, (shape, dtype) in slots.items()}
                    for _ in range(depth)]
        self.stream = torch.cuda.Stream(device=device)
        self.ready = [torch.cuda.Event() for _ in range(depth)]
        self.free = [torch.cuda.Event() for _ in range(depth)]
        for ev in self.free:
            ev.record()
        self.queued = [None] * depth
        self.bytes_per_step = 0

    def fetch(self, step, host):
        b = step % self.depth
        if self.queued[b] == step:
            return
        self.stream.wait_event(self.free[b])
        n = 0
        with torch.cuda.stream(self.stream):
            for k, t in host.items():
                if not t.is_pinned():
                    raise ValueError(f"{k}: host tensors must be pinned (pageable memory makes the copy synchronous)")
                self.dev[b][k].copy_(t, non_blocking=True)
                n += t.numel() * t.element_size()
            self.ready[b].record(self.stream)
        self.queued[b] = step
        self.bytes_per_step = n

    def get(self, step):
        b = step % self.depth
        if self.queued[b] != step:
            raise RuntimeError(f"step {step} was never fetched")
        torch.cuda.current_stream().wait_event(self.ready[b])
        return self.dev[b]

    def done(self, step):
        self.free[step % self.depth].record()


CAMERA_FLOATS = 35   # world_view_transform 16, full_proj_transform 16, camera_center 3


def pack_camera(cam):
    """The three camera tensors a render needs, as one pinned 35-float block (one copy instead of four)."""
    return torch.cat([cam.world_view_transform.flatten().cpu(), cam.full_proj_transform.flatten().cpu(),
                      cam.camera_center.flatten().cpu()]).float().pin_memory()


def camera_view(cam, block):
    """A camera object whose matrices are views into a device copy of ``pack_camera(cam)``."""
    return SimpleNamespace(image_width=cam.image_width, image_height=cam.image_height, FoVx=cam.FoVx, FoVy=cam.FoVy,
                           world_view_transform=block[:16].view(4, 4), full_proj_transform=block[16:32].view(4, 4),
                           camera_center=block[32:35], image_name=getattr(cam, "image_name", None))


class CutAhead:
    """The cut of step i+1 on a side stream while step i's backward runs.

    ``launch(step, cam, cam_block_host, limit)`` copies the camera block and computes the cut into buffer step % 2;
    ``get(step)`` makes the current stream wait and returns ``(camera_view, CutBuffers)``; ``done(step)`` marks the
    buffer reusable once the render that reads it has been back-propagated (the backward reads the index tensors)."""

    def __init__(self, gh, nodes, boxes, n_store, device):
        self.gh, self.nodes, self.boxes = gh, nodes, boxes
        self.cuts = [CutBuffers(n_store, device) for _ in range(2)]
        self.cam_dev = [torch.empty(CAMERA_FLOATS, device=device) for _ in range(2)]
        self.stream = torch.cuda.Stream(device=device)
        self.ready = [torch.cuda.Event() for _ in range(2)]
        self.free = [torch.cuda.Event() for _ in range(2)]
        for ev in self.free:
            ev.record()
        self.queued = [None, None]
        self.cams = [None, None]

    def launch(self, step, cam, cam_block_host, limit):
        b = step % 2
        if self.queued[b] == step:
            return
        self.stream.wait_event(self.free[b])
        with torch.cuda.stream(self.stream):
            self.cam_dev[b].copy_(cam_block_host, non_blocking=True)
            view = camera_view(cam, self.cam_dev[b])
            self.cuts[b].compute(self.gh, self.nodes, self.boxes, view.camera_center, limit,
                                 camera_center_host=cam_block_host[32:35])
            self.ready[b].record(self.stream)
        self.queued[b], self.cams[b] = step, view

    def get(self, step):
        b = step % 2
        if self.queued[b] != step:
            raise RuntimeError(f"the cut of step {step} was never launched")
        torch.cuda.current_stream().wait_event(self.ready[b])
        return self.cams[b], self.cuts[b]

    def done(self, step):
        self.free[step % 2].record()


class LossReadback:
    """``push(loss)`` queues an asynchronous copy of the scalar into pinned memory; ``pop_ready()`` returns the values
    of the steps whose copies have completed so far WITHOUT blocking on the newest one; ``drain()`` waits for all."""

    def __init__(self, depth=4):
        self.host = [torch.zeros(1).pin_memory() for _ in range(depth)]
        self.events = [torch.cuda.Event() for _ in range(depth)]
        self.pending = []   # (slot) in push order
        self.n = 0
        self.depth = depth
        self.values = []

    def push(self, loss):
        if len(self.pending) == self.depth:   # ring full: the oldest copy must finish first (it is depth steps old)
            self._take(block=True)
        slot = self.n % self.depth
        self.host[slot].copy_(loss.detach().reshape(1), non_blocking=True)
        self.events[slot].record()
        self.pending.append(slot)
        self.n += 1

    def _take(self, block):
        slot = self.pending[0]
        if block:
            self.events[slot].synchronize()
        elif not self.events[slot].query():
            return False
        self.values.append(float(self.host[slot]))
        self.pending.pop(0)
        return True

    def pop_ready(self, keep_newest=1):
        """Collect finished values, never waiting for the `keep_newest` most recent pushes."""
        while len(self.pending) > keep_newest:
            self._take(block=True)
        out, self.values = self.values, []
        return out

    def drain(self):
        while self.pending:
            self._take(block=True)
        out, self.values = self.values, []
        return out
