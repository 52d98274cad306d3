"""Screen-space tile partition over the GPUs of one box (one process per GPU, torch.distributed for the plumbing).

The reference renders one frame in one address space (Renderer::draw, raytracing.cpp:67-102; rows are handed to
threads dynamically, :77-80).  Here the frame is cut into bands of `band_rows` memory rows dealt round-robin to the
ranks (hit density is concentrated mid-screen, so contiguous slabs would load-imbalance).  The scene is replicated
(it is read-only and small); there is no exchange during rendering.  The only data movement is the framebuffer
reaching the root rank, in one of two ways:

  * "p2p"  (default, fused): the root allocates the framebuffer, exports it with CUDA IPC, every rank maps it and
            its render kernel stores its tiles straight into the root's memory over NVLink / NVSwitch;
  * "nccl" (baseline): every rank renders into a local framebuffer and ships its bands with NCCL send/recv.

The pure-host parts (partition arithmetic, handle exchange, band packing) run on CPU with the gloo backend too:
tests/test_multigpu_host.py.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib as L

DEFAULT_BAND_ROWS = 8  # = the block height of the tiles kernels; 135 bands at 1080p balance 8 ranks to 6 %


def band_rows_of(height: int, band_rows: int, band_count: int, band_index: int) -> np.ndarray:
    """Memory rows of the frame that partition (band_rows, band_count, band_index) renders, ascending.
    Mirrors rt_frame's rule: row r belongs to index (r // band_rows) % band_count."""
    r = np.arange(height)
    if band_count <= 1:
        return r
    return r[(r // band_rows) % band_count == band_index]


def partition_is_exact(height: int, band_rows: int, band_count: int) -> bool:
    """Every row is rendered by exactly one rank."""
    seen = np.zeros(height, np.int32)
    for i in range(band_count):
        seen[band_rows_of(height, band_rows, band_count, i)] += 1
    return bool((seen == 1).all())


def exchange_bytes(payload: bytes | None, src: int, group=None) -> bytes:
    """Broadcast a small opaque byte string (an IPC handle) from rank `src` to every rank."""
    import torch
    import torch.distributed as dist
    n = L_IPC_BYTES
    backend = dist.get_backend(group)
    dev = torch.device("cuda", torch.cuda.current_device()) if backend == "nccl" else torch.device("cpu")
    buf = torch.zeros(n, dtype=torch.uint8, device=dev)
    if dist.get_rank(group) == src:
        assert payload is not None and len(payload) == n
        buf.copy_(torch.frombuffer(bytearray(payload), dtype=torch.uint8))
    dist.broadcast(buf, src=src, group=group)
    return bytes(buf.cpu().numpy().tobytes())


L_IPC_BYTES = 64  # RT_IPC_HANDLE_BYTES


class SharedFramebuffer:
    """colour (u32) + depth (f32) of a W x H frame living on the root rank's GPU, mapped into every rank."""

    def __init__(self, width: int, height: int, root: int = 0, group=None):
        import torch.distributed as dist
        self.W, self.H, self.root = width, height, root
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.bytes = 4 * width * height
        lib = L.lib()
        self.color = C.c_void_p()
        self.t = C.c_void_p()
        self._owned = self.rank == root
        if self._owned:
            L.check(lib.rt_device_malloc(self.bytes, C.byref(self.color)))
            L.check(lib.rt_device_malloc(self.bytes, C.byref(self.t)))
        if self.world > 1:
            handles = []
            for p in (self.color, self.t):
                h = None
                if self._owned:
                    raw = C.create_string_buffer(L_IPC_BYTES)
                    L.check(lib.rt_ipc_export(p, raw))
                    h = raw.raw
                handles.append(exchange_bytes(h, root, group))
            if not self._owned:
                L.check(lib.rt_ipc_open(handles[0], C.byref(self.color)))
                L.check(lib.rt_ipc_open(handles[1], C.byref(self.t)))

    def download(self, stream: int = 0):
        """Root only: (color [H,W] u32, t [H,W] f32) on the host."""
        assert self._owned
        c = np.empty((self.H, self.W), np.uint32)
        t = np.empty((self.H, self.W), np.float32)
        st = C.c_void_p(stream) if stream else None
        L.check(L.lib().rt_memcpy_d2h(C.c_void_p(c.ctypes.data), self.color, self.bytes, st))
        L.check(L.lib().rt_memcpy_d2h(C.c_void_p(t.ctypes.data), self.t, self.bytes, st))
        return c, t

    def close(self):
        lib = L.lib()
        for p in (self.color, self.t):
            if p and p.value:
                (lib.rt_device_free if self._owned else lib.rt_ipc_close)(p)
        self.color, self.t = C.c_void_p(), C.c_void_p()


def pack_bands(image: "np.ndarray", rows: np.ndarray):
    """Rows of `image` [H, W] owned by one rank as one contiguous block (what the NCCL gather ships)."""
    return np.ascontiguousarray(image[rows])


def unpack_bands(image: "np.ndarray", rows: np.ndarray, block: "np.ndarray"):
    image[rows] = block
    return image
