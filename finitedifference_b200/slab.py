"""One grid over several GPUs: slab decomposition of the FOM along x (second row;
BASELINE.json config 5, the synthetic 8192 x 8192 single-mu run).

The reference has no counterpart (it is a single-process NumPy code); what is mirrored is the
result: the snapshots of ``inviscid_burgers_implicit2D`` (hypernet2D.py:72-131), here held as
column blocks, one per rank.

Each rank owns a contiguous block of the grid's 32-column strips and runs the same fused sweep
kernel as the single-GPU path over them.  Upwinding makes the coupling one-directional: the last
strip of rank g writes its right-edge hand-off records straight into rank g+1's halo buffer with
peer stores over NVLink while it computes (include/burgers_b200.h, ``bfd_fom_slab``), so there is
no exchange phase and no collective; ``torch.distributed`` is used to trade the 64-byte IPC
handles and for barriers around initialisation.
"""
from __future__ import annotations

import ctypes as C
import threading

import numpy as np
import torch

from . import _lib
from .discretisation import is_uniform, parameter_terms, stencil_coefficients
from .fom import _require_cuda

STRIP = 32


def slab_partition(nx: int, world: int):
    """``[(strip_begin, strip_end)]`` per rank: contiguous, sizes differing by at most one strip,
    the larger blocks first.  Needs at least one strip per rank."""
    ns = (int(nx) + STRIP - 1) // STRIP
    if world < 1 or ns < world:
        raise ValueError(f"{nx} columns make {ns} strips: cannot split over {world} ranks")
    base, extra = divmod(ns, world)
    out, lo = [], 0
    for r in range(world):
        hi = lo + base + (1 if r < extra else 0)
        out.append((lo, hi))
        lo = hi
    return out


def slab_columns(nx: int, world: int):
    """``[(col_begin, col_end)]`` per rank."""
    return [(b * STRIP, min(e * STRIP, int(nx))) for b, e in slab_partition(nx, world)]


class FomSlab:
    """This rank's slab of one FOM trajectory batch (usually a single mu)."""

    def __init__(self, grid_x, grid_y, dt, num_steps, mus, rank, world, device=None, max_its=100, rel_tol=1e-12,
                 strict=True, chunk_steps=0, newton_hint=0, lib=None, host_memory=False, kernel="chunk"):
        self.lib = lib or _lib.get_lib()
        self.host_memory = host_memory                      # emulator tests: NumPy buffers, no CUDA
        self.device = torch.device("cpu") if host_memory else _require_cuda(device)
        gx, gy = np.asarray(grid_x, dtype=np.float64), np.asarray(grid_y, dtype=np.float64)
        self.mus = np.atleast_2d(np.asarray(mus, dtype=np.float64))
        self.nx, self.ny = gx.size - 1, gy.size - 1
        self.rank, self.world = int(rank), int(world)
        self.batch, self.num_steps = self.mus.shape[0], int(num_steps)
        self.strips = slab_partition(self.nx, world)[rank]
        self.c0, self.c1 = slab_columns(self.nx, world)[rank]
        self.nx_local = self.c1 - self.c0
        cx, cxl, cy, cyd = stencil_coefficients(gx, gy, float(dt))
        src, lbc = parameter_terms(gx, gy, float(dt), self.mus)
        self.desc = _lib.fom_desc(self.nx, self.ny, self.batch, self.num_steps, max_its, rel_tol, chunk_steps,
                                  newton_hint, strict, is_uniform(cy, cyd), kernel={"chunk": 0, "tile": 1}[kernel])
        self.slab = _lib.FomSlab(self.rank, self.world, self.strips[0], self.strips[1], None, None)
        self.nbytes = self.lib.bfd_fom_slab_workspace_bytes(C.byref(self.desc), C.byref(self.slab))
        if self.nbytes == 0:
            raise _lib.BurgersError("invalid slab description")
        self.halo_offset = self.lib.bfd_fom_slab_halo_offset(C.byref(self.desc), C.byref(self.slab))
        p = C.c_void_p()
        with self._dev():
            _lib.check(self.lib.bfd_mem_alloc(self.nbytes, C.byref(p)), self.lib, "bfd_mem_alloc")
        self.ws = p.value
        self._coef_t = [self._to_dev(a) for a in (cx[self.c0:self.c1], cxl[self.c0:self.c1], cy, cyd,
                                                   src[:, self.c0:self.c1], lbc)]
        self.coef = _lib.FomCoef(*[self._ptr(t) for t in self._coef_t])
        self._pinned = torch.zeros(16, dtype=torch.int32)
        if not host_memory:
            self._pinned = self._pinned.pin_memory()
        self._mapped = []
        self.launches = self.steps_redone = 0

    # -- plumbing -------------------------------------------------------------------------
    def _dev(self):
        return torch.cuda.device(self.device) if not self.host_memory else _Null()

    def _to_dev(self, a):
        return torch.from_numpy(np.ascontiguousarray(a, dtype=np.float64)).to(self.device)

    @staticmethod
    def _ptr(t):
        return t.data_ptr()

    def _stream(self):
        return None if self.host_memory else torch.cuda.current_stream(self.device).cuda_stream

    def _check(self, rc, what):
        _lib.check(rc, self.lib, what)

    def ipc_handle(self) -> bytes:
        buf = C.create_string_buffer(_lib.IPC_HANDLE_BYTES)
        with self._dev():
            self._check(self.lib.bfd_ipc_export(self.ws, buf), "bfd_ipc_export")
        return buf.raw

    def _open(self, handle: bytes) -> int:
        p = C.c_void_p()
        with self._dev():
            self._check(self.lib.bfd_ipc_open(handle, C.byref(p)), "bfd_ipc_open")
        self._mapped.append(p.value)
        return p.value

    def connect(self, home=None, next_ws=None, next_halo_offset=0, handles=False):
        """Tell the slab where rank 0's workspace and rank+1's halo buffer are: raw addresses valid
        in this process, or (``handles=True``) IPC handles of the other ranks' workspaces."""
        if self.rank > 0:
            self.slab.home = self._open(home) if handles else home
        else:
            self.slab.home = self.ws
        if self.rank < self.world - 1:
            base = self._open(next_ws) if handles else next_ws
            self.slab.next_halo = base + next_halo_offset

    # -- the FomSolver interface, on the slab's columns ---------------------------------------
    def init(self):
        with self._dev():
            self._check(self.lib.bfd_fom_slab_init(C.byref(self.desc), C.byref(self.slab), self.ws, self.nbytes,
                                                   self._stream()), "bfd_fom_slab_init")
            if not self.host_memory:
                torch.cuda.current_stream(self.device).synchronize()

    def set_initial(self, w0_local):
        """``w0_local``: this slab's columns in the natural layout ``(2, ny, nx_local)`` (one state for
        every sample)."""
        w0 = w0_local if isinstance(w0_local, torch.Tensor) else self._to_dev(np.asarray(w0_local).ravel())
        w0 = w0.to(self.device, dtype=torch.float64).contiguous().reshape(-1)
        if w0.numel() != 2 * self.ny * self.nx_local:
            raise ValueError(f"w0_local must hold {2 * self.ny * self.nx_local} values")
        with self._dev():
            self._check(self.lib.bfd_fom_slab_set_initial(C.byref(self.desc), C.byref(self.slab), self.ws,
                                                          w0.data_ptr(), 0, self._stream()), "bfd_fom_slab_set_initial")
            if not self.host_memory:
                torch.cuda.current_stream(self.device).synchronize()

    def run(self):
        with self._dev():
            rc = self.lib.bfd_fom_slab_run(C.byref(self.desc), C.byref(self.slab), self.ws, C.byref(self.coef),
                                           self._pinned.data_ptr(), self._stream())
        self.launches, self.steps_redone = int(self._pinned[2]), int(self._pinned[3])
        self._check(rc, "bfd_fom_slab_run")

    def snapshots_device(self, m=0, t0=0, t1=None, layout=0):
        """Snapshots of this slab's columns: layout 0 -> ``(t1-t0, 2, ny, nx_local)``, layout 1 ->
        ``(2*ny*nx_local, t1-t0)``."""
        t1 = self.num_steps + 1 if t1 is None else t1
        nt, n2 = t1 - t0, 2 * self.ny * self.nx_local
        out = torch.empty((n2, nt) if layout == 1 else (nt, 2, self.ny, self.nx_local), dtype=torch.float64,
                          device=self.device)
        with self._dev():
            self._check(self.lib.bfd_fom_slab_export(C.byref(self.desc), C.byref(self.slab), self.ws, int(m), int(t0),
                                                     int(t1), int(layout), nt, out.data_ptr(), self._stream()),
                        "bfd_fom_slab_export")
        return out

    def diagnostics(self):
        its = torch.zeros((self.batch, self.num_steps), dtype=torch.int32, device=self.device)
        rel = torch.zeros((self.batch, self.num_steps), dtype=torch.float64, device=self.device)
        with self._dev():
            self._check(self.lib.bfd_fom_slab_diagnostics(C.byref(self.desc), C.byref(self.slab), self.ws,
                                                          its.data_ptr(), rel.data_ptr(), self._stream()),
                        "bfd_fom_slab_diagnostics")
        return its.cpu().numpy(), rel.cpu().numpy()

    def close(self):
        with self._dev():
            if not self.host_memory:
                torch.cuda.synchronize(self.device)
            for p in self._mapped:
                self.lib.bfd_ipc_close(p)
            self._mapped = []
            if self.ws:
                self.lib.bfd_mem_free(self.ws)
                self.ws = None


class _Null:
    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False


def connect_distributed(slab: FomSlab):
    """Trade IPC handles over the default process group (one process per GPU) and wire the slab up."""
    import torch.distributed as dist
    mine = (slab.ipc_handle(), slab.halo_offset)
    everyone = [None] * slab.world
    dist.all_gather_object(everyone, mine)
    nxt = everyone[slab.rank + 1] if slab.rank < slab.world - 1 else (None, 0)
    slab.connect(home=everyone[0][0], next_ws=nxt[0], next_halo_offset=nxt[1], handles=True)


class SlabGroup:
    """All slabs of a run inside ONE process (streams of one device, or one device per slab): used by
    the tests and by single-process callers; every slab's blocking ``run`` goes on its own thread."""

    def __init__(self, grid_x, grid_y, dt, num_steps, mus, world, devices=None, **kw):
        devices = devices if devices is not None else [None] * world
        self.slabs = [FomSlab(grid_x, grid_y, dt, num_steps, mus, r, world, device=devices[r], **kw)
                      for r in range(world)]
        for r, s in enumerate(self.slabs):
            nxt = self.slabs[r + 1] if r + 1 < world else None
            s.connect(home=self.slabs[0].ws, next_ws=nxt.ws if nxt else None,
                      next_halo_offset=nxt.halo_offset if nxt else 0)
        self.nx, self.ny = self.slabs[0].nx, self.slabs[0].ny
        self._streams = None
        if not self.slabs[0].host_memory:
            self._streams = [torch.cuda.Stream(device=s.device) for s in self.slabs]

    def _each(self, fn, threads=False):
        errs = []

        def call(r, s):
            try:
                if self._streams is not None:
                    with torch.cuda.device(s.device), torch.cuda.stream(self._streams[r]):
                        fn(s)
                else:
                    fn(s)
            except Exception as e:                      # noqa: BLE001 - re-raised below
                errs.append(e)

        if threads:
            ts = [threading.Thread(target=call, args=(r, s)) for r, s in enumerate(self.slabs)]
            [t.start() for t in ts]
            [t.join() for t in ts]
        else:
            for r, s in enumerate(self.slabs):
                call(r, s)
        if errs:
            raise errs[0]

    def set_initial(self, w0):
        w0 = np.asarray(w0, dtype=np.float64).reshape(2, self.ny, self.nx)
        self._each(lambda s: s.init())
        self._each(lambda s: s.set_initial(np.ascontiguousarray(w0[:, :, s.c0:s.c1])))

    def run(self):
        self._each(lambda s: s.run(), threads=True)

    def snapshots(self, m=0):
        """``(2N, T+1)`` NumPy array of the whole grid, assembled from the slabs."""
        T1 = self.slabs[0].num_steps + 1
        out = np.empty((T1, 2, self.ny, self.nx))
        box = {}
        self._each(lambda s: box.__setitem__(s.rank, s.snapshots_device(m).cpu().numpy()))
        for s in self.slabs:
            out[:, :, :, s.c0:s.c1] = box[s.rank]
        return np.ascontiguousarray(out.reshape(T1, -1).T)

    def diagnostics(self):
        return self.slabs[0].diagnostics()

    def close(self):
        for s in self.slabs:
            s.close()
