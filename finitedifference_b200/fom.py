"""Batched full-order solver on one B200: host side of the FOM path.

``FomSolver`` owns the device workspace of ``bfd_fom_*`` (include/burgers_b200.h) for a batch
of parameter samples and mirrors what ``inviscid_burgers_implicit2D``
(reference hypernet2D.py:72-131) does for one.  PyTorch is used for device memory and streams
only; all arithmetic happens in the hand-written kernels of csrc/fom.cu.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _lib
from .discretisation import is_uniform, parameter_terms, stencil_coefficients


def _require_cuda(device=None) -> torch.device:
    if not torch.cuda.is_available():
        raise _lib.BurgersError("finitedifference_b200 needs a CUDA device (sm_100a); there is no CPU path")
    dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    if dev.type != "cuda":
        raise _lib.BurgersError(f"device must be a CUDA device, got {dev}")
    return dev


def fresh_host_array(shape) -> np.ndarray:
    """A new float64 array for a result the caller keeps, with transparent huge pages requested
    (``madvise(MADV_HUGEPAGE)``): first-touching 4.5 GB in 4 KB pages costs more than computing the
    trajectory (measured: 2.1 s per (2N, T+1) array at 750^2, T = 500; pinning it costs 3.3 s)."""
    a = np.empty(shape, dtype=np.float64)
    try:
        libc = C.CDLL("libc.so.6", use_errno=True)
        huge = 1 << 21
        lo = (a.ctypes.data + huge - 1) & ~(huge - 1)
        hi = (a.ctypes.data + a.nbytes) & ~(huge - 1)
        if hi > lo:
            libc.madvise(C.c_void_p(lo), C.c_size_t(hi - lo), 14)      # MADV_HUGEPAGE
    except OSError:
        pass
    return a


def _dev(a: np.ndarray, device) -> torch.Tensor:
    return torch.from_numpy(np.ascontiguousarray(a, dtype=np.float64)).to(device, non_blocking=False)


class FomSolver:
    """Advance ``len(mus)`` samples ``num_steps`` trapezoidal steps with Newton (rel. tol 1e-12).

    Snapshots stay resident in HBM (skewed strip layout, csrc/common.cuh) until exported.
    """

    def __init__(self, grid_x, grid_y, dt, num_steps, mus, device=None, max_its=100, rel_tol=1e-12,
                 strict=True, chunk_steps=0, newton_hint=0, kernel="chunk"):
        self.lib = _lib.get_lib()
        self.device = _require_cuda(device)
        self.grid_x = np.asarray(grid_x, dtype=np.float64)
        self.grid_y = np.asarray(grid_y, dtype=np.float64)
        self.mus = np.atleast_2d(np.asarray(mus, dtype=np.float64))
        self.nx, self.ny = self.grid_x.size - 1, self.grid_y.size - 1
        self.batch, self.num_steps, self.dt = self.mus.shape[0], int(num_steps), float(dt)
        cx, cxl, cy, cyd = stencil_coefficients(self.grid_x, self.grid_y, self.dt)
        self.uniform_y = is_uniform(cy, cyd)
        self.desc = _lib.fom_desc(self.nx, self.ny, self.batch, self.num_steps, max_its, rel_tol,
                                  chunk_steps, newton_hint, strict, self.uniform_y,
                                  kernel={"chunk": 0, "tile": 1}[kernel])
        nbytes = self.lib.bfd_fom_workspace_bytes(C.byref(self.desc))
        if nbytes == 0:
            raise _lib.BurgersError("invalid FOM description")
        self.workspace_bytes = nbytes
        with torch.cuda.device(self.device):
            self.ws = torch.empty(nbytes, dtype=torch.uint8, device=self.device)
            src, lbc = parameter_terms(self.grid_x, self.grid_y, self.dt, self.mus)
            self._coef_t = [_dev(a, self.device) for a in (cx, cxl, cy, cyd, src, lbc)]
            self.coef = _lib.FomCoef(*[t.data_ptr() for t in self._coef_t])
            self._pinned = torch.zeros(16, dtype=torch.int32).pin_memory()
        self.launches = 0
        self.steps_redone = 0
        self._ready = False

    # -- helpers ---------------------------------------------------------------------------
    def _stream(self):
        return torch.cuda.current_stream(self.device).cuda_stream

    def _check(self, rc, what):
        _lib.check(rc, self.lib, what)

    # -- API -------------------------------------------------------------------------------
    def set_parameters(self, mus) -> None:
        """Replace the parameter samples (same batch size): the workspace and the grid coefficients
        are kept, only the per-sample source / inlet terms are recomputed and uploaded."""
        mus = np.atleast_2d(np.asarray(mus, dtype=np.float64))
        if mus.shape[0] != self.batch:
            raise ValueError(f"{mus.shape[0]} samples given, the solver holds {self.batch}")
        self.mus = mus
        src, lbc = parameter_terms(self.grid_x, self.grid_y, self.dt, self.mus)
        with torch.cuda.device(self.device):
            st = torch.cuda.current_stream(self.device)
            for t, a in ((self._coef_t[4], src), (self._coef_t[5], lbc)):
                h = torch.from_numpy(np.ascontiguousarray(a, dtype=np.float64))
                t.copy_(h.to(self.device, non_blocking=False).view_as(t))
            st.synchronize()
        self._ready = False

    def set_initial(self, w0) -> None:
        """``w0``: (2N,) broadcast to every sample, or (batch, 2N); NumPy or CUDA tensor."""
        n2 = 2 * self.nx * self.ny
        with torch.cuda.device(self.device):
            if isinstance(w0, torch.Tensor):
                w = w0.to(self.device, dtype=torch.float64).contiguous()
            else:
                w = _dev(np.asarray(w0, dtype=np.float64), self.device)
            if w.numel() == n2:
                stride = 0
            elif w.numel() == n2 * self.batch:
                stride = n2
            else:
                raise ValueError(f"w0 has {w.numel()} entries, expected {n2} or {self.batch}x{n2}")
            st = self._stream()
            self._check(self.lib.bfd_fom_init(C.byref(self.desc), self.ws.data_ptr(),
                                              self.workspace_bytes, st), "bfd_fom_init")
            self._check(self.lib.bfd_fom_set_initial(C.byref(self.desc), self.ws.data_ptr(),
                                                     w.data_ptr(), stride, st), "bfd_fom_set_initial")
            self._w0_keepalive = w
        self._ready = True

    def run(self) -> None:
        """Run all time steps of all samples; returns when the device has finished."""
        if not self._ready:
            raise _lib.BurgersError("set_initial() first")
        with torch.cuda.device(self.device):
            rc = self.lib.bfd_fom_run(C.byref(self.desc), self.ws.data_ptr(), C.byref(self.coef),
                                      self._pinned.data_ptr(), self._stream())
        self.launches = int(self._pinned[2])
        self.steps_redone = int(self._pinned[3])
        self._check(rc, "bfd_fom_run")
        self._ready = False

    def snapshots_device(self, m: int, t0: int = 0, t1: int | None = None, layout: int = 1,
                         out: torch.Tensor | None = None) -> torch.Tensor:
        """Snapshots of sample ``m`` as a CUDA tensor: layout 1 -> (2N, t1-t0) as the reference
        returns them, layout 0 -> (t1-t0, 2N)."""
        t1 = self.num_steps + 1 if t1 is None else t1
        n2, nt = 2 * self.nx * self.ny, t1 - t0
        shape = (n2, nt) if layout == 1 else (nt, n2)
        with torch.cuda.device(self.device):
            if out is None:
                out = torch.empty(shape, dtype=torch.float64, device=self.device)
            self._check(self.lib.bfd_fom_export(C.byref(self.desc), self.ws.data_ptr(), int(m), int(t0),
                                                int(t1), int(layout), nt, out.data_ptr(),
                                                self._stream()), "bfd_fom_export")
        return out

    _PIN_LOCK = __import__("threading").Lock()
    _PINS = {}      # (device, rows, cols) -> two small pinned staging buffers, shared by all solvers of the process

    def _staged_to_host(self, dev: torch.Tensor, host: np.ndarray, chunk_bytes: int = 64 << 20) -> None:
        """Device-to-host through two small pinned buffers: the DMA of chunk i overlaps the (multi-threaded)
        CPU copy of chunk i-1 into the caller's pageable array, instead of the driver's serial bounce copy."""
        rows, cols = dev.shape
        step = max(1, min(rows, chunk_bytes // (8 * cols)))
        key = (str(self.device), step, cols)
        pins = FomSolver._PINS.get(key)
        if pins is None:
            pins = FomSolver._PINS[key] = [torch.empty((step, cols), dtype=torch.float64).pin_memory() for _ in range(2)]
        events = [torch.cuda.Event(), torch.cuda.Event()]
        out = torch.from_numpy(host)
        FomSolver._PIN_LOCK.acquire()           # the staging buffers are shared by the solvers of the process
        try:
            self._staged_chunks(dev, out, pins, events, step, rows)
        finally:
            FomSolver._PIN_LOCK.release()

    def _staged_chunks(self, dev, out, pins, events, step, rows):
        bounds = [(lo, min(lo + step, rows)) for lo in range(0, rows, step)]
        with torch.cuda.device(self.device):
            for i in range(len(bounds) + 1):
                if i < len(bounds):
                    lo, hi = bounds[i]
                    pins[i % 2][:hi - lo].copy_(dev[lo:hi], non_blocking=True)
                    events[i % 2].record()
                if i > 0:
                    lo, hi = bounds[i - 1]
                    events[(i - 1) % 2].synchronize()
                    out[lo:hi].copy_(pins[(i - 1) % 2][:hi - lo])

    def snapshots(self, m: int, pinned: bool = False, hugepages: bool = True, staged: bool = True) -> np.ndarray:
        """(2N, T+1) float64 C-order NumPy array, the reference's return value (:89, :126).

        A fresh array every call (the caller keeps it).  Default: pageable memory with huge pages
        requested, filled through two small pinned staging buffers (0.21 s for the 4.5 GB array of the
        fine grid; a plain pageable copy takes 1.2 s, pinning a fresh buffer 3.3-3.7 s:
        ``profiles/r01_host_array_probe.json``).  Sweeps that recycle pinned buffers go through
        ``sweep.SnapshotSweep``."""
        dev = self.snapshots_device(m, layout=1)
        if pinned:
            host = torch.empty(dev.shape, dtype=torch.float64, pin_memory=True)
        else:
            host = torch.from_numpy(fresh_host_array(tuple(dev.shape)) if hugepages
                                    else np.empty(tuple(dev.shape), dtype=np.float64))
        if staged and not pinned:
            self._staged_to_host(dev, host.numpy())
        else:
            host.copy_(dev, non_blocking=False)
        return host.numpy()

    def diagnostics(self):
        """(its, relnorm): per sample and step, len(resnorms) of newton_raphson and the last
        relative residual."""
        with torch.cuda.device(self.device):
            its = torch.empty((self.batch, self.num_steps), dtype=torch.int32, device=self.device)
            rel = torch.empty((self.batch, self.num_steps), dtype=torch.float64, device=self.device)
            self._check(self.lib.bfd_fom_diagnostics(C.byref(self.desc), self.ws.data_ptr(), its.data_ptr(),
                                                     rel.data_ptr(), self._stream()), "bfd_fom_diagnostics")
            return its.cpu().numpy(), rel.cpu().numpy()
