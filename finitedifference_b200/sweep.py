"""Snapshot generation for many parameter samples with the results on the HOST.

The reference loops ``load_or_compute_snaps`` over the training parameters serially
(run_prom.py:59-71, train_autoencoder.py:63-72): one 4.5 GB ``(2N, T+1)`` array per mu on the fine
grid.  On a B200 the trajectories themselves take 7.5 ms per mu (24 advancing together), so what
a sweep with host output costs is the PCIe link: 85 ms per mu at 53 GB/s.  ``SnapshotSweep`` keeps
that link busy from the first finished group on:

* ``workers`` host threads, each with its own CUDA stream, ``FomSolver`` (``group`` samples) and
  device staging buffers; a worker advances its group, exports it to the reference's
  ``(2N, T+1)`` C-order layout on the device and copies it to pinned host memory, while the other
  workers are in other phases -- copies of one group overlap the sweeps of the next;
* results land in caller-owned host buffers (``out``) or in a small pool of pinned buffers that
  are handed to ``sink(index, mu, snaps)`` and recycled when the sink returns (e.g. an ``.npy``
  writer: ``save_sink``), so a sweep of any length needs ``pool`` x 4.5 GB of pinned memory.

There is no collective and no CPU path: samples are independent, every
arithmetic operation happens in csrc/fom.cu.
"""
from __future__ import annotations

import os
import queue
import threading
import time

import numpy as np
import torch

from . import _lib
from .fom import FomSolver, _require_cuda


def _parse_cpulist(text):
    cpus = set()
    for part in text.strip().split(","):
        lo, _, hi = part.partition("-")
        if lo.strip().isdigit():
            cpus.update(range(int(lo), int(hi or lo) + 1))
    return cpus


def _topo_affinity(index):
    """(cpus, numa node) of GPU `index` from NVML (what ``nvidia-smi topo -m`` prints as CPU / NUMA affinity);
    (set(), None) when the driver does not know either."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(int(index))
        words = (os.cpu_count() + 63) // 64
        mask = pynvml.nvmlDeviceGetCpuAffinity(h, words)
        cpus = {64 * w + b for w, m in enumerate(mask) for b in range(64) if (m >> b) & 1}
        try:
            node = pynvml.nvmlDeviceGetNumaNodeId(h)
        except Exception:                                  # noqa: BLE001 - older NVML
            node = None
        if len(cpus) >= (os.cpu_count() or 0):             # "all CPUs" carries no information
            return set(), node
        return cpus, (node if node is not None else -1)
    except Exception:                                      # noqa: BLE001 - NVML missing or refusing: no binding
        return set(), None


def bind_host_to_gpu_node(device=None):
    """Restrict this process to the CPU cores of the NUMA node the GPU's PCIe root hangs off, so
    that pinned buffers (first touch) and the worker threads live next to the link they feed.
    With several ranks copying 48 GB/s each, buffers on the wrong socket halve the rate.
    Returns the node number, or None when the topology is not exposed (nothing is changed)."""
    try:
        dev = _require_cuda(device)
        bus = torch.cuda.get_device_properties(dev).pci_bus_id
        dom = torch.cuda.get_device_properties(dev).pci_domain_id
        devid = torch.cuda.get_device_properties(dev).pci_device_id
        path = f"/sys/bus/pci/devices/{dom:04x}:{bus:02x}:{devid:02x}.0/numa_node"
        node = int(open(path).read().strip())
        cpus = set()
        if node < 0:
            # virtual machines often report no node for the device: ask the driver for the GPU's CPU affinity
            cpus, node = _topo_affinity(dev.index if dev.index is not None else torch.cuda.current_device())
            if not cpus:
                return None
        else:
            for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
                lo, _, hi = part.partition("-")
                cpus.update(range(int(lo), int(hi or lo) + 1))
        cpus &= os.sched_getaffinity(0)
        if not cpus:
            return None
        os.sched_setaffinity(0, cpus)
        return node
    except (OSError, ValueError, AttributeError):
        return None


def save_sink(snap_folder="param_snaps", name_fn=None, dtype=None, time_stride=1):
    """A sink that writes ``param_snaps/mu1_<..>+mu2_<..>.npy`` files interchangeable with the
    reference's cache (``load_or_compute_snaps``, hypernet2D.py:3187-3221): by default the same
    ``(2N, T+1)`` float64 C-order bytes ``np.save`` gives the reference.

    Two storage options the reference does not have, for sweeps whose files are the
    bottleneck: ``dtype=np.float32`` halves them, ``time_stride=s`` keeps every s-th time level (columns
    ``0, s, 2s, ...``; the last level is kept as well when ``T`` is not a multiple of ``s``).  Such files carry
    a ``.f32`` / ``.every<s>`` tag before ``.npy`` so that the reference's loader never mistakes them for a
    full trajectory."""
    from .hypernet2D import param_to_snap_fn
    os.makedirs(snap_folder, exist_ok=True)
    time_stride = int(time_stride)
    if time_stride < 1:
        raise ValueError("time_stride must be >= 1")
    dt_ = None if dtype is None else np.dtype(dtype)
    if dt_ is not None and dt_ not in (np.dtype(np.float64), np.dtype(np.float32)):
        raise ValueError("dtype must be float64 or float32")

    def sink(index, mu, snaps):
        fn = name_fn(mu) if name_fn else param_to_snap_fn(list(mu), snap_folder=snap_folder)
        out = snaps
        tag = ""
        if time_stride > 1:
            cols = list(range(0, snaps.shape[1], time_stride))
            if cols[-1] != snaps.shape[1] - 1:
                cols.append(snaps.shape[1] - 1)
            out = snaps[:, cols]
            tag += f".every{time_stride}"
        if dt_ is not None and dt_ != out.dtype:
            out = out.astype(dt_)
            tag += ".f32"
        if tag:
            root, ext = os.path.splitext(fn)
            fn = root + tag + ext
        np.save(fn, np.ascontiguousarray(out))
    return sink


class SnapshotSweep:
    """FOM trajectories of a list of parameter samples, delivered to the host as they finish."""

    def __init__(self, grid_x, grid_y, dt, num_steps, device=None, group=2, workers=3, pool=None, strict=True,
                 **solver_kw):
        self.device = _require_cuda(device)
        self.grid_x, self.grid_y = np.asarray(grid_x, dtype=np.float64), np.asarray(grid_y, dtype=np.float64)
        self.dt, self.num_steps = float(dt), int(num_steps)
        self.group, self.workers = max(1, int(group)), max(1, int(workers))
        self.pool_size = int(pool) if pool else self.group * self.workers
        self.solver_kw = dict(strict=strict, **solver_kw)
        self.n2 = 2 * (self.grid_x.size - 1) * (self.grid_y.size - 1)
        self._pool = None
        self._pool_lock = threading.Lock()
        self._solvers = {}        # (worker, batch) -> FomSolver
        self._stage = {}          # worker -> list of device staging tensors
        self._streams = [torch.cuda.Stream(device=self.device) for _ in range(self.workers)]
        self._copy_stream = torch.cuda.Stream(device=self.device)
        self.copy_stats = None
        self.launches = 0
        self.steps_redone = 0

    # -- buffers -----------------------------------------------------------------------------
    def host_buffer(self):
        """One pinned ``(2N, T+1)`` float64 buffer (what ``out`` entries should be for full speed)."""
        return torch.empty((self.n2, self.num_steps + 1), dtype=torch.float64).pin_memory()

    def _get_pool(self):
        with self._pool_lock:
            if self._pool is None:
                pool = queue.Queue()
                for _ in range(self.pool_size):
                    pool.put(self.host_buffer())
                self._pool = pool
        return self._pool

    def _solver(self, wid, batch, mus):
        key = (wid, batch)
        s = self._solvers.get(key)
        if s is None:
            s = FomSolver(self.grid_x, self.grid_y, self.dt, self.num_steps, mus, device=self.device,
                          **self.solver_kw)
            self._solvers[key] = s
        else:
            s.set_parameters(mus)
        return s

    # -- the sweep ---------------------------------------------------------------------------
    def run(self, w0, mus, out=None, sink=None):
        """Advance every sample of ``mus`` from ``w0`` ((2N,) NumPy, host).

        ``out``: list of host buffers, one per sample (pinned torch tensors from ``host_buffer()`` or
        C-contiguous NumPy arrays), filled in place and returned.  Without ``out`` the trajectories
        go through the pinned pool to ``sink(index, mu, snaps_ndarray)``, called from the worker
        threads, the buffer being reused after the sink returns; with neither, fresh pinned
        buffers are allocated and returned as NumPy arrays (slow: pinning 4.5 GB per sample).
        Returns the list of host arrays (``out`` mode) or ``None`` (sink mode).
        """
        mus = [list(m)[:2] for m in mus]
        n = len(mus)
        if out is not None and len(out) != n:
            raise ValueError("one output buffer per parameter sample")
        if out is None and sink is None:
            out = [self.host_buffer() for _ in range(n)]
        w0 = np.ascontiguousarray(np.asarray(w0, dtype=np.float64).ravel())
        if w0.size != self.n2:
            raise ValueError(f"w0 has {w0.size} entries, expected {self.n2}")
        w0_pinned = torch.from_numpy(w0).pin_memory()
        groups = queue.Queue()
        for lo in range(0, n, self.group):
            groups.put(list(range(lo, min(lo + self.group, n))))
        errors = []
        stats = {"launches": 0, "redone": 0, "sweeps": 0.0, "stage_wait": 0.0, "pool_wait": 0.0}
        lock = threading.Lock()

        # Device-to-host copies go through ONE stream, issued back to back by a copier thread (a single copy
        # saturates the PCIe link; concurrent ones only contend for the host's ingest bandwidth, and a gap
        # between two copies is link time lost: with the copies issued by the workers under a lock the link
        # ran at 45 of the 57 GB/s a bare copy loop reaches on the same box, profiles/r02_d2h_probe_n1.json).
        # A delivery thread waits for each copy and hands the array to the sink, so a slow sink does not hold
        # the next copy back.
        copy_q, done_q = queue.Queue(), queue.Queue()
        copy_stream = self._copy_stream
        spans = []                                  # (begin, done, bytes) of every copy, on the copy stream

        def copier():
            try:
                with torch.cuda.device(self.device):
                    while True:
                        item = copy_q.get()
                        if item is None:
                            break
                        exported, src, dst, i, pooled, free = item
                        copy_stream.wait_event(exported)
                        with torch.cuda.stream(copy_stream):
                            begin = torch.cuda.Event(enable_timing=True)
                            begin.record(copy_stream)
                            dst.copy_(src, non_blocking=True)
                            done = torch.cuda.Event(enable_timing=True)
                            done.record(copy_stream)
                        spans.append((begin, done, src.numel() * src.element_size()))
                        done_q.put((done, dst, i, pooled, free))
            except BaseException as e:          # noqa: BLE001 - re-raised in the caller's thread
                errors.append(e)
            finally:
                done_q.put(None)

        def deliver():
            try:
                while True:
                    item = done_q.get()
                    if item is None:
                        break
                    done, dst, i, pooled, free = item
                    done.synchronize()
                    free.set()                      # the staging buffer may take the next export
                    try:
                        if pooled:
                            sink(i, mus[i], dst.numpy())
                    finally:
                        if pooled:
                            self._pool.put(dst)
            except BaseException as e:          # noqa: BLE001
                errors.append(e)

        def work(wid):
            stream = self._streams[wid]
            try:
                with torch.cuda.device(self.device), torch.cuda.stream(stream):
                    w0_dev = w0_pinned.to(self.device, non_blocking=True)
                    while not errors:
                        try:
                            idx = groups.get_nowait()
                        except queue.Empty:
                            break
                        solver = self._solver(wid, len(idx), [mus[i] for i in idx])
                        tw = time.perf_counter()
                        solver.set_initial(w0_dev)
                        solver.run()
                        with lock:
                            stats["launches"] += solver.launches
                            stats["redone"] += solver.steps_redone
                            stats["sweeps"] += time.perf_counter() - tw
                        stage = self._stage.setdefault(wid, [])
                        while len(stage) < len(idx):
                            ev = threading.Event()
                            ev.set()
                            stage.append((torch.empty((self.n2, self.num_steps + 1), dtype=torch.float64,
                                                      device=self.device), ev))
                        for j, i in enumerate(idx):
                            buf, free = stage[j]
                            tw = time.perf_counter()
                            free.wait()                 # its previous content has reached the host
                            free.clear()
                            with lock:
                                stats["stage_wait"] += time.perf_counter() - tw
                            solver.snapshots_device(j, layout=1, out=buf)
                            exported = torch.cuda.Event()
                            exported.record(stream)
                            pooled = out is None
                            if pooled:
                                tw = time.perf_counter()
                                dst = self._get_pool().get()
                                with lock:
                                    stats["pool_wait"] += time.perf_counter() - tw
                            else:
                                dst = out[i] if isinstance(out[i], torch.Tensor) else torch.from_numpy(out[i])
                            copy_q.put((exported, buf, dst, i, pooled, free))
            except BaseException as e:          # noqa: BLE001 - re-raised in the caller's thread
                errors.append(e)

        movers = [threading.Thread(target=copier, daemon=True), threading.Thread(target=deliver, daemon=True)]
        for t in movers:
            t.start()
        threads = [threading.Thread(target=work, args=(w,), daemon=True) for w in range(min(self.workers, groups.qsize()))]
        for t in threads:
            t.start()
        for t in threads:
            t.join()
        copy_q.put(None)
        for t in movers:
            t.join()
        if errors:
            raise errors[0]
        self.launches, self.steps_redone = stats["launches"], stats["redone"]
        if spans:
            # how busy the link was: time inside copies against first-copy-start .. last-copy-end, and the wall
            # time of the call before the first copy could start (the first group's sweeps)
            torch.cuda.synchronize(self.device)
            busy = sum(b.elapsed_time(d) for b, d, _ in spans) * 1e-3
            span = spans[0][0].elapsed_time(spans[-1][1]) * 1e-3
            self.copy_stats = {"copies": len(spans), "bytes": int(sum(n for _, _, n in spans)),
                               "seconds_in_copies": busy, "seconds_first_to_last": span,
                               "GBps_while_copying": sum(n for _, _, n in spans) / busy / 1e9 if busy > 0 else None,
                               "worker_seconds_in_sweeps": stats["sweeps"], "worker_seconds_waiting_for_staging": stats["stage_wait"],
                               "worker_seconds_waiting_for_pinned": stats["pool_wait"]}
        if out is None:
            return None
        return [o.numpy() if isinstance(o, torch.Tensor) else o for o in out]
