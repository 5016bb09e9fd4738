"""Device-resident particle state + the whole time step.

The reference keeps its state as module-level numpy arrays in the driver
(/root/reference/sph/code_running.py:121-167) and advances it inline (:217-540); there is no function
for it.  `Session` holds the same struct-of-arrays in HBM and advances it with
`sphb200_session_step` so nothing crosses PCIe between steps.  Field names are the driver's
variable names.
"""
import ctypes
import os

import numpy as np
import torch

from . import _lib
from . import constants as C

STATE_FIELDS = {"points": 3, "velocities": 3, "total_accel": 3, "mass": 1, "sizes": 1, "T": 1, "E_internal": 1,
                "mu_array": 1, "gamma_array": 1, "particle_type": 1, "densities_0": 1}
RESULT_FIELDS = {"densities": 1, "num_densities": 1, "dust_densities": 1, "visc_heat": 1, "hydro_accel": 3,
                 "visc_accel": 3, "pressure_accel": 3, "grav_accel": 3}
# owner-computed per-particle arrays the slab-decomposed driver reads / overwrites for ghosts (domain.py)
AUX_FIELDS = {"densities": 1, "visc_pi": 1, "knn_h": 1, "dist_sq": 1, "gid": 1}
# host-resident stepping (step_host): what crossing time / clamp / cell build / k-NN read, and what must
# be re-uploaded after the early sort because "sort_keys" does not carry it into the new order
HOST_EARLY_FIELDS = ("points", "velocities", "sizes", "particle_type")
HOST_LATE_FIELDS = ("total_accel", "mass", "T", "E_internal", "mu_array", "gamma_array", "densities_0")
PHASES = ("crossing", "dt_clamp", "sort", "knn", "pack", "density", "pi", "force", "integrate")


class Session(object):
    def __init__(self, n, k=C.N_NEIGH, dist=np.inf, device=None, dt_0=C.dt_0):
        # a context of its own: the session's scratch buffers (neighbour list, records, cell list) are never
        # shared with the function-level nsc.* calls (they use the process-wide _lib.context()) nor with
        # another Session on the same GPU
        if not torch.cuda.is_available():
            raise _lib.Sphb200Error("no CUDA device visible: sphb200 runs on a B200 only (no CPU fallback)")
        self.ctx = _lib.Context(torch.cuda.current_device() if device is None else int(device))
        self.lib = _lib.load()
        # tunables from the environment, "name=value,name=value" (sphb200_set_option), e.g. SPHB200_OPTIONS=knn_margin=0.1
        for item in filter(None, os.environ.get("SPHB200_OPTIONS", "").split(",")):
            name, _, val = item.partition("=")
            self.ctx.set_option(name.strip(), float(val))
        self.n = int(n)
        self.k = int(k)
        self.device = torch.device("cuda", self.ctx.device)
        self.dist, self.dt_0 = float(dist), float(dt_0)
        _lib.check(self.lib.sphb200_session_create(self.ctx.handle, self.n, self.k, float(dist)))
        for name, val in (("dt_0", dt_0), ("MAX_AGE", C.MAX_AGE), ("N_INT_PER_PARTICLE", C.N_INT_PER_PARTICLE),
                          ("raise_factor", C.raise_factor), ("AU", C.AU)):
            self.set_scalar(name, val)

    def close(self):
        """Free the session and its context (device buffers)."""
        if getattr(self, "ctx", None) is not None:
            self.ctx.close()
            self.ctx = None

    def resize(self, n, n_owned=-1):
        """Same run, new particle set (slab decomposition): keeps scalars and run parameters."""
        _lib.check(self.lib.sphb200_session_resize(self.ctx.handle, int(n), int(n_owned), self._stream()))
        self.n = int(n)

    # -- plumbing ---------------------------------------------------------------------------
    def _stream(self):
        return ctypes.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def set_scalar(self, name, value):
        _lib.check(self.lib.sphb200_session_set_scalar(self.ctx.handle, name.encode(), float(value)))
        if name in ("dist", "dt_0"):
            setattr(self, name, float(value))

    def scalar(self, name):
        out = ctypes.c_double(0.0)
        _lib.check(self.lib.sphb200_session_scalar(self.ctx.handle, name.encode(), ctypes.byref(out), self._stream()))
        return float(out.value)

    def upload_tensor(self, name, t):
        """t: contiguous float64 CUDA tensor in the caller's particle order."""
        w = STATE_FIELDS.get(name) or AUX_FIELDS[name]
        if t.numel() != self.n * w or t.dtype != torch.float64 or not t.is_cuda:
            raise ValueError("field %s needs a float64 CUDA tensor of %d values" % (name, self.n * w))
        _lib.check(self.lib.sphb200_session_upload(self.ctx.handle, name.encode(), ctypes.c_void_p(t.data_ptr()),
                                                   t.numel(), self._stream()))

    def upload(self, state):
        """state: dict of host arrays keyed by the driver's variable names (missing keys keep zeros)."""
        for name in STATE_FIELDS:
            if name in state and state[name] is not None:
                a = np.ascontiguousarray(np.asarray(state[name]), dtype=np.float64)
                self.upload_tensor(name, torch.from_numpy(a).to(self.device))
        if "d" in state:
            self.set_scalar("d", state["d"])

    def download_tensor(self, name):
        if name == "neighbor":
            out = torch.empty((self.n, self.k), dtype=torch.int64, device=self.device)
        elif name == "num_nontrivial":
            out = torch.empty((self.n,), dtype=torch.int64, device=self.device)
        else:
            w = STATE_FIELDS.get(name) or RESULT_FIELDS.get(name) or AUX_FIELDS[name]
            out = torch.empty((self.n, w) if w > 1 else (self.n,), dtype=torch.float64, device=self.device)
        _lib.check(self.lib.sphb200_session_download(self.ctx.handle, name.encode(), ctypes.c_void_p(out.data_ptr()),
                                                     out.numel(), self._stream()))
        return out

    def download_raw(self, name, count):
        """First `count` doubles of an internal array that has no particle order ("dist_sq_sorted")."""
        out = torch.empty((count,), dtype=torch.float64, device=self.device)
        _lib.check(self.lib.sphb200_session_download(self.ctx.handle, name.encode(), ctypes.c_void_p(out.data_ptr()),
                                                     int(count), self._stream()))
        return out

    def upload_raw(self, name, t):
        """Device-to-device write of "scalar:<name>" from a 1-element float64 CUDA tensor (no host sync)."""
        _lib.check(self.lib.sphb200_session_upload(self.ctx.handle, name.encode(), ctypes.c_void_p(t.data_ptr()),
                                                   int(t.numel()), self._stream()))

    def buffer_tensor(self, name):
        """Zero-copy tensor over an internal device array: "length_hist" (int64), "cover" (float64)."""
        ptr = ctypes.c_void_p(0); cnt = ctypes.c_int64(0)
        _lib.check(self.lib.sphb200_session_buffer(self.ctx.handle, name.encode(), ctypes.byref(ptr), ctypes.byref(cnt)))

        class _Raw(object):
            __cuda_array_interface__ = {"shape": (int(cnt.value),), "typestr": "<f8" if name == "cover" else "<i8",
                                        "data": (int(ptr.value), False), "version": 2, "strides": None}
        return torch.as_tensor(_Raw(), device=self.device)

    def cover_check(self, lo, hi):
        _lib.check(self.lib.sphb200_session_cover_check(self.ctx.handle, float(lo), float(hi), self._stream()))

    def download_rows_tagged(self, n_rows, ncol, tag):
        """download_rows with column 18 taken from `tag` (float64 CUDA tensor, caller order)."""
        rows = torch.empty((int(n_rows), int(ncol)), dtype=torch.float64, device=self.device)
        _lib.check(self.lib.sphb200_session_download_rows_tagged(self.ctx.handle, ctypes.c_void_p(rows.data_ptr()), int(n_rows),
                                                                 int(ncol), ctypes.c_void_p(tag.data_ptr()), self._stream()))
        return rows

    def upload_rows_at(self, rows, first):
        if rows.dtype != torch.float64 or not rows.is_cuda or not rows.is_contiguous():
            raise ValueError("upload_rows_at needs a contiguous float64 CUDA tensor")
        _lib.check(self.lib.sphb200_session_upload_rows_at(self.ctx.handle, ctypes.c_void_p(rows.data_ptr()), int(first),
                                                           int(rows.shape[0]), int(rows.shape[1]), self._stream()))

    def gather(self, name, idx):
        """Field `name` at caller indices idx (int64 CUDA tensor) -> (len(idx), width) float64 tensor."""
        w = STATE_FIELDS.get(name) or RESULT_FIELDS.get(name) or AUX_FIELDS[name]
        out = torch.empty((idx.numel(), w), dtype=torch.float64, device=self.device)
        _lib.check(self.lib.sphb200_session_gather(self.ctx.handle, name.encode(), ctypes.c_void_p(idx.data_ptr()),
                                                   int(idx.numel()), ctypes.c_void_p(out.data_ptr()), self._stream()))
        return out

    def scatter(self, name, idx, values):
        _lib.check(self.lib.sphb200_session_scatter(self.ctx.handle, name.encode(), ctypes.c_void_p(idx.data_ptr()),
                                                    int(idx.numel()), ctypes.c_void_p(values.data_ptr()), self._stream()))

    def download_rows(self, n_rows, ncol):
        """(n_rows, ncol) float64 CUDA tensor: state fields + trial radius of caller rows [0, n_rows)."""
        out = torch.zeros((n_rows, ncol), dtype=torch.float64, device=self.device)
        _lib.check(self.lib.sphb200_session_download_rows(self.ctx.handle, ctypes.c_void_p(out.data_ptr()), int(n_rows),
                                                          int(ncol), self._stream()))
        return out

    def upload_rows(self, rows):
        if rows.shape[0] != self.n or rows.dtype != torch.float64 or not rows.is_cuda or not rows.is_contiguous():
            raise ValueError("upload_rows needs a contiguous float64 CUDA tensor of %d rows" % self.n)
        _lib.check(self.lib.sphb200_session_upload_rows(self.ctx.handle, ctypes.c_void_p(rows.data_ptr()), self.n,
                                                        int(rows.shape[1]), self._stream()))

    def append_rows(self, rows):
        """SURVEY row F4 (code_running.py:361-381): append particles (rows in the layout of upload_rows) as
        caller indices N .. N+n_new-1.  Call search() before the next step, as the driver calls
        nsc.neighbors at :386."""
        if (not torch.is_tensor(rows) or rows.dim() != 2 or rows.dtype != torch.float64 or not rows.is_cuda
                or not rows.is_contiguous() or rows.shape[1] < 18):
            raise ValueError("append_rows needs a contiguous float64 CUDA tensor (n_new, >=18)")
        _lib.check(self.lib.sphb200_session_append_rows(self.ctx.handle, ctypes.c_void_p(rows.data_ptr()),
                                                        int(rows.shape[0]), int(rows.shape[1]), self._stream()))
        torch.cuda.current_stream(self.device).synchronize()      # `rows` may be freed by the caller
        self.n += int(rows.shape[0])

    def append(self, state):
        """append() for a dict of host arrays keyed like upload() (missing fields are zero)."""
        m = len(state["points"])
        rows = np.zeros((m, 18))
        col = 0
        for name, w in STATE_FIELDS.items():
            if state.get(name) is not None:
                rows[:, col:col + w] = np.asarray(state[name], dtype=np.float64).reshape(m, w)
            col += w
        self.append_rows(torch.from_numpy(rows).to(self.device))

    def search(self):
        """Neighbour query on the current positions (nsc.neighbors, code_running.py:386): fills
        neighbor / num_nontrivial and overwrites sizes with the k-th neighbour distance."""
        self.phase("sort")
        self.phase("knn")

    def download(self, name):
        return self.download_tensor(name).cpu().numpy()

    # -- the path ---------------------------------------------------------------------------
    def upload_neighbor(self, neighbor):
        """Restore a saved (N,k) neighbour array (caller's indices; N = missing neighbour)."""
        nb = torch.as_tensor(np.ascontiguousarray(np.asarray(neighbor), dtype=np.int64)).to(self.device)
        if tuple(nb.shape) != (self.n, self.k):
            raise ValueError("neighbor must be (%d, %d)" % (self.n, self.k))
        _lib.check(self.lib.sphb200_session_upload(self.ctx.handle, b"neighbor", ctypes.c_void_p(nb.data_ptr()),
                                                   self.n * self.k, self._stream()))
        torch.cuda.current_stream(self.device).synchronize()
        self.check()

    def set_drag(self, f_un):
        """Dust-gas drag of code_running.py:455,462-463,473 in the force phase (`net_impulse`, nsc:719-742):
        f_un is the (N, n_species) composition in the caller's order; None switches the drag off."""
        if f_un is None:
            _lib.check(self.lib.sphb200_session_set_drag(self.ctx.handle, None, 0, None, None, self._stream()))
            return
        from . import navier_stokes_cleaned as nsc
        f = torch.as_tensor(np.ascontiguousarray(np.asarray(f_un, dtype=np.float64)) if not torch.is_tensor(f_un) else f_un)
        f = f.to(device=self.device, dtype=torch.float64).contiguous()
        if f.dim() != 2 or f.shape[0] != self.n:
            raise ValueError("f_un must be (N, n_species)")
        ns = int(f.shape[1])
        meff = np.ascontiguousarray(np.broadcast_to(nsc.grain_mass(C.mineral_densities, C.mrn_constants), (ns,)), dtype=np.float64)
        seff = np.ascontiguousarray(np.broadcast_to(nsc.sigma_effective(C.mineral_densities, C.mrn_constants, C.mu_specie), (ns,)),
                                    dtype=np.float64)
        _lib.check(self.lib.sphb200_session_set_drag(self.ctx.handle, f.data_ptr(), ns, meff.ctypes.data_as(ctypes.c_void_p),
                                                     seff.ctypes.data_as(ctypes.c_void_p), self._stream()))
        torch.cuda.current_stream(self.device).synchronize()       # f lives until the means are computed

    def set_gravity(self, on=True, softening=None):
        """self-gravity in the step (code_running.py:448,477).  softening=None follows the reference
        (nsc:359): the median of `sizes` right after the neighbour search, recomputed every step (one host
        synchronisation per step); a number fixes the Plummer length.  Direct O(N^2) sum: for N up to ~10^5."""
        self._gravity = bool(on)
        self._grav_soft = softening
        if not on:
            _lib.check(self.lib.sphb200_session_set_gravity(self.ctx.handle, C.G, 0.0))
        elif softening is not None:
            _lib.check(self.lib.sphb200_session_set_gravity(self.ctx.handle, C.G, float(softening)))

    def startup(self):
        """code_running.py:171-179: first neighbour query + hydro_update; densities_0 <- densities."""
        _lib.check(self.lib.sphb200_session_startup(self.ctx.handle, self._stream()))

    def step(self):
        """One pass of the while-body (code_running.py:217-540), asynchronous on the current stream."""
        if getattr(self, "_gravity", False) and self._grav_soft is None:
            # the reference's softening is the median of the sizes the search just returned (nsc:359 after :437)
            for ph in ("crossing", "dt_clamp", "sort", "knn"):
                self.phase(ph)
            hs = torch.sort(self.download_tensor("sizes")).values         # np.median: mean of the two middle values
            eps = float((0.5 * (hs[(self.n - 1) // 2] + hs[self.n // 2])).item())
            _lib.check(self.lib.sphb200_session_set_gravity(self.ctx.handle, C.G, eps))
            for ph in ("pack", "density", "pi", "force", "gravity", "integrate"):
                self.phase(ph)
            return
        _lib.check(self.lib.sphb200_session_step(self.ctx.handle, self._stream()))

    def capture_step(self, steps=2):
        """`steps` time steps captured into ONE CUDA graph (torch.cuda.CUDAGraph); `replay()` runs them with a single
        launch.  The step makes no host synchronisation and chooses its grid, its time step and its work lists
        on the device, so the capture holds every kernel of the step; buffers must have reached their size first
        (run at least one ordinary step before capturing: a capture cannot call cudaMalloc).  `steps` must be even:
        the sort writes the state into the other half of a pair of buffers and the library flips which half is
        current on the host, so the captured pointers are right again after two steps.  Not available with
        self-gravity, whose softening is read back by the host."""
        if getattr(self, "_gravity", False):
            raise _lib.Sphb200Error("capture_step: the gravity step reads its softening back on the host")
        if int(steps) < 2 or int(steps) % 2:
            raise ValueError("capture_step: an even number of steps (the state buffers alternate every step)")
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            for _ in range(int(steps)):
                _lib.check(self.lib.sphb200_session_step(self.ctx.handle, self._stream()))
        return graph

    # -- slab decomposition (domain.py drives the collectives) ------------------------------------
    def enable_gid(self):
        _lib.check(self.lib.sphb200_session_enable_gid(self.ctx.handle))

    def migrate_pack(self, phase, bounds, world, rank, rows_out, counts_out):
        _lib.check(self.lib.sphb200_session_migrate_pack(
            self.ctx.handle, int(phase), ctypes.c_void_p(bounds.data_ptr()), int(world), int(rank),
            ctypes.c_void_p(rows_out.data_ptr()) if rows_out is not None and rows_out.numel() else None,
            int(rows_out.shape[0]) if rows_out is not None else 0, 19,
            ctypes.c_void_p(counts_out.data_ptr()) if counts_out is not None else None, self._stream()))

    def set_count(self, n, n_owned):
        _lib.check(self.lib.sphb200_session_set_count(self.ctx.handle, int(n), int(n_owned)))
        self.n = int(n)

    def receive_rows(self, rows, as_ghosts):
        _lib.check(self.lib.sphb200_session_receive_rows(
            self.ctx.handle, ctypes.c_void_p(rows.data_ptr()) if rows.numel() else None, int(rows.shape[0]), int(rows.shape[1]),
            1 if as_ghosts else 0, self._stream()))
        self.n += int(rows.shape[0])

    def halo_pack(self, phase, bounds, world, rank, w, rows_out, idx_out, counts_out):
        _lib.check(self.lib.sphb200_session_halo_pack(
            self.ctx.handle, int(phase), ctypes.c_void_p(bounds.data_ptr()), int(world), int(rank), float(w),
            ctypes.c_void_p(rows_out.data_ptr()) if rows_out is not None and rows_out.numel() else None,
            ctypes.c_void_p(idx_out.data_ptr()) if idx_out is not None and idx_out.numel() else None,
            int(rows_out.shape[0]) if rows_out is not None else 0, 19,
            ctypes.c_void_p(counts_out.data_ptr()) if counts_out is not None else None, self._stream()))

    def perturb(self, amplitude, seed):
        """points += amplitude * u(id, seed) (synthetic motion between steps; ics.perturbation is the host mirror)."""
        _lib.check(self.lib.sphb200_session_perturb(self.ctx.handle, float(amplitude), int(seed), self._stream()))

    def phase(self, name):
        _lib.check(self.lib.sphb200_session_phase(self.ctx.handle, name.encode(), self._stream()))

    def check(self):
        """Raise if the last k-NN searches did not converge."""
        over = self.ctx.counter("knn_overflow", self._stream())
        if over:
            raise _lib.Sphb200Error("k-NN search did not converge for %d particles" % over)

    def launches(self):
        return self.ctx.counter("launches")

    def counters(self, reset=True):
        """Statistics of the neighbour searches since the last read: extra scans (stale trial radius), targets
        handed to the general kernel (k-th neighbour outside the cell block), length of the last work list."""
        st = self._stream()
        return {"knn_rescan": self.ctx.counter("knn_rescan", st), "knn_widen": self.ctx.counter("knn_widen", st),
                "knn_worklist_last": self.ctx.counter("knn_worklist", st)}

    def median_h(self):
        """Median smoothing length of the current state (device-side; one host sync)."""
        return float(self.download_tensor("sizes").median().item())

    def fp64_peak(self):
        """Measured fp64 rate of this GPU (DFMA thread-instructions per second of a register-resident chain):
        the denominator of the fp64-pipe fractions bench.py reports."""
        out = ctypes.c_double(0.0)
        _lib.check(self.lib.sphb200_fp64_peak(self.ctx.handle, ctypes.byref(out), self._stream()))
        return {"dfma_per_s": float(out.value), "tflops": 2.0 * float(out.value) / 1e12,
                "how": "sphb200_fp64_peak: 8 independent DFMA chains per thread, 592 CTAs x 512 threads, CUDA events"}

    def phases(self):
        return PHASES

    def n_owned(self):
        return self.n

    # -- host-resident state (the call a user of the reference makes: numpy arrays in, arrays out) --
    def make_host_state(self, upload_fields, download_fields):
        """Pinned host buffers holding the current state (inputs) and room for the results, plus the
        device staging buffers and the copy stream step_host() uses."""
        hs = {"in": {}, "out": {}, "stage": {}, "upload": tuple(upload_fields), "download": tuple(download_fields)}
        for f in upload_fields:
            hs["in"][f] = self.download_tensor(f).cpu().pin_memory()
            hs["stage"][f] = torch.empty_like(hs["in"][f], device=self.device)
        for f in download_fields:
            w = STATE_FIELDS.get(f) or RESULT_FIELDS[f]
            hs["out"][f] = torch.empty((self.n, w) if w > 1 else (self.n,), dtype=torch.float64).pin_memory()
        hs["h2d_bytes"] = sum(t.numel() * 8 for t in hs["in"].values())
        hs["d2h_bytes"] = sum(t.numel() * 8 for t in hs["out"].values())
        hs["copy_stream"] = torch.cuda.Stream(device=self.device)
        # fields the step reads before / during the neighbour search; the rest may arrive while it runs
        hs["early"] = tuple(f for f in upload_fields if f in HOST_EARLY_FIELDS)
        hs["late"] = tuple(f for f in upload_fields if f not in HOST_EARLY_FIELDS)
        hs["overlap"] = set(HOST_EARLY_FIELDS) <= set(upload_fields) and set(HOST_LATE_FIELDS) <= set(upload_fields)
        return hs

    def step_host(self, hs, perturb=None):
        """One time step of a host-resident state: every input field is copied from pinned memory, the step
        runs, the advanced fields are copied back to pinned memory and become the next step's inputs.
        The copy of the fields that are first read after the neighbour search (72 of the 136 B/particle)
        runs on a second stream underneath crossing time, cell build and k-NN: the early phase "sort_keys"
        carries only points / velocities / particle_type into the new order, and the late fields are then
        uploaded straight into it."""
        cur = torch.cuda.current_stream(self.device)
        if not hs.get("overlap"):
            for f in hs["upload"]:
                self.upload_tensor(f, hs["in"][f].to(self.device, non_blocking=True))
            if perturb:
                self.perturb(*perturb)
            self.step()
        else:
            cs = hs["copy_stream"]
            cs.wait_stream(cur)                                   # the staging buffers are free again
            with torch.cuda.stream(cs):
                for f in hs["early"]:
                    hs["stage"][f].copy_(hs["in"][f], non_blocking=True)
                ev_early = cs.record_event()
                for f in hs["late"]:
                    hs["stage"][f].copy_(hs["in"][f], non_blocking=True)
                ev_late = cs.record_event()
            cur.wait_event(ev_early)
            for f in hs["early"]:
                self.upload_tensor(f, hs["stage"][f])
            if perturb:
                self.perturb(*perturb)                            # synthetic motion (bench / tests), as before step()
            for ph in ("crossing", "dt_clamp", "sort_keys", "knn"):
                self.phase(ph)
            cur.wait_event(ev_late)
            for f in hs["late"]:
                self.upload_tensor(f, hs["stage"][f])             # gathers through the new id: lands in the new order
            for ph in ("pack", "density", "pi", "force", "integrate"):
                self.phase(ph)
        for f in hs["download"]:
            hs["out"][f].copy_(self.download_tensor(f), non_blocking=True)
        cur.synchronize()                                         # the caller reads the result
        for f in hs["download"]:
            if f in hs["in"]:
                hs["in"][f], hs["out"][f] = hs["out"][f], hs["in"][f]
