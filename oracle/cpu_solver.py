"""ctypes front end of oracle/mgcpu.c -- TEST INFRASTRUCTURE: the threaded C restatement of the reference's CPU algorithm
(HMG-global hierarchies without hanging nodes).  Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may use it.
The index maps are the product's (inputs shared by both sides); the arithmetic is mgcpu.c's own."""
import ctypes as C
import os
import subprocess

import numpy as np

from . import reference as ref

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


_NATIVE = None
_NATIVE_TAG = ""


def build_native():
    """build the library for THIS machine's CPU (-march=native) when a compiler is present -- the timed host baseline should
    use the vector width of the box it runs on; returns the description of the build that load() will use"""
    global _NATIVE
    if _LIB is not None:
        return _NATIVE or "-march=x86-64-v3 build"
    global _NATIVE_TAG
    _NATIVE_TAG = _cpu_tag()
    try:
        subprocess.check_call(["make", "-C", _HERE, "native", "NATIVE_TAG=" + _NATIVE_TAG], stdout=subprocess.DEVNULL,
                              stderr=subprocess.DEVNULL)
        _NATIVE = "-O3 -march=native build of this machine"
    except (OSError, subprocess.CalledProcessError):
        _NATIVE = None
    return _NATIVE or "-march=x86-64-v3 build"


def _cpu_tag():
    """names the -march=native library after this CPU's model and feature flags: the directory travels between machines"""
    import hashlib
    model = flags = ""
    try:
        with open("/proc/cpuinfo") as f:
            for line in f:
                if line.startswith("model name") and not model:
                    model = line
                elif line.startswith("flags") and not flags:
                    flags = line
                if model and flags:
                    break
    except OSError:
        pass
    return "_" + hashlib.sha1((model + flags).encode()).hexdigest()[:10]


def load():
    global _LIB
    if _LIB is not None:
        return _LIB
    path = os.path.join(_HERE, "_build", f"libmgcpu_native{_NATIVE_TAG}.so" if _NATIVE else "libmgcpu.so")
    if not os.path.exists(path):
        subprocess.check_call(["make", "-C", _HERE])
        path = os.path.join(_HERE, "_build", "libmgcpu.so")
    lib = C.CDLL(path)
    P = C.c_void_p
    lib.mgcpu_create.restype = P
    lib.mgcpu_create.argtypes = [C.c_int, C.c_int, C.c_double, C.c_int]
    lib.mgcpu_set_level.argtypes = [P, C.c_int, C.c_int, C.c_long, C.c_long, P, P, C.c_long, P, P, P, P, C.c_int, P, P]
    lib.mgcpu_set_transfer.argtypes = [P, C.c_int, P, P, P, P, P]
    lib.mgcpu_setup.argtypes = [P]
    lib.mgcpu_vmult.argtypes = [P, C.c_int, P, P]
    lib.mgcpu_lambda_max.restype = C.c_double
    lib.mgcpu_lambda_max.argtypes = [P, C.c_int]
    lib.mgcpu_transfer.argtypes = [P, C.c_int, P, P, C.c_int]
    lib.mgcpu_vcycle.argtypes = [P, P, P]
    lib.mgcpu_solve.restype = C.c_int
    lib.mgcpu_solve.argtypes = [P, P, P, C.c_double, C.POINTER(C.c_double), C.POINTER(C.c_long)]
    lib.mgcpu_set_threads.argtypes = [C.c_int]
    lib.mgcpu_num_threads.restype = C.c_int
    lib.mgcpu_set_batched.argtypes = [C.c_int]
    lib.mgcpu_vector_lanes.restype = C.c_int
    lib.mgcpu_set_even_odd.argtypes = [C.c_int]
    lib.mgcpu_set_blocks.argtypes = [C.c_int]
    lib.mgcpu_set_vmult_blocks.restype = C.c_int
    lib.mgcpu_set_vmult_blocks.argtypes = [P, C.c_int, C.c_long, C.c_int, P, P]
    lib.mgcpu_even_odd_available.restype = C.c_int
    lib.mgcpu_even_odd_available.argtypes = [P, C.c_int]
    lib.mgcpu_destroy.argtypes = [P]
    _LIB = lib
    return lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


class CpuMultigrid:
    """HMG-global V-cycle-preconditioned CG on the host cores.  levels: list of (mesh, dofmap) coarse -> fine from the
    product's generator, tmaps: product TransferMaps (index l connects l-1 and l)."""

    def __init__(self, dofmaps, tmaps, smoother_degree=3, smoothing_range=20.0, eig_its=20):
        lib = load()
        self.lib = lib
        self.maps = [ref.LevelMaps.from_dofmap(d) for d in dofmaps]
        self.h = lib.mgcpu_create(len(dofmaps), smoother_degree, smoothing_range, eig_its)
        self._keep = []
        for l, (d, m) in enumerate(zip(dofmaps, self.maps)):
            if np.any(m.mask >> 3):
                raise ValueError("mgcpu.c covers meshes without hanging nodes only")
            idx = (m.eval_idx.astype(np.uint32) | (m.eval_flag.astype(np.uint32) << 31)).astype(np.uint32)
            idx = np.ascontiguousarray(idx)
            hcell = np.ascontiguousarray(m.h, dtype=np.float64)
            con = np.ascontiguousarray(m.constrained, dtype=np.uint32)
            xyz = m.cell_xyz
            uniform = len(np.unique(m.cell_level)) == 1
            if uniform:
                colour = (xyz[:, 0] & 1) | ((xyz[:, 1] & 1) << 1) | ((xyz[:, 2] & 1) << 2)
                order = np.argsort(colour, kind="stable").astype(np.uint32)
                begin = np.searchsorted(colour[order], np.arange(9)).astype(np.int64)
                ncol = 8
            else:
                order, begin, ncol = np.zeros(1, np.uint32), np.zeros(2, np.int64), 0
            fe = m.fe
            B, Dc, w = (np.ascontiguousarray(a, dtype=np.float64) for a in (fe.B, fe.Dc, fe.wq))
            self._keep += [idx, hcell, con, order, begin, B, Dc, w]
            lib.mgcpu_set_level(self.h, l, d.degree, m.n_cells, m.n_dofs, _p(idx), _p(hcell), len(con), _p(con), _p(B), _p(Dc), _p(w),
                                ncol, _p(begin), _p(order))
            if uniform:
                # operator loop: colour Morton-aligned cubes of 8^k cells by the parity of their block coordinates; the largest
                # k <= 3 that leaves 64 blocks per colour (the cells are in Morton order, so a stable sort keeps blocks whole)
                k = 3
                while k > 0 and m.n_cells // 8 ** k < 8 * 64:
                    k -= 1
                bcol = ((xyz[:, 0] >> k) & 1) | (((xyz[:, 1] >> k) & 1) << 1) | (((xyz[:, 2] >> k) & 1) << 2)
                border = np.ascontiguousarray(np.argsort(bcol, kind="stable").astype(np.uint32))
                bbegin = np.ascontiguousarray(np.searchsorted(bcol[border], np.arange(9)).astype(np.int64))
                if k >= 1 and m.n_cells >= 8 ** (k + 1):   # k = 0 is the cell colouring above
                    self._keep += [border, bbegin]
                    if lib.mgcpu_set_vmult_blocks(self.h, l, 8 ** k, 8, _p(bbegin), _p(border)) != 0:
                        raise RuntimeError("mgcpu_set_vmult_blocks: colour lists are not whole blocks")
        for l in range(1, len(dofmaps)):
            fine, coarse, tm = self.maps[l], self.maps[l - 1], tmaps[l]
            first = np.ascontiguousarray(tm.fine_first(), dtype=np.uint32)
            nch = np.ascontiguousarray(tm.n_children(), dtype=np.uint8)
            w = np.ascontiguousarray(ref.transfer_weights_uniform(fine, coarse, first, nch), dtype=np.float64)
            own = np.ascontiguousarray(fine.own_idx, dtype=np.uint32)
            P1 = np.ascontiguousarray(coarse.fe.Ph, dtype=np.float64)
            self._keep += [first, nch, w, own, P1]
            lib.mgcpu_set_transfer(self.h, l, _p(first), _p(nch), _p(w), _p(own), _p(P1))
        lib.mgcpu_setup(self.h)

    def autotune_threads(self, candidates=None):
        """pick the OpenMP thread count that applies the finest-level operator fastest (containers often expose more
        hardware threads than the cgroup lets the process use; oversubscribed OpenMP teams are far slower than one thread)"""
        import time
        ncpu = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
        if candidates is None:
            candidates = sorted({1, 2, 4, 8, 16, 32, 64, 128, ncpu, max(1, ncpu // 2)})
        candidates = [c for c in candidates if c <= ncpu]
        l = len(self.maps) - 1
        x = np.ones(self.maps[l].n_dofs)
        best, best_t = 1, 1e300
        for c in candidates:
            self.lib.mgcpu_set_threads(c)
            self.vmult(l, x)
            t0 = time.perf_counter()
            for _ in range(3):
                self.vmult(l, x)
            dt = time.perf_counter() - t0
            if dt < 0.97 * best_t:   # more threads only if they measurably help (hyper-threads often do not)
                best, best_t = c, dt
        self.lib.mgcpu_set_threads(best)
        return best

    def set_batched(self, on):
        """cell kernel on mgcpu_vector_lanes() cells at once (default, what deal.II's VectorizedArray does) or cell by cell"""
        self.lib.mgcpu_set_batched(1 if on else 0)

    def set_blocks(self, on):
        """operator loop over coloured Morton-aligned blocks of cells (default; one pass over the vectors) or coloured cells"""
        self.lib.mgcpu_set_blocks(1 if on else 0)

    def set_even_odd(self, on):
        """1-D sweeps in even-odd form (default; deal.II's evaluate_evenodd) or as full (p+1) x (p+1) products"""
        self.lib.mgcpu_set_even_odd(1 if on else 0)

    def even_odd_available(self, level):
        return bool(self.lib.mgcpu_even_odd_available(self.h, level))

    @property
    def vector_lanes(self):
        return int(self.lib.mgcpu_vector_lanes())

    @property
    def n_threads(self):
        return int(self.lib.mgcpu_num_threads())

    def vmult(self, l, x):
        x = np.ascontiguousarray(x, dtype=np.float64)
        y = np.zeros_like(x)
        self.lib.mgcpu_vmult(self.h, l, _p(y), _p(x))
        return y

    def lambda_max(self, l):
        return float(self.lib.mgcpu_lambda_max(self.h, l))

    def prolongate_and_add(self, l, dst_fine, src_coarse):
        src = np.ascontiguousarray(src_coarse, dtype=np.float64)
        self.lib.mgcpu_transfer(self.h, l, _p(dst_fine), _p(src), 1)

    def restrict_and_add(self, l, dst_coarse, src_fine):
        src = np.ascontiguousarray(src_fine, dtype=np.float64)
        self.lib.mgcpu_transfer(self.h, l, _p(dst_coarse), _p(src), 0)

    def vcycle(self, r):
        r = np.ascontiguousarray(r, dtype=np.float64)
        y = np.zeros_like(r)
        self.lib.mgcpu_vcycle(self.h, _p(y), _p(r))
        return y

    def solve(self, b, reltol=1e-4):
        b = np.ascontiguousarray(b, dtype=np.float64)
        x = np.zeros_like(b)
        res, cits = C.c_double(0), C.c_long(0)
        its = self.lib.mgcpu_solve(self.h, _p(x), _p(b), reltol, C.byref(res), C.byref(cits))
        return x, int(its), float(res.value), int(cits.value)

    def __del__(self):
        if getattr(self, "h", None):
            self.lib.mgcpu_destroy(self.h)
            self.h = None
