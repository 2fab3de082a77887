"""ctypes view of libufe_host.so: synthetic type_mesh inputs + ownership ranges (CPU only).

Arrays are zero-copy numpy views in the Fortran layout the C ABI expects
(column-major, 1-based indices; see host/ufe_mesh.hpp).
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


def lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(_HERE, "host", "libufe_host.so")
        if not os.path.exists(path):
            raise RuntimeError(f"{path} missing: run `python -c 'import __graft_entry__ as g; g.build()'`")
        L = C.CDLL(path)
        L.ufe_host_mesh_synth.restype = C.c_void_p
        L.ufe_host_mesh_synth.argtypes = [C.c_double] * 6 + [C.c_ulonglong] + [C.c_double] * 3 + [C.c_int] * 3 + [C.c_double]
        L.ufe_host_mesh_dummy5.restype = C.c_void_p
        L.ufe_host_mesh_dummy5.argtypes = [C.c_double] * 4
        L.ufe_host_mesh_from_primary.restype = C.c_void_p
        L.ufe_host_mesh_from_primary.argtypes = [C.c_int] * 5 + [C.c_double] * 5 + [C.c_void_p] * 9
        L.ufe_host_mesh_free.argtypes = [C.c_void_p]
        L.ufe_host_mesh_error.restype = C.c_char_p
        L.ufe_host_mesh_error.argtypes = [C.c_void_p]
        L.ufe_host_mesh_dims.argtypes = [C.c_void_p, C.POINTER(C.c_int)]
        L.ufe_host_mesh_domain.argtypes = [C.c_void_p, C.POINTER(C.c_double)]
        L.ufe_host_mesh_array.restype = C.c_void_p
        L.ufe_host_mesh_array.argtypes = [C.c_void_p, C.c_char_p, C.POINTER(C.c_int), C.POINTER(C.c_longlong)]
        L.ufe_host_partition_list.argtypes = [C.c_int] * 3 + [C.POINTER(C.c_int)] * 2
        L.ufe_host_determine_halos.argtypes = [C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_int)]
        _LIB = L
    return _LIB


_SHAPES = {  # name -> number of columns (Fortran second dimension); rows are implied by the count
    "V": 2, "Tricc": 2, "TriGC": 2, "E": 2, "Tri": 3, "TriC": 3, "TriE": 3, "EV": 4, "ETri": 2,
    "C": "nC_mem", "iTri": "nC_mem", "VE": "nC_mem", "Cw": "nC_mem", "D_x": "nC_mem", "D_y": "nC_mem", "D": "nC_mem",
}
_NAMES = ["V", "nC", "C", "niTri", "iTri", "VBI", "Tri", "Tricc", "TriC", "A", "Cw", "D_x", "D_y", "D", "TriBI",
          "TriGC", "TriA", "E", "VE", "EV", "ETri", "TriE", "EBI", "zeta"]


class HostMesh:
    """Owns one host mesh; attributes are Fortran-ordered numpy views, e.g. m.C has shape (nV, nC_mem)."""

    def __init__(self, handle):
        L = lib()
        self._h = handle
        err = L.ufe_host_mesh_error(handle)
        if err:
            msg = err.decode()
            L.ufe_host_mesh_free(handle)
            self._h = None
            raise RuntimeError("mesh generation failed: " + msg)
        d = (C.c_int * 5)()
        L.ufe_host_mesh_dims(handle, d)
        self.nV, self.nTri, self.nE, self.nC_mem, self.nz = [int(x) for x in d]
        dom = (C.c_double * 4)()
        L.ufe_host_mesh_domain(handle, dom)
        self.xmin, self.xmax, self.ymin, self.ymax = [float(x) for x in dom]
        for name in _NAMES:
            dt, cnt = C.c_int(), C.c_longlong()
            p = L.ufe_host_mesh_array(handle, name.encode(), C.byref(dt), C.byref(cnt))
            n = cnt.value
            if n == 0:
                arr = np.zeros(0, dtype=np.int32 if dt.value == 0 else np.float64)
            else:
                ct = C.c_int32 if dt.value == 0 else C.c_double
                arr = np.ctypeslib.as_array(C.cast(p, C.POINTER(ct)), shape=(n,))
            cols = _SHAPES.get(name)
            if cols is not None:
                cols = self.nC_mem if cols == "nC_mem" else cols
                arr = arr.reshape((n // cols, cols), order="F")
            setattr(self, name, arr)

    @classmethod
    def synthetic(cls, xmin, xmax, ymin, ymax, h, jitter=0.2, seed=20261104, grade_x0=0.0, grade_width=0.0,
                  grade_depth=0.0, nC_mem=32, nz=12, zeta_kind=0, zeta_R=10.0):
        return cls(lib().ufe_host_mesh_synth(xmin, xmax, ymin, ymax, h, jitter, seed, grade_x0, grade_width,
                                             grade_depth, nC_mem, nz, zeta_kind, zeta_R))

    @classmethod
    def dummy5(cls, xmin=-1.0, xmax=1.0, ymin=-1.0, ymax=1.0):
        return cls(lib().ufe_host_mesh_dummy5(xmin, xmax, ymin, ymax))

    @classmethod
    def from_primary(cls, V, nC, C_, niTri, iTri, VBI, Tri, Tricc, TriC, xmin, xmax, ymin, ymax, nz=12, zeta_kind=0, zeta_R=10.0):
        """A mesh from primary data the caller holds (Fortran shapes: V (nV,2), C (nV,nC_mem), Tri (nTri,3), ...; 1-based indices)."""
        f64 = lambda a: np.asfortranarray(a, dtype=np.float64)
        i32 = lambda a: np.asfortranarray(a, dtype=np.int32)
        V, Tricc = f64(V), f64(Tricc)
        nC, C_, niTri, iTri, VBI, Tri, TriC = (i32(a) for a in (nC, C_, niTri, iTri, VBI, Tri, TriC))
        p = lambda a: a.ctypes.data_as(C.c_void_p)
        return cls(lib().ufe_host_mesh_from_primary(V.shape[0], Tri.shape[0], C_.shape[1], nz, zeta_kind, zeta_R, xmin, xmax, ymin, ymax,
                                                    p(V), p(nC), p(C_), p(niTri), p(iTri), p(VBI), p(Tri), p(Tricc), p(TriC)))

    def halos(self, rank, nranks):
        out = (C.c_int * 36)()
        lib().ufe_host_determine_halos(self._h, rank, nranks, out)
        keys = ["i1", "i2", "i1_nih", "i2_nih", "i1_hle", "i2_hle", "i1_hli", "i2_hli", "i1_hre", "i2_hre", "i1_hri", "i2_hri"]
        o = list(out)
        return {g: dict(zip(keys, o[12 * k:12 * k + 12])) for k, g in enumerate("abc")}

    def __del__(self):
        if getattr(self, "_h", None):
            try:
                lib().ufe_host_mesh_free(self._h)
            except Exception:
                pass
            self._h = None


def partition_list(ntot, i, n):
    a, b = C.c_int(), C.c_int()
    lib().ufe_host_partition_list(ntot, i, n, C.byref(a), C.byref(b))
    return a.value, b.value
