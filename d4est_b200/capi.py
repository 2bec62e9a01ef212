"""ctypes binding of the C ABI (include/d4est_b200.h) -- the same stub a d4est maintainer's
Python tooling would use.  No compute happens in Python; there is no CPU
fallback: loading fails loudly when libd4est_b200.so is missing, and every compute call
fails with D4B_ENODEV when no CUDA device is present.
"""
import ctypes as C
import os
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libd4est_b200.so")


class D4BError(RuntimeError):
    pass


class HM(C.Structure):  # d4b_hm_t
    _fields_ = [("nbr", C.c_int32), ("f_m", C.c_int8), ("f_p", C.c_int8), ("tcode", C.c_int8),
                ("child_m", C.c_int8), ("child_p", C.c_int8), ("degq", C.c_int8),
                ("face_slot", C.c_int8), ("pad", C.c_int8), ("pad0", C.c_int32),
                ("fac_off", C.c_int64), ("g_off", C.c_int64), ("pad2", C.c_int64)]


_P = C.c_void_p


class MeshDesc(C.Structure):  # d4b_mesh_desc_t
    _fields_ = [("n_local", C.c_int32), ("n_ghost", C.c_int32), ("quad_type", C.c_int32),
                ("reserved", C.c_int32), ("deg", _P), ("node_off", _P), ("degq", _P), ("quad_off", _P),
                ("G", _P), ("hm_begin", _P), ("hm", _P), ("fac", _P), ("n_fac", C.c_int64),
                ("bc_g", _P), ("n_bc_g", C.c_int64), ("n_peers", C.c_int32), ("reserved2", C.c_int32),
                ("peer_rank", _P), ("send_begin", _P), ("send_elem", _P), ("recv_begin", _P)]


class BrickParams(C.Structure):  # d4b_brick_params_t
    _fields_ = [("nx", C.c_int32), ("ny", C.c_int32), ("nz", C.c_int32), ("deg", C.c_int32),
                ("degq", C.c_int32), ("quad_type", C.c_int32),
                ("x0", C.c_double), ("x1", C.c_double), ("y0", C.c_double), ("y1", C.c_double),
                ("z0", C.c_double), ("z1", C.c_double), ("penalty_prefactor", C.c_double),
                ("face_h", C.c_double)]


# every symbol include/d4est_b200.h declares: name -> (restype, argtypes)
SYMBOLS = {
    "d4b_last_error": (C.c_char_p, []),
    "d4b_version": (C.c_char_p, []),
    "d4b_device_count": (C.c_int, [C.POINTER(C.c_int)]),
    "d4b_set_device": (C.c_int, [C.c_int]),
    "d4b_mesh_create": (C.c_int, [C.POINTER(MeshDesc), C.POINTER(_P)]),
    "d4b_mesh_destroy": (None, [_P]),
    "d4b_brick_create": (C.c_int, [C.POINTER(BrickParams), C.POINTER(_P)]),
    "d4b_brick_create_slab": (C.c_int, [C.POINTER(BrickParams), C.c_int, C.c_int, C.POINTER(_P)]),
    "d4b_mesh_local_nodes": (C.c_int64, [_P]),
    "d4b_mesh_ghost_nodes": (C.c_int64, [_P]),
    "d4b_mesh_bytes": (C.c_int64, [_P]),
    "d4b_mesh_algorithmic_bytes": (C.c_int64, [_P]),
    "d4b_apply_host": (C.c_int, [_P, _P, _P]),
    "d4b_apply_dev": (C.c_int, [_P, _P, _P, _P]),
    "d4b_set_ghost_host": (C.c_int, [_P, _P]),
    "d4b_kernel_launches": (C.c_int64, []),
    "d4b_mesh_kernel_split": (C.c_int, [_P, C.POINTER(C.c_int64)]),
    "d4b_dev_alloc": (C.c_int, [C.c_size_t, C.POINTER(_P)]),
    "d4b_dev_free": (C.c_int, [_P]),
    "d4b_h2d": (C.c_int, [_P, _P, C.c_size_t, _P]),
    "d4b_d2h": (C.c_int, [_P, _P, C.c_size_t, _P]),
    "d4b_sync": (C.c_int, [_P]),
    "d4b_host_alloc_pinned": (C.c_int, [C.c_size_t, C.POINTER(_P)]),
    "d4b_host_free_pinned": (C.c_int, [_P]),
    "d4b_comm_unique_id": (C.c_int, [C.c_char_p]),
    "d4b_comm_init": (C.c_int, [C.c_int, C.c_int, C.c_char_p]),
    "d4b_comm_finalize": (C.c_int, []),
    "d4b_comm_size": (C.c_int, []),
    "d4b_halo_exchange": (C.c_int, [_P, _P, _P]),
    "d4b_vec_axpy": (C.c_int, [C.c_int64, C.c_double, _P, _P, _P]),
    "d4b_vec_xpby": (C.c_int, [C.c_int64, _P, C.c_double, _P, _P]),
    "d4b_vec_axpyeqz": (C.c_int, [C.c_int64, C.c_double, _P, _P, _P, _P]),
    "d4b_vec_scale": (C.c_int, [C.c_int64, C.c_double, _P, _P]),
    "d4b_vec_copy": (C.c_int, [C.c_int64, _P, _P, _P]),
    "d4b_vec_zero": (C.c_int, [C.c_int64, _P, _P]),
    "d4b_vec_dot": (C.c_int, [C.c_int64, _P, _P, C.POINTER(C.c_double), _P]),
    "d4b_vec_dot_noalloc": (C.c_int, [C.c_int64, _P, _P, _P, _P]),
    "d4b_cheby_iterate": (C.c_int, [_P, _P, _P, _P, C.c_int, C.c_double, C.c_double, C.c_int, _P, _P]),
    "d4b_cg_eigs": (C.c_int, [_P, _P, _P, C.c_int, C.c_int, C.POINTER(C.c_double), _P, _P]),
    "d4b_cg_solve": (C.c_int, [_P, _P, _P, C.c_int, C.c_double, C.c_double, C.POINTER(C.c_int), C.POINTER(C.c_double), _P, _P]),
    "d4b_fcg_solve": (C.c_int, [_P, _P, _P, C.c_int, C.c_double, C.c_double, _P, _P, C.POINTER(C.c_int), C.POINTER(C.c_double), _P, _P]),
    "d4b_set_option": (C.c_int, [C.c_char_p, C.c_int]),
    "d4b_mg_transfer_create": (C.c_int, [C.c_int, _P, _P, _P, _P, _P, C.POINTER(_P)]),
    "d4b_mg_transfer_destroy": (None, [_P]),
    "d4b_mg_restrict": (C.c_int, [_P, _P, _P, _P]),
    "d4b_mg_prolong": (C.c_int, [_P, _P, _P, _P]),
    "d4b_amr_transfer_create": (C.c_int, [C.c_int, _P, _P, C.POINTER(_P), C.POINTER(C.c_int64), C.POINTER(C.c_int64)]),
    "d4b_amr_interpolate_field_sizes": (C.c_int, [C.c_int, _P, _P, C.c_int, _P, C.POINTER(C.c_int64), C.POINTER(C.c_int64), C.POINTER(C.c_int64)]),
    "d4b_amr_interpolate_field_dev": (C.c_int, [C.c_int, _P, _P, C.c_int, _P, _P, _P, _P, _P]),
    "d4b_amr_interpolate_field_host": (C.c_int, [C.c_int, _P, _P, C.c_int, _P, _P, C.c_int64, _P, C.c_int64]),
    "d4b_mg_create": (C.c_int, [C.c_int, C.POINTER(_P), C.POINTER(_P), _P, C.POINTER(_P)]),
    "d4b_mg_destroy": (None, [_P]),
    "d4b_mg_get_eigs": (C.c_int, [_P, _P]),
    "d4b_mg_vcycle": (C.c_int, [_P, _P, _P, C.POINTER(C.c_double), _P]),
    "d4b_mg_solve": (C.c_int, [_P, _P, _P, C.POINTER(C.c_int), C.POINTER(C.c_double), _P]),
    "d4b_mg_pc_apply": (C.c_int, [_P, _P, _P, _P]),
    "d4b_mg_num_levels": (C.c_int, [_P]),
    "d4b_residual_dev": (C.c_int, [_P, _P, _P, _P, _P]),
    "d4b_residual_host": (C.c_int, [_P, _P, _P, _P]),
    "d4b_cheby_smooth_host": (C.c_int, [_P, _P, _P, _P, C.c_int, C.c_int, C.c_double, C.c_double, C.c_int, C.c_int, C.POINTER(C.c_double)]),
    "d4b_mass_set_coef": (C.c_int, [_P, _P]),
    "d4b_mass_apply_dev": (C.c_int, [_P, _P, _P, C.c_int, _P]),
    "d4b_mass_apply_host": (C.c_int, [_P, _P, _P, C.c_int]),
    "d4b_apply_jac_dev": (C.c_int, [_P, _P, _P, _P]),
    "d4b_mg_pc_apply_host": (C.c_int, [_P, _P, _P]),
    "d4b_mg_fcg_solve_host": (C.c_int, [_P, _P, _P, C.c_int, C.c_double, C.c_double, C.POINTER(C.c_int), C.POINTER(C.c_double)]),
    "d4b_mg_fcg_solve": (C.c_int, [_P, _P, _P, C.c_int, C.c_double, C.c_double, C.POINTER(C.c_int), C.POINTER(C.c_double), _P, _P]),
}


class MGParams(C.Structure):  # d4b_mg_params_t
    _fields_ = [("vcycle_imax", C.c_int32), ("vcycle_rtol", C.c_double), ("vcycle_atol", C.c_double),
                ("cheby_imax", C.c_int32), ("cheby_eigs_cg_imax", C.c_int32),
                ("cheby_eigs_lmax_lmin_ratio", C.c_double), ("cheby_eigs_max_multiplier", C.c_double),
                ("eigs_reuse_fromdownvcycle", C.c_int32), ("eigs_reuse_fromlastvcycle", C.c_int32),
                ("use_new_cg_eigs", C.c_int32), ("bottom_kind", C.c_int32), ("bottom_imax", C.c_int32),
                ("bottom_eigs_cg_imax", C.c_int32), ("bottom_eigs_lmax_lmin_ratio", C.c_double),
                ("bottom_eigs_max_multiplier", C.c_double), ("bottom_atol", C.c_double), ("bottom_rtol", C.c_double)]


class DevVec:
    """n doubles in HBM (d4b_dev_alloc); host copies only through .set / .get."""

    def __init__(self, n, init=None):
        self.n = int(n); self.ptr = _P()
        check(lib().d4b_dev_alloc(self.n * 8, C.byref(self.ptr)))
        if init is not None:
            self.set(init)
        else:
            check(lib().d4b_vec_zero(self.n, self.ptr, None))

    def set(self, a):
        a = np.ascontiguousarray(a, np.float64); assert a.size <= self.n
        check(lib().d4b_h2d(self.ptr, _ptr(a), a.size * 8, None)); check(lib().d4b_sync(None))

    def get(self, n=None):
        out = np.empty(self.n if n is None else n)
        check(lib().d4b_d2h(out.ctypes.data_as(_P), self.ptr, out.size * 8, None)); check(lib().d4b_sync(None))
        return out

    def __del__(self):
        try:
            if self.ptr:
                lib().d4b_dev_free(self.ptr)
        except Exception:
            pass


class Transfer:
    def __init__(self, tab):
        self._k = [np.ascontiguousarray(tab[k], dt) for k, dt in (("kind", np.int32), ("degH", np.int32),
                   ("coarse_off", np.int64), ("child_deg", np.int32), ("child_off", np.int64))]
        self.handle = _P()
        check(lib().d4b_mg_transfer_create(len(self._k[0]), *[_ptr(a) for a in self._k], C.byref(self.handle)))


class Multigrid:
    """levels: list of MeshLevel, coarsest first; transfers[l] joins level l and l+1."""

    def __init__(self, levels, transfers, **kw):
        p = MGParams(vcycle_imax=1, vcycle_rtol=1e-15, vcycle_atol=0., cheby_imax=15, cheby_eigs_cg_imax=15,
                     cheby_eigs_lmax_lmin_ratio=30., cheby_eigs_max_multiplier=1.1, eigs_reuse_fromdownvcycle=1,
                     eigs_reuse_fromlastvcycle=0, use_new_cg_eigs=0, bottom_kind=0, bottom_imax=15,
                     bottom_eigs_cg_imax=15, bottom_eigs_lmax_lmin_ratio=30., bottom_eigs_max_multiplier=1.1,
                     bottom_atol=1e-15, bottom_rtol=1e-15)
        for k, v in kw.items():
            setattr(p, k, v)
        self.params, self.levels, self.transfers = p, levels, transfers
        L = (_P * len(levels))(*[l.handle for l in levels]); T = (_P * max(1, len(transfers)))(*[t.handle for t in transfers])
        self.handle = _P()
        check(lib().d4b_mg_create(len(levels), L, T, C.byref(p), C.byref(self.handle)))

_lib = None


def _preload_nccl():
    """libd4est_b200.so needs libnccl.so.2.  When the process also hosts torch (tests, bench), both must bind
    to ONE NCCL: load the copy bundled with torch's wheels first (nvidia/nccl/lib) so that a later `import torch`
    does not meet the older system library; without that wheel the system libnccl is used."""
    try:
        import importlib.util
        spec = importlib.util.find_spec("nvidia")
        for base in (spec.submodule_search_locations if spec else []):
            cand = os.path.join(base, "nccl", "lib", "libnccl.so.2")
            if os.path.exists(cand):
                C.CDLL(cand, mode=C.RTLD_GLOBAL)
                return
    except Exception:
        pass


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise D4BError(f"{LIB_PATH} not built: run `python -c 'import __graft_entry__ as g; g.build()'` "
                           "(there is no CPU fallback)")
        _preload_nccl()
        _lib = C.CDLL(LIB_PATH)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(_lib, name)
            fn.restype, fn.argtypes = res, args
    return _lib


def check(rc):
    if rc != 0:
        raise D4BError(f"d4est_b200 error {rc}: {lib().d4b_last_error().decode()}")


def _ptr(a):
    return None if a is None or a.size == 0 else a.ctypes.data_as(_P)


class MeshLevel:
    """Device-resident mesh level built from flat tables (dict of numpy arrays with the field
    names of d4b_mesh_desc_t, e.g. an oracle fixture or d4est_b200.brick.brick_tables())."""

    def __init__(self, t):
        L = lib()
        self.handle = _P()
        if t is None:
            return
        nl, ng, qt, npeer = [int(v) for v in t["hdr"]]
        self._keep = k = {}
        def arr(name, dt):
            a = np.ascontiguousarray(t.get(name, np.zeros(0, dt)))
            if a.dtype != np.dtype(dt):
                a = a.view(np.uint8) if name == "hm" else a.astype(dt)
            k[name] = a
            return a
        d = MeshDesc()
        d.n_local, d.n_ghost, d.quad_type, d.n_peers = nl, ng, qt, npeer
        d.deg = _ptr(arr("deg", np.int32)); d.node_off = _ptr(arr("node_off", np.int64))
        d.degq = _ptr(arr("degq", np.int32)); d.quad_off = _ptr(arr("quad_off", np.int64))
        d.G = _ptr(arr("G", np.float64)); d.hm_begin = _ptr(arr("hm_begin", np.int32))
        hm = np.ascontiguousarray(t["hm"]); k["hm"] = hm
        d.hm = _ptr(hm.view(np.uint8))
        fac = arr("fac", np.float64); d.fac = _ptr(fac); d.n_fac = fac.size
        g = arr("bc_g", np.float64); d.bc_g = _ptr(g); d.n_bc_g = g.size
        d.peer_rank = _ptr(arr("peer_rank", np.int32)); d.send_begin = _ptr(arr("send_begin", np.int32))
        d.send_elem = _ptr(arr("send_elem", np.int32)); d.recv_begin = _ptr(arr("recv_begin", np.int32))
        if k["send_begin"].size == 0:
            k["send_begin"] = np.zeros(1, np.int32); d.send_begin = _ptr(k["send_begin"])
            k["recv_begin"] = np.zeros(1, np.int32); d.recv_begin = _ptr(k["recv_begin"])
        self.handle = _P()
        check(L.d4b_mesh_create(C.byref(d), C.byref(self.handle)))
        self.local_nodes = int(L.d4b_mesh_local_nodes(self.handle))
        self.ghost_nodes = int(L.d4b_mesh_ghost_nodes(self.handle))
        self.n_local = nl

    @classmethod
    def brick(cls, n, deg=3, degq=None, quad_type=0, prefactor=2.0, box=(0., 1., 0., 1., 0., 1.), face_h=0.0,
              rank=0, nranks=1):
        """Synthetic uniform brick generated on the device (d4b_brick_create); with nranks > 1 the z-slab
        `rank` of the brick nx x ny x (nz*nranks), with ghost planes and halo lists (d4b_brick_create_slab)."""
        nx, ny, nz = (n, n, n) if np.isscalar(n) else n
        p = BrickParams(nx, ny, nz, deg, deg if degq is None else degq, quad_type, *box, prefactor, face_h)
        self = cls(None)
        check(lib().d4b_brick_create_slab(C.byref(p), rank, nranks, C.byref(self.handle)))
        self.local_nodes = int(lib().d4b_mesh_local_nodes(self.handle))
        self.ghost_nodes = int(lib().d4b_mesh_ghost_nodes(self.handle))
        self.n_local = nx * ny * nz
        return self

    @property
    def algorithmic_bytes(self):
        return int(lib().d4b_mesh_algorithmic_bytes(self.handle))

    @property
    def resident_bytes(self):
        return int(lib().d4b_mesh_bytes(self.handle))

    @property
    def kernel_split(self):
        """(elements in the regular-element kernels, in the templated hp kernel, in the completeness kernel)"""
        c = (C.c_int64 * 3)()
        check(lib().d4b_mesh_kernel_split(self.handle, c))
        return tuple(int(v) for v in c)

    def set_ghost(self, ug):
        ug = np.ascontiguousarray(ug, np.float64)
        assert ug.size == self.ghost_nodes
        check(lib().d4b_set_ghost_host(self.handle, _ptr(ug)))

    def apply_host(self, u, out=None):
        """Au = A u with HOST buffers (H2D + kernel + D2H): d4b_apply_host."""
        u = np.ascontiguousarray(u, np.float64)
        assert u.size == self.local_nodes
        Au = np.empty_like(u) if out is None else out
        check(lib().d4b_apply_host(self.handle, _ptr(u), Au.ctypes.data_as(_P)))
        return Au

    def apply_dev(self, u_ptr, Au_ptr, stream=0):
        """Au = A u with DEVICE pointers (ints), on the given CUDA stream: d4b_apply_dev."""
        check(lib().d4b_apply_dev(self.handle, _P(u_ptr), _P(Au_ptr), _P(stream)))

    def close(self):
        if self.handle:
            lib().d4b_mesh_destroy(self.handle)
            self.handle = _P()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
