"""ctypes view of the C ABI (include/tinyfk_b200.h) -- what a non-Python host (or the reference's own
wrapper.cpp) would bind.  Used by the parity tests and by bench.py for the device-resident entry
(tfk_solve_fk_device takes raw device pointers and a cudaStream_t; torch only supplies the memory).
"""
import ctypes
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("TFK_B200_LIB") or os.path.join(_HERE, "libtinyfk_b200.so")  # (env: A/B builds of the library)

_vp = ctypes.c_void_p
_ip = ctypes.POINTER(ctypes.c_int)
_dp = ctypes.POINTER(ctypes.c_double)
_i64 = ctypes.c_int64

SYMBOLS = [
    "tfk_last_error", "tfk_model_create", "tfk_model_destroy", "tfk_num_links", "tfk_num_joints",
    "tfk_root_link_id", "tfk_total_mass", "tfk_link_name", "tfk_joint_name", "tfk_get_link_ids",
    "tfk_get_joint_ids", "tfk_get_joint_limits", "tfk_get_joint_velocity_limits",
    "tfk_get_joint_effort_limits", "tfk_link_info", "tfk_is_relevant", "tfk_set_joint_angles", "tfk_get_joint_angles",
    "tfk_set_base_pose", "tfk_get_base_pose", "tfk_add_new_link", "tfk_solve_fk_host",
    "tfk_solve_fk_device", "tfk_solve_fk_device_noalloc", "tfk_inter_link_sqdists_host", "tfk_inter_link_sqdists_device", "tfk_solve_com_fk_host", "tfk_solve_com_fk_device", "tfk_total_inertia_matrix_host", "tfk_total_inertia_matrix_device", "tfk_specialize", "tfk_set_jit_mode", "tfk_get_jit_mode", "tfk_device_count", "tfk_host_alloc", "tfk_host_free",
    "tfk_kernel_launch_count", "tfk_last_launch_info", "tfk_last_launch_variant", "tfk_plan_cache_size", "tfk_model_tables_from_cache", "tfk_version",
]

_lib = None


def lib():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError("tinyfk_b200: %s is not built (make -C tinyfk_b200/csrc); there is no CPU fallback" % LIB_PATH)
    L = ctypes.CDLL(LIB_PATH)
    L.tfk_last_error.restype = ctypes.c_char_p
    L.tfk_version.restype = ctypes.c_char_p
    L.tfk_model_create.argtypes = [ctypes.c_char_p, ctypes.c_size_t, ctypes.c_int, ctypes.POINTER(_vp)]
    L.tfk_model_destroy.argtypes = [_vp]
    L.tfk_model_destroy.restype = None
    for n in ("tfk_num_links", "tfk_num_joints", "tfk_root_link_id"):
        getattr(L, n).argtypes = [_vp]
    L.tfk_total_mass.argtypes = [_vp]
    L.tfk_total_mass.restype = ctypes.c_double
    for n in ("tfk_link_name", "tfk_joint_name"):
        getattr(L, n).argtypes = [_vp, ctypes.c_int]
        getattr(L, n).restype = ctypes.c_char_p
    for n in ("tfk_get_link_ids", "tfk_get_joint_ids"):
        getattr(L, n).argtypes = [_vp, ctypes.POINTER(ctypes.c_char_p), ctypes.c_int, _ip]
    L.tfk_get_joint_limits.argtypes = [_vp, _ip, ctypes.c_int, _dp, _dp]
    L.tfk_get_joint_velocity_limits.argtypes = [_vp, _ip, ctypes.c_int, _dp]
    L.tfk_get_joint_effort_limits.argtypes = [_vp, _ip, ctypes.c_int, _dp]
    L.tfk_link_info.argtypes = [_vp, ctypes.c_int, _ip, _ip, _ip, _dp, _dp]
    L.tfk_is_relevant.argtypes = [_vp, ctypes.c_int, ctypes.c_int, _ip]
    L.tfk_set_joint_angles.argtypes = [_vp, _ip, ctypes.c_int, _dp]
    L.tfk_get_joint_angles.argtypes = [_vp, _ip, ctypes.c_int, _dp]
    L.tfk_set_base_pose.argtypes = [_vp, _dp]
    L.tfk_get_base_pose.argtypes = [_vp, _dp]
    L.tfk_add_new_link.argtypes = [_vp, ctypes.c_char_p, ctypes.c_int, _dp, _dp, _ip]
    L.tfk_solve_fk_host.argtypes = [_vp, _vp, _i64, _ip, ctypes.c_int, _ip, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                    ctypes.c_int, ctypes.c_int, _ip, ctypes.c_int, _vp, _vp]
    L.tfk_solve_fk_device.argtypes = [_vp, ctypes.c_int, _vp, _vp, _i64, _ip, ctypes.c_int, _ip, ctypes.c_int,
                                      ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int, _vp, _vp, _i64, _i64]
    L.tfk_solve_fk_device_noalloc.argtypes = [_vp, ctypes.c_int, _vp, _vp, _i64, _ip, ctypes.c_int, _ip, ctypes.c_int,
                                              ctypes.c_int, ctypes.c_int, ctypes.c_int, _vp, _vp, _i64, _i64]
    L.tfk_inter_link_sqdists_host.argtypes = [_vp, _vp, _i64, _ip, _ip, ctypes.c_int, _ip, ctypes.c_int, ctypes.c_int,
                                              ctypes.c_int, ctypes.c_int, _vp, _vp]
    L.tfk_inter_link_sqdists_device.argtypes = [_vp, ctypes.c_int, _vp, _vp, _i64, _ip, _ip, ctypes.c_int, _ip, ctypes.c_int,
                                                ctypes.c_int, ctypes.c_int, ctypes.c_int, _vp, _vp, _i64, _i64]
    L.tfk_solve_com_fk_host.argtypes = [_vp, _vp, _i64, _ip, ctypes.c_int, _ip, _dp, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                        ctypes.c_int, _vp, _vp]
    L.tfk_solve_com_fk_device.argtypes = [_vp, ctypes.c_int, _vp, _vp, _i64, _ip, ctypes.c_int, _ip, _dp, ctypes.c_int,
                                          ctypes.c_int, ctypes.c_int, ctypes.c_int, _vp, _vp, _i64, _i64]
    L.tfk_total_inertia_matrix_host.argtypes = [_vp, _vp, _i64, _ip, ctypes.c_int, ctypes.c_int, _vp]
    L.tfk_total_inertia_matrix_device.argtypes = [_vp, ctypes.c_int, _vp, _vp, _i64, _ip, ctypes.c_int, ctypes.c_int, _vp, _i64, _i64]
    L.tfk_specialize.argtypes = [_vp, _ip, ctypes.c_int, _ip, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                 ctypes.c_int, ctypes.POINTER(ctypes.c_int64)]
    L.tfk_set_jit_mode.argtypes = [_vp, ctypes.c_int]
    L.tfk_get_jit_mode.argtypes = [_vp]
    L.tfk_device_count.argtypes = [_ip]
    L.tfk_host_alloc.argtypes = [ctypes.c_size_t, ctypes.POINTER(_vp)]
    L.tfk_host_free.argtypes = [_vp]
    L.tfk_kernel_launch_count.restype = _i64
    L.tfk_last_launch_info.argtypes = [_vp, _ip, _ip, _ip, _ip, _ip]
    L.tfk_plan_cache_size.argtypes = [_vp]
    L.tfk_model_tables_from_cache.argtypes = [_vp]
    _lib = L
    return L


class TfkError(RuntimeError):
    pass


def _check(rc):
    if rc == 0:
        return
    msg = lib().tfk_last_error().decode()
    if rc == 1:
        raise ValueError(msg)
    raise TfkError(msg)


def _ints(v):
    return np.ascontiguousarray(np.asarray(v, dtype=np.int32).reshape(-1))


def pose_dim(rot_type):
    return {0: 3, 1: 6, 2: 7}[int(rot_type)]


def base_dof(base_type):
    return {0: 0, 1: 3, 2: 6}[int(base_type)]


class PinnedArray:
    """F-ordered numpy array on pooled pinned host memory (tfk_host_alloc)."""

    def __init__(self, shape, dtype=np.float64, order="F"):
        self._ptr = _vp()
        nbytes = int(np.prod(shape)) * np.dtype(dtype).itemsize
        _check(lib().tfk_host_alloc(max(nbytes, 1), ctypes.byref(self._ptr)))
        buf = (ctypes.c_char * max(nbytes, 1)).from_address(self._ptr.value)
        self.array = np.frombuffer(buf, dtype=dtype, count=int(np.prod(shape))).reshape(shape, order=order)

    def free(self):
        if self._ptr:
            self.array = None
            lib().tfk_host_free(self._ptr)
            self._ptr = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


class Model:
    """Thin object wrapper over a tfk_model handle."""

    def __init__(self, xml_text, host_libm=0):
        self._L = lib()
        self._h = _vp()
        b = xml_text.encode()
        _check(self._L.tfk_model_create(b, len(b), int(host_libm), ctypes.byref(self._h)))

    def __del__(self):
        if getattr(self, "_h", None):
            self._L.tfk_model_destroy(self._h)
            self._h = None

    n_links = property(lambda s: s._L.tfk_num_links(s._h))
    n_joints = property(lambda s: s._L.tfk_num_joints(s._h))
    root_link_id = property(lambda s: s._L.tfk_root_link_id(s._h))
    total_mass = property(lambda s: s._L.tfk_total_mass(s._h))

    def get_link_names(self):
        return [self._L.tfk_link_name(self._h, i).decode() for i in range(self.n_links)]

    def get_joint_names(self):
        return [self._L.tfk_joint_name(self._h, i).decode() for i in range(self.n_joints)]

    def _ids(self, fn, names):
        arr = (ctypes.c_char_p * len(names))(*[n.encode() for n in names])
        out = np.zeros(len(names), dtype=np.int32)
        _check(fn(self._h, arr, len(names), out.ctypes.data_as(_ip)))
        return [int(v) for v in out]

    def get_link_ids(self, names):
        return self._ids(self._L.tfk_get_link_ids, names)

    def get_joint_ids(self, names):
        return self._ids(self._L.tfk_get_joint_ids, names)

    def get_joint_limits(self, joint_ids):
        ids = _ints(joint_ids)
        lo, hi = np.zeros(ids.size), np.zeros(ids.size)
        _check(self._L.tfk_get_joint_limits(self._h, ids.ctypes.data_as(_ip), ids.size, lo.ctypes.data_as(_dp),
                                            hi.ctypes.data_as(_dp)))
        return list(zip(lo.tolist(), hi.tolist()))

    def link_info(self, link_id):
        p, t, j = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
        axis, origin = np.zeros(3), np.zeros(7)
        _check(self._L.tfk_link_info(self._h, int(link_id), ctypes.byref(p), ctypes.byref(t), ctypes.byref(j),
                                     axis.ctypes.data_as(_dp), origin.ctypes.data_as(_dp)))
        return p.value, t.value, j.value, axis, origin

    def export_table(self):
        """Link table in the layout oracle/fk_oracle.h takes (test use)."""
        n = self.n_links
        info = [self.link_info(i) for i in range(n)]
        return dict(parent=np.array([x[0] for x in info], dtype=np.int32),
                    jtype=np.array([x[1] for x in info], dtype=np.int32),
                    jid=np.array([x[2] for x in info], dtype=np.int32), axis=np.array([x[3] for x in info]),
                    origin=np.array([x[4] for x in info]), n_joints=self.n_joints, root=self.root_link_id)

    def is_relevant(self, link_id, joint_id):
        r = ctypes.c_int()
        _check(self._L.tfk_is_relevant(self._h, int(link_id), int(joint_id), ctypes.byref(r)))
        return bool(r.value)

    def set_joint_angles(self, joint_ids, angles):
        ids = _ints(joint_ids)
        a = np.ascontiguousarray(angles, dtype=np.float64)
        assert a.size == ids.size
        _check(self._L.tfk_set_joint_angles(self._h, ids.ctypes.data_as(_ip), ids.size, a.ctypes.data_as(_dp)))

    def get_joint_angles(self, joint_ids):
        ids = _ints(joint_ids)
        a = np.zeros(ids.size)
        _check(self._L.tfk_get_joint_angles(self._h, ids.ctypes.data_as(_ip), ids.size, a.ctypes.data_as(_dp)))
        return a

    def set_base_pose(self, xyzrpy):
        a = np.ascontiguousarray(xyzrpy, dtype=np.float64)
        if a.size != 6:
            raise ValueError("argument must have size 6")
        _check(self._L.tfk_set_base_pose(self._h, a.ctypes.data_as(_dp)))

    def get_base_pose(self):
        a = np.zeros(6)
        _check(self._L.tfk_get_base_pose(self._h, a.ctypes.data_as(_dp)))
        return a

    def add_new_link(self, name, parent_id, pos, rpy):
        pos = np.ascontiguousarray(pos, dtype=np.float64)
        rpy = np.ascontiguousarray(rpy, dtype=np.float64)
        new_id = ctypes.c_int()
        _check(self._L.tfk_add_new_link(self._h, name.encode(), int(parent_id), pos.ctypes.data_as(_dp),
                                        rpy.ctypes.data_as(_dp), ctypes.byref(new_id)))
        return new_id.value

    def solve_fk(self, qs, elink_ids, joint_ids, rot_type=0, base_type=0, with_jacobian=False, precision=0,
                 devices=None, out=None):
        """Host buffers through tfk_solve_fk_host.  Returns F-ordered (P, J)."""
        dt = np.float32 if precision == 1 else np.float64
        qs = np.ascontiguousarray(np.atleast_2d(np.asarray(qs, dtype=dt)))
        l, j = _ints(elink_ids), _ints(joint_ids)
        n, n_in = qs.shape
        if n_in != j.size + base_dof(base_type):
            raise ValueError("dof mismatch")
        pd = pose_dim(rot_type)
        if out is None:
            P = np.zeros((n * l.size, pd), dtype=dt, order="F")
            J = np.zeros((n * l.size * pd, n_in), dtype=dt, order="F")
        else:
            P, J = out
        dv = _ints(devices) if devices else None
        _check(self._L.tfk_solve_fk_host(self._h, qs.ctypes.data, n, l.ctypes.data_as(_ip), l.size,
                                         j.ctypes.data_as(_ip), j.size, int(rot_type), int(base_type),
                                         int(bool(with_jacobian)), int(precision),
                                         dv.ctypes.data_as(_ip) if dv is not None else None,
                                         dv.size if dv is not None else 0, P.ctypes.data,
                                         J.ctypes.data if with_jacobian else None))
        return P, J

    def inter_link_sqdists(self, qs, link_id_pairs, joint_ids, base_type=0, with_jacobian=False, precision=0):
        """Host buffers through tfk_inter_link_sqdists_host.  Returns (sqdists [n*n_check], grads F-ordered)."""
        dt = np.float32 if precision == 1 else np.float64
        qs = np.ascontiguousarray(np.atleast_2d(np.asarray(qs, dtype=dt)))
        l1, l2, j = _ints([p[0] for p in link_id_pairs]), _ints([p[1] for p in link_id_pairs]), _ints(joint_ids)
        n, n_in = qs.shape
        if n_in != j.size + base_dof(base_type):
            raise ValueError("dof mismatch")
        sq = np.zeros(n * l1.size, dtype=dt)
        G = np.zeros((n * l1.size, n_in), dtype=dt, order="F") if with_jacobian else np.zeros((1, 1), dtype=dt)
        _check(self._L.tfk_inter_link_sqdists_host(self._h, qs.ctypes.data, n, l1.ctypes.data_as(_ip), l2.ctypes.data_as(_ip),
                                                   l1.size, j.ctypes.data_as(_ip), j.size, int(base_type),
                                                   int(bool(with_jacobian)), int(precision), sq.ctypes.data,
                                                   G.ctypes.data if with_jacobian else None))
        return sq, G

    def inter_link_sqdists_device(self, device, stream, d_q, n, link_id_pairs, joint_ids, base_type, with_jacobian, precision,
                                  d_sq, d_G, ld_cfg=None, row0=0):
        l1, l2, j = _ints([p[0] for p in link_id_pairs]), _ints([p[1] for p in link_id_pairs]), _ints(joint_ids)
        _check(self._L.tfk_inter_link_sqdists_device(self._h, int(device), _vp(stream), _vp(d_q), int(n),
                                                     l1.ctypes.data_as(_ip), l2.ctypes.data_as(_ip), l1.size,
                                                     j.ctypes.data_as(_ip), j.size, int(base_type), int(bool(with_jacobian)),
                                                     int(precision), _vp(d_sq), _vp(d_G) if d_G else None,
                                                     int(ld_cfg if ld_cfg is not None else n), int(row0)))

    def specialize(self, elink_ids, joint_ids, rot_type, base_type, with_jacobian, precision=0, wait=True, compile_only=False):
        """tfk_specialize: compile the single-link kernel against this query's plan.  Returns the info dict."""
        l, j = _ints(elink_ids), _ints(joint_ids)
        info = (ctypes.c_int64 * 6)()
        _check(self._L.tfk_specialize(self._h, l.ctypes.data_as(_ip), l.size, j.ctypes.data_as(_ip), j.size, int(rot_type),
                                      int(base_type), int(bool(with_jacobian)), int(precision),
                                      (1 if wait else 0) | (2 if compile_only else 0), info))
        return dict(zip(("state", "block", "blocks_per_sm", "smem_bytes", "cubin_bytes", "from_cache"), [int(x) for x in info]))

    def set_jit_mode(self, mode):
        _check(self._L.tfk_set_jit_mode(self._h, int(mode)))

    def get_jit_mode(self):
        return int(self._L.tfk_get_jit_mode(self._h))

    def solve_com_fk(self, qs, joint_ids, action_link_ids=(), forces=(), base_type=0, with_jacobian=False, precision=0):
        """Host buffers through tfk_solve_com_fk_host.  Returns (com (n, 3), J (3n, n_joint + 0|6) F-ordered or None)."""
        dt = np.float32 if precision == 1 else np.float64
        qs = np.ascontiguousarray(np.atleast_2d(np.asarray(qs, dtype=dt)))
        j, al = _ints(joint_ids), _ints(action_link_ids)
        f = np.ascontiguousarray(np.asarray(forces, dtype=np.float64).reshape(-1))
        if f.size != al.size:
            raise ValueError("action_link_ids and forces differ in length")
        n, n_in = qs.shape
        if n_in != j.size + base_dof(base_type):
            raise ValueError("dof mismatch")
        n_dof = j.size + (6 if base_type else 0)
        com = np.zeros((n, 3), dtype=dt)
        J = np.zeros((3 * n, n_dof), dtype=dt, order="F") if with_jacobian else None
        _check(self._L.tfk_solve_com_fk_host(self._h, qs.ctypes.data, n, j.ctypes.data_as(_ip), j.size, al.ctypes.data_as(_ip),
                                             f.ctypes.data_as(_dp), al.size, int(base_type), int(bool(with_jacobian)),
                                             int(precision), com.ctypes.data, J.ctypes.data if with_jacobian else None))
        return com, J

    def total_inertia_matrix(self, qs, joint_ids, precision=0):
        """Host buffers through tfk_total_inertia_matrix_host.  Returns (n, 3, 3)."""
        dt = np.float32 if precision == 1 else np.float64
        qs = np.ascontiguousarray(np.atleast_2d(np.asarray(qs, dtype=dt)))
        j = _ints(joint_ids)
        n, n_in = qs.shape
        if n_in != j.size:
            raise ValueError("dof mismatch")
        out = np.zeros((n, 3, 3), dtype=dt)
        _check(self._L.tfk_total_inertia_matrix_host(self._h, qs.ctypes.data, n, j.ctypes.data_as(_ip), j.size, int(precision),
                                                     out.ctypes.data))
        return out

    def solve_com_fk_device(self, device, stream, d_q, n, joint_ids, action_link_ids, forces, base_type, with_jacobian,
                            precision, d_com, d_J, ld_cfg=None, row0=0):
        j, al = _ints(joint_ids), _ints(action_link_ids)
        f = np.ascontiguousarray(np.asarray(forces, dtype=np.float64).reshape(-1))
        _check(self._L.tfk_solve_com_fk_device(self._h, int(device), _vp(stream), _vp(d_q), int(n), j.ctypes.data_as(_ip), j.size,
                                               al.ctypes.data_as(_ip), f.ctypes.data_as(_dp), al.size, int(base_type),
                                               int(bool(with_jacobian)), int(precision), _vp(d_com), _vp(d_J) if d_J else None,
                                               int(ld_cfg if ld_cfg is not None else n), int(row0)))

    def solve_fk_device(self, device, stream, d_q, n, elink_ids, joint_ids, rot_type, base_type, with_jacobian,
                        precision, d_P, d_J, ld_cfg=None, row0=0):
        """Raw device pointers (ints) through tfk_solve_fk_device; asynchronous on `stream`."""
        l, j = _ints(elink_ids), _ints(joint_ids)
        _check(self._L.tfk_solve_fk_device(self._h, int(device), _vp(stream), _vp(d_q), int(n), l.ctypes.data_as(_ip),
                                           l.size, j.ctypes.data_as(_ip), j.size, int(rot_type), int(base_type),
                                           int(bool(with_jacobian)), int(precision), _vp(d_P),
                                           _vp(d_J) if d_J else None, int(ld_cfg if ld_cfg is not None else n),
                                           int(row0)))

    def solve_fk_device_noalloc(self, device, stream, d_q, n, elink_ids, joint_ids, rot_type, base_type, with_jacobian,
                                d_P, d_J, ld_cfg=None, row0=0):
        """tfk_solve_fk_device_noalloc: strict FP64, scratch in a static device array, no device-memory allocation"""
        l, j = _ints(elink_ids), _ints(joint_ids)
        _check(self._L.tfk_solve_fk_device_noalloc(self._h, int(device), _vp(stream), _vp(d_q), int(n), l.ctypes.data_as(_ip),
                                                   l.size, j.ctypes.data_as(_ip), j.size, int(rot_type), int(base_type),
                                                   int(bool(with_jacobian)), _vp(d_P), _vp(d_J) if d_J else None,
                                                   int(ld_cfg if ld_cfg is not None else n), int(row0)))

    def plan_cache_size(self):
        return int(self._L.tfk_plan_cache_size(self._h))

    def tables_from_cache(self):
        return bool(self._L.tfk_model_tables_from_cache(self._h))

    def last_launch_info(self):
        v = [ctypes.c_int() for _ in range(5)]
        _check(self._L.tfk_last_launch_info(self._h, *[ctypes.byref(x) for x in v]))
        d = dict(zip(("grid", "block", "smem_bytes", "blocks_per_sm", "rec_reals"), [x.value for x in v]))
        self._L.tfk_last_launch_variant.argtypes = [_vp]
        k = int(self._L.tfk_last_launch_variant(self._h))
        d["kernel"] = {0: "fk_fused_kernel<smem records>", 1: "fk_fused_kernel<global records>",
                       2: "fk_twophase_kernel", 3: "fk_fused_kernel<smem records, 5 blocks/SM>", 4: "fk_sqdist_kernel", 5: "fk_com_kernel", 6: "fk_inertia_kernel", 9: "fk_sqdist_kernel<joint records in shared memory>", 8: "fk_twophase_kernel<joint records in shared memory>", 7: "tfk_spec_kernel (plan-specialised fk_fused_kernel, NVRTC)", 12: "tfk_spec_kernel (plan-specialised fk_twophase_kernel, NVRTC)"}.get(k, "none")
        return d


def kernel_launch_count():
    return int(lib().tfk_kernel_launch_count())


def device_count():
    c = ctypes.c_int()
    _check(lib().tfk_device_count(ctypes.byref(c)))
    return c.value
