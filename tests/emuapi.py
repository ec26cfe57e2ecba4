"""ctypes binding of tests/emu/libfk_emu.so (TEST INFRASTRUCTURE: the kernels' per-thread code run on
the CPU through the product's host model + plan flattening; see tests/emu/emu_driver.cpp)."""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "emu", "libfk_emu.so")
_dp = ctypes.POINTER(ctypes.c_double)
_ip = ctypes.POINTER(ctypes.c_int)
_lib = None

MODE_STRICT64, MODE_F32, MODE_NATIVE64, MODE_FAST64 = 0, 1, 2, 3


def _srcs():
    c = os.path.join(_HERE, "..", "tinyfk_b200", "csrc")
    return [os.path.join(_HERE, "emu", "emu_driver.cpp")] + [os.path.join(c, f) for f in os.listdir(c)]


def lib():
    global _lib
    if _lib is not None:
        return _lib
    stale = not os.path.exists(_LIB) or any(os.path.getmtime(s) > os.path.getmtime(_LIB) for s in _srcs())
    if stale:
        subprocess.run(["sh", os.path.join(_HERE, "emu", "build.sh")], check=True)
    L = ctypes.CDLL(_LIB)
    L.emu_create.restype = ctypes.c_void_p
    L.emu_create.argtypes = [ctypes.c_char_p, ctypes.c_int]
    L.emu_destroy.argtypes = [ctypes.c_void_p]
    L.emu_last_error.restype = ctypes.c_char_p
    L.emu_add_link.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_int, _dp, _dp]
    L.emu_set_joint_angles.argtypes = [ctypes.c_void_p, _ip, ctypes.c_int, _dp]
    L.emu_set_base_pose.argtypes = [ctypes.c_void_p, _dp]
    for n in ("emu_num_links", "emu_num_joints", "emu_root"):
        getattr(L, n).argtypes = [ctypes.c_void_p]
    for n in ("emu_link_name", "emu_joint_name"):
        getattr(L, n).argtypes = [ctypes.c_void_p, ctypes.c_int]
        getattr(L, n).restype = ctypes.c_char_p
    L.emu_link_info.argtypes = [ctypes.c_void_p, ctypes.c_int, _ip, _ip, _ip, _dp, _dp]
    L.emu_is_relevant.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int]
    L.emu_solve.argtypes = [ctypes.c_void_p, _dp, ctypes.c_long, _ip, ctypes.c_int, _ip, ctypes.c_int, ctypes.c_int,
                            ctypes.c_int, ctypes.c_int, ctypes.c_int, _dp, _dp]
    L.emu_sqdists.argtypes = [ctypes.c_void_p, _dp, ctypes.c_long, _ip, _ip, ctypes.c_int, _ip, ctypes.c_int, ctypes.c_int,
                              ctypes.c_int, ctypes.c_int, _dp, _dp]
    _lib = L
    return L


def _ints(v):
    return np.ascontiguousarray(np.asarray(v, dtype=np.int32).reshape(-1))


class EmuModel:
    def __init__(self, xml_text, system_libm=False):
        self._L = lib()
        self._h = self._L.emu_create(xml_text.encode(), int(system_libm))
        if not self._h:
            raise ValueError(self._L.emu_last_error().decode())

    def __del__(self):
        if getattr(self, "_h", None):
            self._L.emu_destroy(self._h)
            self._h = None

    @property
    def n_links(self):
        return self._L.emu_num_links(self._h)

    @property
    def n_joints(self):
        return self._L.emu_num_joints(self._h)

    @property
    def root_link_id(self):
        return self._L.emu_root(self._h)

    def get_link_names(self):
        return [self._L.emu_link_name(self._h, i).decode() for i in range(self.n_links)]

    def get_joint_names(self):
        return [self._L.emu_joint_name(self._h, i).decode() for i in range(self.n_joints)]

    def get_link_ids(self, names):
        d = {n: i for i, n in enumerate(self.get_link_names())}
        return [d[n] for n in names]

    def get_joint_ids(self, names):
        d = {n: i for i, n in enumerate(self.get_joint_names())}
        return [d[n] for n in names]

    def link_info(self, i):
        p, t, j = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
        axis, origin = np.zeros(3), np.zeros(7)
        self._L.emu_link_info(self._h, i, ctypes.byref(p), ctypes.byref(t), ctypes.byref(j), axis.ctypes.data_as(_dp),
                              origin.ctypes.data_as(_dp))
        return p.value, t.value, j.value, axis, origin

    def is_relevant(self, l, j):
        return bool(self._L.emu_is_relevant(self._h, int(l), int(j)))

    def add_new_link(self, name, parent_id, pos, rpy):
        pos = np.ascontiguousarray(pos, dtype=np.float64)
        rpy = np.ascontiguousarray(rpy, dtype=np.float64)
        r = self._L.emu_add_link(self._h, name.encode(), int(parent_id), pos.ctypes.data_as(_dp), rpy.ctypes.data_as(_dp))
        if r < 0:
            raise ValueError(self._L.emu_last_error().decode())
        return r

    def set_joint_angles(self, ids, angles):
        ids = _ints(ids)
        a = np.ascontiguousarray(angles, dtype=np.float64)
        self._L.emu_set_joint_angles(self._h, ids.ctypes.data_as(_ip), ids.size, a.ctypes.data_as(_dp))

    def set_base_pose(self, xyzrpy):
        a = np.ascontiguousarray(xyzrpy, dtype=np.float64)
        self._L.emu_set_base_pose(self._h, a.ctypes.data_as(_dp))

    def solve_fk(self, qs, elink_ids, joint_ids, rot_type=0, base_type=0, with_jacobian=False, mode=MODE_STRICT64):
        qs = np.ascontiguousarray(np.atleast_2d(np.asarray(qs, dtype=np.float64)))
        l, j = _ints(elink_ids), _ints(joint_ids)
        n, n_in = qs.shape
        pd = {0: 3, 1: 6, 2: 7}[int(rot_type)]
        assert n_in == j.size + {0: 0, 1: 3, 2: 6}[int(base_type)]
        P = np.zeros((n * l.size, pd), order="F")
        J = np.zeros((n * l.size * pd, n_in), order="F")
        rc = self._L.emu_solve(self._h, qs.ctypes.data_as(_dp), n, l.ctypes.data_as(_ip), l.size, j.ctypes.data_as(_ip),
                               j.size, int(rot_type), int(base_type), int(bool(with_jacobian)), int(mode),
                               P.ctypes.data_as(_dp), J.ctypes.data_as(_dp))
        if rc:
            raise ValueError(self._L.emu_last_error().decode())
        return P, J

    def inter_link_sqdists(self, qs, link_id_pairs, joint_ids, base_type=0, with_jacobian=False, mode=MODE_STRICT64):
        qs = np.ascontiguousarray(np.atleast_2d(np.asarray(qs, dtype=np.float64)))
        l1, l2, j = _ints([p[0] for p in link_id_pairs]), _ints([p[1] for p in link_id_pairs]), _ints(joint_ids)
        n, n_in = qs.shape
        assert n_in == j.size + {0: 0, 1: 3, 2: 6}[int(base_type)]
        sq = np.zeros(n * l1.size)
        G = np.zeros((n * l1.size, n_in), order="F")
        rc = self._L.emu_sqdists(self._h, qs.ctypes.data_as(_dp), n, l1.ctypes.data_as(_ip), l2.ctypes.data_as(_ip), l1.size,
                                 j.ctypes.data_as(_ip), j.size, int(base_type), int(bool(with_jacobian)), int(mode),
                                 sq.ctypes.data_as(_dp), G.ctypes.data_as(_dp))
        if rc:
            raise ValueError(self._L.emu_last_error().decode())
        return sq, G

    def solve_com_fk(self, qs, joint_ids, action_link_ids=(), forces=(), base_type=0, with_jacobian=False, mode=MODE_STRICT64):
        """base_type 0 or 2 (the PLANER pad is host glue of the product's com_host, tested on the GPU)"""
        qs = np.ascontiguousarray(np.atleast_2d(np.asarray(qs, dtype=np.float64)))
        j, al = _ints(joint_ids), _ints(action_link_ids)
        f = np.ascontiguousarray(np.asarray(forces, dtype=np.float64).reshape(-1))
        n, n_in = qs.shape
        assert n_in == j.size + {0: 0, 2: 6}[int(base_type)]
        com = np.zeros((n, 3))
        J = np.zeros((3 * n, n_in), order="F")
        self._L.emu_com.argtypes = [ctypes.c_void_p, _dp, ctypes.c_long, _ip, ctypes.c_int, _ip, _dp, ctypes.c_int, ctypes.c_int,
                                    ctypes.c_int, ctypes.c_int, _dp, _dp]
        rc = self._L.emu_com(self._h, qs.ctypes.data_as(_dp), n, j.ctypes.data_as(_ip), j.size, al.ctypes.data_as(_ip),
                             f.ctypes.data_as(_dp), al.size, int(base_type), int(bool(with_jacobian)), int(mode),
                             com.ctypes.data_as(_dp), J.ctypes.data_as(_dp))
        if rc:
            raise ValueError(self._L.emu_last_error().decode())
        return com, (J if with_jacobian else None)

    def total_inertia_matrix(self, qs, joint_ids, mode=MODE_STRICT64):
        qs = np.ascontiguousarray(np.atleast_2d(np.asarray(qs, dtype=np.float64)))
        j = _ints(joint_ids)
        assert qs.shape[1] == j.size
        out = np.zeros((qs.shape[0], 3, 3))
        self._L.emu_inertia.argtypes = [ctypes.c_void_p, _dp, ctypes.c_long, _ip, ctypes.c_int, ctypes.c_int, _dp]
        rc = self._L.emu_inertia(self._h, qs.ctypes.data_as(_dp), qs.shape[0], j.ctypes.data_as(_ip), j.size, int(mode),
                                 out.ctypes.data_as(_dp))
        if rc:
            raise ValueError(self._L.emu_last_error().decode())
        return out
