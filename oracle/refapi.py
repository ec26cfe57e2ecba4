"""ctypes bindings of the CPU checkers (TEST INFRASTRUCTURE -- never on the product path).

  RefModel(xml, strict=False)   -> oracle/_ref/libtinyfk_ref[_strict].so : the reference's own C++
                                   (src/tinyfk.cpp, src/kinematics.cpp, urdf_parser/) behind the loop
                                   of src/wrapper.cpp:60-124 (see oracle/ref_driver.cpp).
  PortModel(table, strict=False) -> oracle/build/libfk_oracle[_strict].so : plain-C restatement
                                   (oracle/fk_oracle.c) driven by a link table.

Both expose solve_fk(qs, elink_ids, joint_ids, rot_type, base_type, with_jacobian) returning
(P, J) as F-ordered float64 numpy arrays with exactly the shapes tinyfk's Python facade returns
(python/tinyfk/__init__.py:105-159).
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_dp = ctypes.POINTER(ctypes.c_double)
_ip = ctypes.POINTER(ctypes.c_int)

ROT_IGNORE, ROT_RPY, ROT_XYZW = 0, 1, 2
BASE_FIXED, BASE_PLANER, BASE_FLOATING = 0, 1, 2


def pose_dim(rot_type):
    return {0: 3, 1: 6, 2: 7}[int(rot_type)]


def base_dof(base_type):
    return {0: 0, 1: 3, 2: 6}[int(base_type)]


def ref_lib_path(strict=False):
    return os.path.join(_HERE, "_ref", "libtinyfk_ref_strict.so" if strict else "libtinyfk_ref.so")


def port_lib_path(strict=False):
    return os.path.join(_HERE, "build", "libfk_oracle_strict.so" if strict else "libfk_oracle.so")


def build_port():
    subprocess.run(["make", "-s", "-C", _HERE, "port"], check=True)


def build_ref(reference="/root/reference"):
    if not os.path.isdir(reference):
        return False
    subprocess.run(["make", "-s", "-C", _HERE, "ref", "REF=" + reference], check=True)
    return True


def have_ref(strict=False):
    return os.path.exists(ref_lib_path(strict))


def _as_int_array(v):
    return np.ascontiguousarray(np.asarray(v, dtype=np.int32).reshape(-1))


_ref_libs = {}


def _load_ref(strict):
    if strict in _ref_libs:
        return _ref_libs[strict]
    lib = ctypes.CDLL(ref_lib_path(strict))
    lib.tfkref_create.restype = ctypes.c_void_p
    lib.tfkref_create.argtypes = [ctypes.c_char_p]
    lib.tfkref_destroy.argtypes = [ctypes.c_void_p]
    lib.tfkref_last_error.restype = ctypes.c_char_p
    for name in ("tfkref_num_links", "tfkref_num_joints", "tfkref_root_link_id"):
        getattr(lib, name).argtypes = [ctypes.c_void_p]
        getattr(lib, name).restype = ctypes.c_int
    lib.tfkref_total_mass.argtypes = [ctypes.c_void_p]
    lib.tfkref_total_mass.restype = ctypes.c_double
    for name in ("tfkref_link_name", "tfkref_joint_name"):
        getattr(lib, name).argtypes = [ctypes.c_void_p, ctypes.c_int]
        getattr(lib, name).restype = ctypes.c_char_p
    for name in ("tfkref_link_id", "tfkref_joint_id"):
        getattr(lib, name).argtypes = [ctypes.c_void_p, ctypes.c_char_p]
        getattr(lib, name).restype = ctypes.c_int
    lib.tfkref_link_info.argtypes = [ctypes.c_void_p, ctypes.c_int, _ip, _ip, _ip, _dp, _dp]
    lib.tfkref_joint_limits.argtypes = [ctypes.c_void_p, ctypes.c_int, _dp, _dp]
    lib.tfkref_is_relevant.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int]
    lib.tfkref_is_relevant.restype = ctypes.c_int
    lib.tfkref_add_link.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_int, _dp, _dp]
    lib.tfkref_add_link.restype = ctypes.c_int
    lib.tfkref_set_joint_angles.argtypes = [ctypes.c_void_p, _ip, ctypes.c_int, _dp]
    lib.tfkref_get_joint_angles.argtypes = [ctypes.c_void_p, _ip, ctypes.c_int, _dp]
    lib.tfkref_set_base_pose.argtypes = [ctypes.c_void_p, _dp]
    lib.tfkref_get_base_pose.argtypes = [ctypes.c_void_p, _dp]
    lib.tfkref_solve_fk.argtypes = [ctypes.c_void_p, _dp, ctypes.c_long, _ip, ctypes.c_int, _ip, ctypes.c_int,
                                    ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int, _dp, _dp]
    lib.tfkref_solve_fk.restype = ctypes.c_int
    lib.tfkref_sqdists.argtypes = [ctypes.c_void_p, _dp, ctypes.c_long, _ip, _ip, ctypes.c_int, _ip, ctypes.c_int, ctypes.c_int,
                                   ctypes.c_int, ctypes.c_int, _dp, _dp]
    lib.tfkref_sqdists.restype = ctypes.c_int
    lib.tfkref_com.argtypes = [ctypes.c_void_p, _dp, ctypes.c_long, _ip, ctypes.c_int, _ip, _dp, ctypes.c_int, ctypes.c_int,
                               ctypes.c_int, ctypes.c_int, _dp, _dp]
    lib.tfkref_com.restype = ctypes.c_int
    lib.tfkref_inertia.argtypes = [ctypes.c_void_p, _dp, ctypes.c_long, _ip, ctypes.c_int, ctypes.c_int, _dp]
    lib.tfkref_inertia.restype = ctypes.c_int
    lib.tfkref_probe_sin.argtypes = [ctypes.c_double]
    lib.tfkref_probe_sin.restype = ctypes.c_double
    lib.tfkref_probe_atan2.argtypes = [ctypes.c_double, ctypes.c_double]
    lib.tfkref_probe_atan2.restype = ctypes.c_double
    _ref_libs[strict] = lib
    return lib


class RefModel:
    """The reference KinematicModel (its own object code) behind a C handle."""

    def __init__(self, xml_text, strict=False):
        self._lib = _load_ref(strict)
        self.strict = strict
        self._h = self._lib.tfkref_create(xml_text.encode())
        if not self._h:
            raise ValueError("reference model: " + self._lib.tfkref_last_error().decode())

    def __del__(self):
        if getattr(self, "_h", None):
            self._lib.tfkref_destroy(self._h)
            self._h = None

    @property
    def n_links(self):
        return self._lib.tfkref_num_links(self._h)

    @property
    def n_joints(self):
        return self._lib.tfkref_num_joints(self._h)

    @property
    def root_link_id(self):
        return self._lib.tfkref_root_link_id(self._h)

    @property
    def total_mass(self):
        return self._lib.tfkref_total_mass(self._h)

    def get_link_names(self):
        return [self._lib.tfkref_link_name(self._h, i).decode() for i in range(self.n_links)]

    def get_joint_names(self):
        return [self._lib.tfkref_joint_name(self._h, i).decode() for i in range(self.n_joints)]

    def get_link_ids(self, names):
        out = []
        for n in names:
            i = self._lib.tfkref_link_id(self._h, n.encode())
            if i < 0:
                raise ValueError(self._lib.tfkref_last_error().decode())
            out.append(i)
        return out

    def get_joint_ids(self, names):
        out = []
        for n in names:
            i = self._lib.tfkref_joint_id(self._h, n.encode())
            if i < 0:
                raise ValueError(self._lib.tfkref_last_error().decode())
            out.append(i)
        return out

    def link_info(self, link_id):
        parent, jtype, jid = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
        axis = np.zeros(3)
        origin = np.zeros(7)
        self._lib.tfkref_link_info(self._h, link_id, ctypes.byref(parent), ctypes.byref(jtype), ctypes.byref(jid),
                                   axis.ctypes.data_as(_dp), origin.ctypes.data_as(_dp))
        return parent.value, jtype.value, jid.value, axis, origin

    def get_joint_limits(self, joint_ids):
        out = []
        for j in joint_ids:
            lo, hi = ctypes.c_double(), ctypes.c_double()
            self._lib.tfkref_joint_limits(self._h, int(j), ctypes.byref(lo), ctypes.byref(hi))
            out.append((lo.value, hi.value))
        return out

    def is_relevant(self, link_id, joint_id):
        return bool(self._lib.tfkref_is_relevant(self._h, int(link_id), int(joint_id)))

    def add_new_link(self, name, parent_id, pos, rpy):
        pos = np.ascontiguousarray(pos, dtype=np.float64)
        rpy = np.ascontiguousarray(rpy, dtype=np.float64)
        if self._lib.tfkref_add_link(self._h, name.encode(), int(parent_id), pos.ctypes.data_as(_dp),
                                     rpy.ctypes.data_as(_dp)):
            raise ValueError(self._lib.tfkref_last_error().decode())

    def set_joint_angles(self, joint_ids, angles):
        ids = _as_int_array(joint_ids)
        a = np.ascontiguousarray(angles, dtype=np.float64)
        self._lib.tfkref_set_joint_angles(self._h, ids.ctypes.data_as(_ip), ids.size, a.ctypes.data_as(_dp))

    def get_joint_angles(self, joint_ids):
        ids = _as_int_array(joint_ids)
        a = np.zeros(ids.size)
        self._lib.tfkref_get_joint_angles(self._h, ids.ctypes.data_as(_ip), ids.size, a.ctypes.data_as(_dp))
        return a

    def set_base_pose(self, xyzrpy):
        a = np.ascontiguousarray(xyzrpy, dtype=np.float64)
        assert a.size == 6
        self._lib.tfkref_set_base_pose(self._h, a.ctypes.data_as(_dp))

    def get_base_pose(self):
        a = np.zeros(6)
        self._lib.tfkref_get_base_pose(self._h, a.ctypes.data_as(_dp))
        return a

    def solve_fk(self, qs, elink_ids, joint_ids, rot_type=0, base_type=0, with_jacobian=False, use_cache=False):
        qs = np.ascontiguousarray(np.atleast_2d(np.asarray(qs, dtype=np.float64)))
        lids = _as_int_array(elink_ids)
        jids = _as_int_array(joint_ids)
        n_wp, n_in = qs.shape
        assert n_in == jids.size + base_dof(base_type), (n_in, jids.size, base_type)
        pd = pose_dim(rot_type)
        P = np.zeros((n_wp * lids.size, pd), order="F")
        J = np.zeros((n_wp * lids.size * pd, n_in), order="F")
        rc = self._lib.tfkref_solve_fk(self._h, qs.ctypes.data_as(_dp), n_wp, lids.ctypes.data_as(_ip), lids.size,
                                       jids.ctypes.data_as(_ip), jids.size, int(rot_type), int(base_type),
                                       int(bool(with_jacobian)), int(bool(use_cache)), P.ctypes.data_as(_dp),
                                       J.ctypes.data_as(_dp) if with_jacobian else None)
        if rc:
            raise ValueError(self._lib.tfkref_last_error().decode())
        return P, J

    def inter_link_sqdists(self, qs, link_id_pairs, joint_ids, base_type=0, with_jacobian=False, use_cache=False):
        """compute_inter_link_sqdists (python/tinyfk/__init__.py:233-266): (sqdists [n*n_check], grads F-order)"""
        qs = np.ascontiguousarray(np.atleast_2d(np.asarray(qs, dtype=np.float64)))
        l1 = _as_int_array([p[0] for p in link_id_pairs])
        l2 = _as_int_array([p[1] for p in link_id_pairs])
        jids = _as_int_array(joint_ids)
        n_wp, n_in = qs.shape
        assert n_in == jids.size + base_dof(base_type), (n_in, jids.size, base_type)
        sq = np.zeros(n_wp * l1.size)
        G = np.zeros((n_wp * l1.size, n_in), order="F") if with_jacobian else np.zeros((1, 1))
        rc = self._lib.tfkref_sqdists(self._h, qs.ctypes.data_as(_dp), n_wp, l1.ctypes.data_as(_ip), l2.ctypes.data_as(_ip),
                                      l1.size, jids.ctypes.data_as(_ip), jids.size, int(base_type),
                                      int(bool(with_jacobian)), int(bool(use_cache)), sq.ctypes.data_as(_dp),
                                      G.ctypes.data_as(_dp) if with_jacobian else None)
        if rc:
            raise ValueError(self._lib.tfkref_last_error().decode())
        return sq, G

    def solve_com_fk(self, qs, joint_ids, action_link_ids=(), forces=(), base_type=0, with_jacobian=False, clear_each=True):
        """solve_com_fk (python/tinyfk/__init__.py:160-195): (com (n, 3), J (3n, n_joint + 0|6) F-order or None).
        clear_each=False reproduces the reference's stale-cache behaviour with a FIXED base (ref_driver.cpp)."""
        qs = np.ascontiguousarray(np.atleast_2d(np.asarray(qs, dtype=np.float64)))
        jids, al = _as_int_array(joint_ids), _as_int_array(action_link_ids)
        f = np.ascontiguousarray(np.asarray(forces, dtype=np.float64).reshape(-1))
        assert f.size == al.size
        n_wp, n_in = qs.shape
        assert n_in == jids.size + base_dof(base_type), (n_in, jids.size, base_type)
        n_dof = jids.size + (6 if base_type else 0)
        com = np.zeros((n_wp, 3))
        J = np.zeros((3 * n_wp, n_dof), order="F") if with_jacobian else None
        rc = self._lib.tfkref_com(self._h, qs.ctypes.data_as(_dp), n_wp, jids.ctypes.data_as(_ip), jids.size,
                                  al.ctypes.data_as(_ip), f.ctypes.data_as(_dp), al.size, int(base_type),
                                  int(bool(with_jacobian)), int(bool(clear_each)), com.ctypes.data_as(_dp),
                                  J.ctypes.data_as(_dp) if with_jacobian else None)
        if rc:
            raise ValueError(self._lib.tfkref_last_error().decode())
        return com, J

    def total_inertia_matrix(self, qs, joint_ids, clear_each=True):
        """compute_total_inertia_matrix (src/wrapper.cpp:181-187) per row of qs: (n, 3, 3)"""
        qs = np.ascontiguousarray(np.atleast_2d(np.asarray(qs, dtype=np.float64)))
        jids = _as_int_array(joint_ids)
        assert qs.shape[1] == jids.size
        out = np.zeros((qs.shape[0], 3, 3))
        rc = self._lib.tfkref_inertia(self._h, qs.ctypes.data_as(_dp), qs.shape[0], jids.ctypes.data_as(_ip), jids.size,
                                      int(bool(clear_each)), out.ctypes.data_as(_dp))
        if rc:
            raise ValueError(self._lib.tfkref_last_error().decode())
        return out

    def probe_sin(self, x):
        return self._lib.tfkref_probe_sin(float(x))

    def export_table(self):
        """Link table in the layout oracle/fk_oracle.h expects (see PortModel)."""
        n = self.n_links
        parent = np.zeros(n, dtype=np.int32)
        jtype = np.zeros(n, dtype=np.int32)
        jid = np.zeros(n, dtype=np.int32)
        axis = np.zeros((n, 3))
        origin = np.zeros((n, 7))
        for i in range(n):
            p, t, j, a, o = self.link_info(i)
            parent[i], jtype[i], jid[i] = p, t, j
            axis[i], origin[i] = a, o
        return dict(parent=parent, jtype=jtype, jid=jid, axis=axis, origin=origin, n_joints=self.n_joints,
                    root=self.root_link_id)


_port_libs = {}


def _load_port(strict):
    if strict in _port_libs:
        return _port_libs[strict]
    if not os.path.exists(port_lib_path(strict)):
        build_port()
    lib = ctypes.CDLL(port_lib_path(strict))
    lib.fko_create.restype = ctypes.c_void_p
    lib.fko_create.argtypes = [ctypes.c_int, ctypes.c_int, _ip, _ip, _ip, _dp, _dp, ctypes.c_int]
    lib.fko_destroy.argtypes = [ctypes.c_void_p]
    lib.fko_set_joint_angles.argtypes = [ctypes.c_void_p, _ip, ctypes.c_int, _dp]
    lib.fko_set_base_pose.argtypes = [ctypes.c_void_p, _dp]
    lib.fko_solve_fk.argtypes = [ctypes.c_void_p, _dp, ctypes.c_long, _ip, ctypes.c_int, _ip, ctypes.c_int,
                                 ctypes.c_int, ctypes.c_int, ctypes.c_int, _dp, _dp]
    _port_libs[strict] = lib
    return lib


class PortModel:
    """oracle/fk_oracle.c (plain-C restatement) driven by a link table:
    dict(parent, jtype, jid, axis[n,3], origin[n,7], n_joints, root) -- see RefModel.export_table()."""

    def __init__(self, table, strict=False):
        self._lib = _load_port(strict)
        self.strict = strict
        parent = _as_int_array(table["parent"])
        jtype = _as_int_array(table["jtype"])
        jid = _as_int_array(table["jid"])
        axis = np.ascontiguousarray(table["axis"], dtype=np.float64)
        origin = np.ascontiguousarray(table["origin"], dtype=np.float64)
        self._h = self._lib.fko_create(parent.size, int(table["n_joints"]), parent.ctypes.data_as(_ip),
                                       jtype.ctypes.data_as(_ip), jid.ctypes.data_as(_ip), axis.ctypes.data_as(_dp),
                                       origin.ctypes.data_as(_dp), int(table["root"]))

    def __del__(self):
        if getattr(self, "_h", None):
            self._lib.fko_destroy(self._h)
            self._h = None

    def set_joint_angles(self, joint_ids, angles):
        ids = _as_int_array(joint_ids)
        a = np.ascontiguousarray(angles, dtype=np.float64)
        self._lib.fko_set_joint_angles(self._h, ids.ctypes.data_as(_ip), ids.size, a.ctypes.data_as(_dp))

    def set_base_pose(self, xyzrpy):
        a = np.ascontiguousarray(xyzrpy, dtype=np.float64)
        self._lib.fko_set_base_pose(self._h, a.ctypes.data_as(_dp))

    def solve_fk(self, qs, elink_ids, joint_ids, rot_type=0, base_type=0, with_jacobian=False, use_cache=False):
        qs = np.ascontiguousarray(np.atleast_2d(np.asarray(qs, dtype=np.float64)))
        lids = _as_int_array(elink_ids)
        jids = _as_int_array(joint_ids)
        n_wp, n_in = qs.shape
        assert n_in == jids.size + base_dof(base_type)
        pd = pose_dim(rot_type)
        P = np.zeros((n_wp * lids.size, pd), order="F")
        J = np.zeros((n_wp * lids.size * pd, n_in), order="F")
        self._lib.fko_solve_fk(self._h, qs.ctypes.data_as(_dp), n_wp, lids.ctypes.data_as(_ip), lids.size,
                               jids.ctypes.data_as(_ip), jids.size, int(rot_type), int(base_type),
                               int(bool(with_jacobian)), P.ctypes.data_as(_dp),
                               J.ctypes.data_as(_dp) if with_jacobian else None)
        return P, J
