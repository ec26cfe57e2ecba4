"""Shared helpers of the parity tests: build the product model and the matching oracle the same way,
and compare with the tolerance tiers of DESIGN.md section "Parity".

Oracle choice: oracle/_ref (the reference's own object code, stock or libm-pinned) when the prebuilt
library is present (it travels to the GPU box), else oracle/fk_oracle.c (the C restatement, validated
against _ref in tests/test_oracle.py) driven by the product's exported link table.
"""
import numpy as np

from oracle import refapi
from tinyfk_b200 import capi
from tinyfk_b200 import workloads as W

TOL_F64 = 1e-12  # north_star: poses and Jacobians within 1e-12 abs in FP64
TOL_F32 = 1e-5   # north_star: 1e-5 in FP32
# reference's own noise floor on its finite-difference columns when libm differs in the last bit
# (SURVEY.md section 7 hard part 1; measured stock-vs-pinned reference: 5e-8)
TOL_FD_STOCK = 1e-6


class Pair:
    """product model (C ABI) + oracle with identical structure and stored state."""

    def __init__(self, urdf, strict=True, links_to_add=()):
        xml = W.urdf_text(urdf)
        self.strict = strict
        self.prod = capi.Model(xml, host_libm=0 if strict else 1)
        self._ref = refapi.RefModel(xml, strict=strict) if refapi.have_ref(strict) else None
        self._port = None
        for name, parent, pos, rpy in links_to_add:
            self.add_new_link(name, parent, pos, rpy)

    @property
    def oracle_kind(self):
        return "reference" if self._ref is not None else "port"

    def add_new_link(self, name, parent_name, pos, rpy):
        pid = self.prod.get_link_ids([parent_name])[0]
        self.prod.add_new_link(name, pid, pos, rpy)
        if self._ref is not None:
            self._ref.add_new_link(name, pid, pos, rpy)
        self._port = None

    def setup(self, cfg):
        if self._ref is not None:
            W.setup(self._ref, cfg)
        self._port = None
        return W.setup(self.prod, cfg)

    def set_state(self, joint_ids, angles, base_xyzrpy):
        self._state = (list(joint_ids), np.array(angles, dtype=np.float64), np.array(base_xyzrpy, dtype=np.float64))
        self.prod.set_joint_angles(joint_ids, angles)
        self.prod.set_base_pose(base_xyzrpy)
        self._apply_state_to_oracle()

    def _oracle(self):
        if self._ref is not None:
            return self._ref
        if self._port is None:
            self._port = refapi.PortModel(self.prod.export_table(), strict=self.strict)
        return self._port

    def _apply_state_to_oracle(self):
        st = getattr(self, "_state", None)
        if st is None:
            return
        o = self._oracle()
        o.set_joint_angles(st[0], st[1])
        o.set_base_pose(st[2])

    def oracle_solve(self, q, lids, jids, rot, base, jac):
        # both sides mutate their stored state during a batch (last configuration wins): re-apply
        self._apply_state_to_oracle()
        return self._oracle().solve_fk(q, lids, jids, rot, base, jac)

    def prod_solve(self, q, lids, jids, rot, base, jac, **kw):
        st = getattr(self, "_state", None)
        if st is not None:
            self.prod.set_joint_angles(st[0], st[1])
            self.prod.set_base_pose(st[2])
        return self.prod.solve_fk(q, lids, jids, rot, base, jac, **kw)


def fd_column_mask(n_joint, base, n_cols):
    """Boolean mask over Jacobian columns: the reference's finite-difference columns."""
    m = np.zeros(n_cols, dtype=bool)
    if base == W.BASE_PLANER:
        m[n_joint + 2] = True
    elif base == W.BASE_FLOATING:
        m[n_joint + 3:n_joint + 6] = True
    return m


def gimbal_rows(P_ref, rot):
    """rows whose pitch sits on the getRPY gimbal branch (pose.h:178-185): |pitch| == pi/2 exactly"""
    if rot != W.ROT_RPY:
        return np.zeros(P_ref.shape[0], dtype=bool)
    return np.abs(np.abs(P_ref[:, 4]) - np.pi / 2) < 1e-12


def rel_err(a, b):
    return np.abs(a - b) / np.maximum(1.0, np.abs(b))


def assert_bitwise(P, J, Pr, Jr, what=""):
    """identical numbers (-0.0 == +0.0 allowed: see DESIGN.md 'signed zeros'); NaN must match NaN"""
    for name, a, b in (("P", P, Pr), ("J", J, Jr)):
        same = (a == b) | (np.isnan(a) & np.isnan(b))
        if not same.all():
            idx = np.argwhere(~same)[0]
            raise AssertionError("%s %s differs at %s: %r vs %r (%d entries)" % (what, name, tuple(idx), a[tuple(idx)],
                                                                             b[tuple(idx)], (~same).sum()))


def assert_close_stock(P, J, Pr, Jr, n_joint, base, rot, n_link, what=""):
    """FP64 vs the stock-glibc reference: 1e-12 on poses and analytic entries (relative above 1 for the
    1/cos(pitch)-amplified RPY rows), the reference's own FD noise floor on its FD columns."""
    finite = np.isfinite(Pr).all(axis=1)
    ep = rel_err(P[finite], Pr[finite]).max() if finite.any() else 0.0
    assert ep <= TOL_F64, "%s P err %.3e" % (what, ep)
    if J is None or Jr is None or J.size == 0:
        return
    fd = fd_column_mask(n_joint, base, J.shape[1])
    pd = W.pose_dim(rot)
    bad_rows = np.repeat(gimbal_rows(Pr, rot) | ~finite, pd)
    Ja, Jb = J[~bad_rows], Jr[~bad_rows]
    ok = np.isfinite(Jb)
    if (~fd).any():
        e = np.where(ok[:, ~fd], rel_err(Ja[:, ~fd], Jb[:, ~fd]), 0.0).max()
        assert e <= TOL_F64, "%s analytic J err %.3e" % (what, e)
    if fd.any():
        e = np.where(ok[:, fd], rel_err(Ja[:, fd], Jb[:, fd]), 0.0).max()
        assert e <= TOL_FD_STOCK, "%s FD J err %.3e" % (what, e)
