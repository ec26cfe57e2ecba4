"""tinyfk_b200 -- B200-native batched forward kinematics behind tinyfk's Python interface.

Drop-in for the FK path of ``tinyfk`` (reference python/tinyfk/__init__.py:42-290): the same
``KinematicModel``, ``BaseType``, ``RotationType``, ``get_joint_ids`` / ``get_link_ids`` and
``solve_fk`` (shapes, ordering and F-contiguous layout of the returned arrays included).  All
numerics run in hand-written CUDA kernels for sm_100a through the C ABI in
``include/tinyfk_b200.h``; there is no CPU compute path -- importing this package without the built
extension raises ImportError, and solving without a CUDA device raises RuntimeError.

New relative to the reference: ``solve_fk(..., precision="fp64" | "fp32" | "fp64_fast", devices=[...])`` -- an
(N, n_dof) ndarray of configurations is solved in one launch per chunk and sharded over the listed
GPUs in contiguous slices.
"""
import copy
import os
from enum import Enum
from typing import Any, List, Optional, Sequence, Tuple, Union

import numpy as np

try:
    from . import _tinyfk  # type: ignore
except ImportError as e:  # fail loudly: nothing below works without the CUDA extension
    raise ImportError(
        "tinyfk_b200: the native extension (_tinyfk / libtinyfk_b200.so) is not built. "
        "Run `python -c 'import __graft_entry__ as g; g.build()'` or `make -C tinyfk_b200/csrc`. "
        "There is no pure-Python or CPU fallback. (%s)" % e
    ) from e

_data_dir = os.path.join(os.path.dirname(os.path.abspath(__file__)), "data")


def pr2_urdfpath() -> str:
    """reference python/tinyfk/__init__.py:21-22 (bundled instead of downloaded at import)."""
    return os.path.join(_data_dir, "pr2.urdf")


def fetch_urdfpath() -> str:
    return os.path.join(_data_dir, "fetch.urdf")


def kuka_urdfpath() -> str:
    return os.path.join(_data_dir, "kuka.urdf")


class BaseType(Enum):  # reference python/tinyfk/__init__.py:29-32
    FIXED = 0
    PLANER = 1
    FLOATING = 2


class RotationType(Enum):  # reference python/tinyfk/__init__.py:35-38
    IGNORE = _tinyfk.RotationType.IGNORE
    RPY = _tinyfk.RotationType.RPY
    XYZW = _tinyfk.RotationType.XYZW


_ROT_CODE = {RotationType.IGNORE: 0, RotationType.RPY: 1, RotationType.XYZW: 2}
_PRECISION = {"fp64": 0, "f64": 0, "double": 0, "strict": 0, "fp32": 1, "f32": 1, "float": 1, "fp64_fast": 2, "f64_fast": 2}


class KinematicModel:
    """reference python/tinyfk/__init__.py:42-290."""

    def __init__(self, urdfpath: Optional[str] = None, xml_text: Optional[str] = None, host_libm: str = "portable"):
        assert (urdfpath is None) ^ (xml_text is None)
        if not xml_text:
            with open(urdfpath, "r") as reader:
                xml_text = reader.read()
        self._xml_text = xml_text  # solely for pickling & unpickling
        self._host_libm = host_libm
        self._robot = _tinyfk.RobotModel(xml_text, {"portable": 0, "system": 1}[host_libm])
        self._add_new_link_history: List[Any] = []
        self._set_joint_angles_history: List[Any] = []

    @property
    def root_link_name(self) -> str:
        return self._robot.get_root_link_name()

    def get_q(self, joint_ids: List[int], base_type: BaseType = BaseType.FIXED) -> np.ndarray:
        if base_type == BaseType.FIXED:
            base_pose_vec = np.array([])
        elif base_type == BaseType.PLANER:
            xyzrpy = self._robot.get_base_pose()
            base_pose_vec = np.array([xyzrpy[0], xyzrpy[1], xyzrpy[-1]])
        elif base_type == BaseType.FLOATING:
            base_pose_vec = np.array(self._robot.get_base_pose())
        else:
            assert False
        return np.hstack([self._robot.get_joint_angles(list(joint_ids)), base_pose_vec])

    def set_q(self, joint_ids: List[int], q: np.ndarray, base_type: BaseType = BaseType.FIXED) -> None:
        self._set_joint_angles_history.append([joint_ids, q, base_type])
        q = np.asarray(q, dtype=np.float64)
        if base_type == BaseType.PLANER:
            assert len(q) == len(joint_ids) + 3, "{} != {}".format(len(q), len(joint_ids) + 3)
            joint_angles, b = q[:-3], q[-3:]
            self._robot.set_joint_angles(list(joint_ids), list(joint_angles))
            self._robot.set_base_pose([b[0], b[1], 0.0, 0.0, 0.0, b[2]])
        elif base_type == BaseType.FLOATING:
            assert len(q) == len(joint_ids) + 6, "{} != {}".format(len(q), len(joint_ids) + 6)
            self._robot.set_joint_angles(list(joint_ids), list(q[:-6]))
            self._robot.set_base_pose(list(q[-6:]))
        else:
            assert len(q) == len(joint_ids), "{} != {}".format(len(q), len(joint_ids))
            self._robot.set_joint_angles(list(joint_ids), list(q))

    def solve_fk(
        self,
        qs: Union[List[np.ndarray], np.ndarray],
        elink_ids: Union[List[int], int],
        joint_ids: List[int],
        rot_type: RotationType = RotationType.IGNORE,
        base_type: BaseType = BaseType.FIXED,
        with_jacobian: bool = False,
        use_cache: bool = False,
        precision: str = "fp64",
        devices: Optional[Sequence[int]] = None,
    ) -> Tuple[np.ndarray, np.ndarray]:
        """reference python/tinyfk/__init__.py:105-159.  Returns (P, J):
        P (n_seq*n_link, pose_dim), J (n_seq*n_link*pose_dim, n_dof), both F-contiguous; J is all
        zeros when with_jacobian is False.  The PLANER padding and column pick of the reference
        (:98-103, :154-158) happen inside the kernel.  ``use_cache`` is accepted and ignored: there is
        no transform cache, every batch is a pure function of its inputs and the stored state."""
        if isinstance(qs, np.ndarray):
            if qs.ndim == 1:
                qs = np.expand_dims(qs, axis=0)
        else:
            qs = np.array(qs)
            if qs.ndim == 1:
                qs = np.expand_dims(qs, axis=0)
        assert isinstance(qs, np.ndarray)
        n_seq, n_dof = qs.shape
        if isinstance(elink_ids, (int, np.integer)):
            elink_ids = [int(elink_ids)]
        elink_ids = [int(e) for e in elink_ids]
        joint_ids = [int(j) for j in joint_ids]
        n_joint = len(joint_ids)
        if base_type == BaseType.PLANER:
            assert n_dof == n_joint + 3
        elif base_type == BaseType.FLOATING:
            assert n_dof == n_joint + 6
        elif base_type == BaseType.FIXED:
            assert n_dof == n_joint
        else:
            assert False
        P, J = self._robot.solve_forward_kinematics_batch(
            qs, elink_ids, joint_ids, _ROT_CODE[rot_type], base_type.value, bool(with_jacobian),
            _PRECISION[precision], list(devices) if devices else [], True,
        )
        return P, J

    def specialize(self, elink_ids, joint_ids, rot_type=RotationType.IGNORE, base_type=BaseType.FIXED, with_jacobian=False,
                   precision="fp64", wait=True):
        """Not in the reference.  Compile the kernel for THIS query signature (one requested link) at run time: same
        results bit for bit, ~1.25x the throughput; later solve_fk calls with the signature use it.  Without this call
        the engine does it by itself in the background for signatures that keep coming back (TFK_JIT=off|auto|on)."""
        return self._robot.specialize(list(elink_ids), list(joint_ids), rot_type.value, base_type.value, bool(with_jacobian),
                                      _PRECISION[precision], bool(wait), False)

    def solve_com_fk(self, qs, joint_ids, action_link_ids=None, action_forces=None, base_type=BaseType.FIXED,
                     with_jacobian=False, precision="fp64"):
        """reference python/tinyfk/__init__.py:160-195 (batched on the GPU).  Returns (coms (n, 3), J (3n, n_dof));
        n_dof = n_joint (+6 for PLANER and FLOATING: the reference pads a PLANER row and keeps all six base columns
        on this entry); J is all zeros when with_jacobian is false.  Every row is evaluated from its own
        configuration (the reference forgets clear_cache() here for a FIXED base and repeats row 0)."""
        qs = np.atleast_2d(np.asarray(qs))
        if action_link_ids is None:
            action_link_ids, action_forces = [], []
        return self._robot.solve_com_forward_kinematics_batch(
            qs, list(joint_ids), list(action_link_ids), [float(f) for f in action_forces], base_type.value,
            bool(with_jacobian), _PRECISION[precision])

    def compute_total_inertia_matrix(self, q, joint_ids, precision="fp64"):
        """reference python/tinyfk/__init__.py:197-198: total inertia matrix about the centre of mass for one
        configuration (3, 3); a 2-D `q` of several configurations returns (n, 3, 3).  The stored base pose applies."""
        q = np.asarray(q)
        out = self._robot.compute_total_inertia_matrix_batch(np.atleast_2d(q), list(joint_ids), _PRECISION[precision])
        return out[0] if q.ndim == 1 else out

    def compute_inter_link_sqdists(self, qs, link_id_pairs, joint_ids, base_type=BaseType.FIXED, with_jacobian=False,
                                   use_cache=False, precision="fp64"):
        """reference python/tinyfk/__init__.py:233-266 (batched on the GPU; the PLANER pad and column pick are
        fused into the kernel).  Returns (sqdists [n*n_pairs], grads (n*n_pairs, n_dof)); grads is a 1x1 zero
        matrix when with_jacobian is false, as in the reference.  use_cache is accepted and ignored."""
        qs = np.atleast_2d(np.asarray(qs))
        link_ids1, link_ids2 = (list(x) for x in zip(*link_id_pairs)) if len(link_id_pairs) else ([], [])
        return self._robot.compute_inter_link_squared_dists_batch(
            qs, link_ids1, link_ids2, list(joint_ids), base_type.value, bool(with_jacobian), _PRECISION[precision])

    def get_joint_names(self) -> List[str]:
        return self._robot.get_joint_names()

    def get_link_names(self) -> List[str]:
        return self._robot.get_link_names()

    def get_joint_ids(self, joint_names) -> List[int]:
        return self._robot.get_joint_ids(list(joint_names))

    def get_link_ids(self, link_names) -> List[int]:
        return self._robot.get_link_ids(list(link_names))

    def get_joint_limits(self, joint_ids) -> List[Tuple[Optional[float], Optional[float]]]:
        limits: List[Tuple[Optional[float], Optional[float]]] = []
        for lower, upper in self._robot.get_joint_limits([int(j) for j in joint_ids]):
            if lower == -np.inf and upper == np.inf:
                limits.append((None, None))  # reference python/tinyfk/__init__.py:209-220
            else:
                limits.append((lower, upper))
        return limits

    def add_new_link(self, link_name: str, parent_id: int, position, rpy=None):
        self._add_new_link_history.append([link_name, parent_id, position, rpy])
        if rpy is None:
            rpy = np.zeros(3)
        self._robot.add_new_link(link_name, int(parent_id), [float(v) for v in position], [float(v) for v in rpy])

    def clear_cache(self) -> None:
        self._robot.clear_cache()

    @property
    def total_mass(self) -> float:
        return self._robot.total_mass_

    def __getstate__(self):  # pickling: rebuild from the XML and replay the histories
        state = self.__dict__.copy()
        state["_robot"] = None
        return state

    def __setstate__(self, state):
        self.__dict__.update(state)
        self._robot = _tinyfk.RobotModel(self._xml_text, {"portable": 0, "system": 1}[self._host_libm])
        links, angles = copy.deepcopy(self._add_new_link_history), copy.deepcopy(self._set_joint_angles_history)
        self._add_new_link_history, self._set_joint_angles_history = [], []
        for arg in links:
            self.add_new_link(*arg)
        for arg in angles:
            self.set_q(*arg)
