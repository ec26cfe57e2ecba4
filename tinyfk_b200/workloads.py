"""The five BASELINE.json configurations as data (robot, joints, links, rot/base type) plus the
synthetic input generator of SURVEY.md section 8(d).  Pure numpy / names only -- works with any model
object that has get_joint_ids / get_link_ids / get_joint_limits / add_new_link (the product's, the
ctypes Model, or the oracle's RefModel), so tests, smoke() and bench.py all draw identical inputs.
"""
import os

import numpy as np

_DATA = os.path.join(os.path.dirname(os.path.abspath(__file__)), "data")

ROT_IGNORE, ROT_RPY, ROT_XYZW = 0, 1, 2
BASE_FIXED, BASE_PLANER, BASE_FLOATING = 0, 1, 2

PR2_RARM = ["r_shoulder_pan_joint", "r_shoulder_lift_joint", "r_upper_arm_roll_joint", "r_elbow_flex_joint",
            "r_forearm_roll_joint", "r_wrist_flex_joint", "r_wrist_roll_joint"]
PR2_LARM = ["l" + n[1:] for n in PR2_RARM]
KUKA_JOINTS = ["lbr_iiwa_joint_%d" % i for i in range(1, 8)]
FETCH_JOINTS = ["torso_lift_joint", "shoulder_pan_joint", "shoulder_lift_joint", "upperarm_roll_joint",
                "elbow_flex_joint", "forearm_roll_joint", "wrist_flex_joint", "wrist_roll_joint"]
FETCH_SPHERE_PARENTS = ["torso_lift_link", "shoulder_pan_link", "shoulder_lift_link", "upperarm_roll_link",
                        "elbow_flex_link", "forearm_roll_link", "wrist_flex_link", "wrist_roll_link", "gripper_link",
                        "head_pan_link"]
_SIDE = ["shoulder_pan_link", "shoulder_lift_link", "upper_arm_roll_link", "upper_arm_link", "elbow_flex_link",
         "forearm_roll_link", "forearm_link", "wrist_flex_link", "wrist_roll_link", "gripper_palm_link",
         "gripper_l_finger_link", "gripper_l_finger_tip_link", "gripper_r_finger_link", "gripper_r_finger_tip_link",
         "gripper_tool_frame", "gripper_led_frame", "gripper_motor_accelerometer_link", "gripper_l_finger_tip_frame",
         "forearm_cam_frame", "gripper_motor_slider_link", "gripper_motor_screw_link"]
PR2_BODY_LINKS = (["torso_lift_link", "head_pan_link", "head_tilt_link", "head_plate_frame", "sensor_mount_link",
                   "double_stereo_link", "laser_tilt_mount_link", "imu_link"]
                  + ["r_" + n for n in _SIDE] + ["l_" + n for n in _SIDE])
assert len(PR2_BODY_LINKS) == 50

README_Q = [0.564, 0.35, -0.74, -0.7, -0.7, -0.17, -0.63, 0.1, 0.2, 0.3]  # reference README.md:41

CONFIGS = {
    # name: (urdf, joint names, link names or None (=cfg4 spheres), rot, base, nominal N)
    "cfg1": ("pr2.urdf", PR2_RARM, ["r_gripper_tool_frame"], ROT_RPY, BASE_PLANER, 1),
    "cfg2": ("pr2.urdf", PR2_RARM, ["r_gripper_tool_frame"], ROT_RPY, BASE_PLANER, 1 << 20),
    "cfg3": ("kuka.urdf", KUKA_JOINTS, ["lbr_iiwa_link_7"], ROT_XYZW, BASE_FIXED, 1 << 24),
    "cfg4": ("fetch.urdf", FETCH_JOINTS, None, ROT_IGNORE, BASE_FIXED, 1 << 22),
    "cfg5": ("pr2.urdf", ["torso_lift_joint"] + PR2_RARM + PR2_LARM, PR2_BODY_LINKS, ROT_RPY, BASE_FLOATING, 1 << 23),
}


def urdf_text(name):
    with open(os.path.join(_DATA, name), "r") as f:
        return f.read()


def pose_dim(rot):
    return {0: 3, 1: 6, 2: 7}[int(rot)]


def base_dof(base):
    return {0: 0, 1: 3, 2: 6}[int(base)]


def fetch_sphere_specs():
    """30 collision-sphere centres: 3 per parent link, fixed offsets (cfg4)."""
    specs = []
    for p in FETCH_SPHERE_PARENTS:
        for k in range(3):
            specs.append(("sphere_%s_%d" % (p, k), p, [0.05 * k, 0.02 * (-1) ** k, 0.03 * (k - 1)]))
    return specs


def setup(model, cfg):
    """Resolve a config on `model`: adds cfg4's sphere links, returns (joint_ids, link_ids, rot, base)."""
    urdf, jn, ln, rot, base, _ = CONFIGS[cfg]
    if ln is None:
        ln = []
        for name, parent, pos in fetch_sphere_specs():
            pid = model.get_link_ids([parent])[0]
            model.add_new_link(name, pid, pos, [0.0, 0.0, 0.0])
            ln.append(name)
    return list(model.get_joint_ids(jn)), list(model.get_link_ids(ln)), rot, base


def algorithmic_bytes(cfg_or_shape, real_bytes=8):
    """Bytes per configuration: input row + P + J."""
    if isinstance(cfg_or_shape, str):
        _, jn, ln, rot, base, _ = CONFIGS[cfg_or_shape]
        n_joint, n_link = len(jn), (30 if ln is None else len(ln))
    else:
        n_joint, n_link, rot, base = cfg_or_shape
    n_dof = n_joint + base_dof(base)
    pd = pose_dim(rot)
    return real_bytes * (n_dof + n_link * pd + n_link * pd * n_dof)


def sample_q(limits, base, n, seed=0, wide_pitch=False):
    """uniform in the URDF limits ([-pi, pi] for continuous joints); base x,y,z in [-1,1],
    roll/pitch in [-0.5,0.5] (wide_pitch: pitch in [-pi/2, pi/2] to exercise the gimbal branch),
    yaw in [-pi, pi].  limits: list of (lo, hi) with None or +-inf for continuous joints."""
    rng = np.random.default_rng(seed)
    cols = []
    for lo, hi in limits:
        if lo is None or hi is None or not np.isfinite(lo) or not np.isfinite(hi):
            lo, hi = -np.pi, np.pi
        cols.append(rng.uniform(lo, hi, n))
    if base == BASE_PLANER:
        cols += [rng.uniform(-1, 1, n), rng.uniform(-1, 1, n), rng.uniform(-np.pi, np.pi, n)]
    elif base == BASE_FLOATING:
        pl = np.pi / 2 if wide_pitch else 0.5
        cols += [rng.uniform(-1, 1, n) for _ in range(3)]
        cols += [rng.uniform(-0.5, 0.5, n), rng.uniform(-pl, pl, n), rng.uniform(-np.pi, np.pi, n)]
    if not cols:
        return np.zeros((n, 0))
    return np.ascontiguousarray(np.stack(cols, axis=1))
