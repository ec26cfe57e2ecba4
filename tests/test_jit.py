"""Plan specialisation (tinyfk_b200/csrc/jit.cpp): the fused kernel compiled at run time against one plan must be the
generic kernel bit for bit.  The compile itself needs no GPU (NVRTC), so it is checked on the CPU too."""
import os

import numpy as np
import pytest

from tinyfk_b200 import capi
from tinyfk_b200 import workloads as W

CASES = [  # urdf, joints, link, rot, base, jac, precision
    ("pr2.urdf", W.PR2_RARM, "r_gripper_tool_frame", 1, 1, True, 0),   # cfg2
    ("kuka.urdf", W.KUKA_JOINTS, "lbr_iiwa_link_7", 2, 0, True, 0),    # cfg3
    ("pr2.urdf", W.PR2_RARM, "r_gripper_tool_frame", 1, 2, True, 0),   # FLOATING: three finite-difference columns
    ("fetch.urdf", W.FETCH_JOINTS, "gripper_link", 0, 1, True, 0),     # prismatic torso, IGNORE
    ("pr2.urdf", W.PR2_RARM, "r_gripper_tool_frame", 2, 0, False, 0),  # pose only
    ("pr2.urdf", W.PR2_RARM, "r_gripper_tool_frame", 1, 1, True, 1),   # FP32 mode
]


@pytest.fixture(autouse=True)
def _cache(tmp_path, monkeypatch):
    monkeypatch.setenv("TFK_CACHE_DIR", str(tmp_path / "cache"))


def _have_nvrtc():
    m = capi.Model(W.urdf_text("kuka.urdf"), host_libm=0)
    j, l = list(m.get_joint_ids(W.KUKA_JOINTS)), list(m.get_link_ids(["lbr_iiwa_link_7"]))
    try:
        m.specialize(l, j, 0, 0, False, 0, wait=True, compile_only=True)
        return True
    except RuntimeError:
        return False


needs_nvrtc = pytest.mark.skipif(not _have_nvrtc(), reason="NVRTC (libnvrtc.so.12) not found")


@needs_nvrtc
def test_specialisation_compiles_without_a_gpu_and_is_cached(tmp_path):
    m = capi.Model(W.urdf_text("pr2.urdf"), host_libm=0)
    j, l = list(m.get_joint_ids(W.PR2_RARM)), list(m.get_link_ids(["r_gripper_tool_frame"]))
    a = m.specialize(l, j, 1, 1, True, 0, wait=True, compile_only=True)
    assert a["state"] == 2 and a["block"] == 256 and a["blocks_per_sm"] == 2 and a["cubin_bytes"] > 10000 and a["from_cache"] == 0
    assert any(f.endswith(".cubin") for f in os.listdir(os.environ["TFK_CACHE_DIR"]))
    m2 = capi.Model(W.urdf_text("pr2.urdf"), host_libm=0)
    b = m2.specialize(l, j, 1, 1, True, 0, wait=True, compile_only=True)
    assert b["from_cache"] == 1 and b["cubin_bytes"] == a["cubin_bytes"]
    # several links: the two-phase kernel is specialised (128-thread blocks)
    l2 = list(m.get_link_ids(["r_gripper_tool_frame", "r_elbow_flex_link"]))
    c = m.specialize(l2, j, 0, 1, True, 0, wait=True, compile_only=True)
    assert c["state"] == 2 and c["block"] == 128 and c["cubin_bytes"] > 10000
    c = m.specialize(l2, j, 1, 1, True, 1, wait=True, compile_only=True)  # orientation rows: FP32 mode only
    assert c["state"] == 2 and c["block"] == 128
    with pytest.raises(ValueError):
        m.specialize(l2, j, 1, 1, True, 0, wait=True, compile_only=True)
    # more joints than the unrolled build takes
    with pytest.raises(ValueError):
        m.specialize(l2, list(range(m.n_joints)), 1, 1, True, 0, wait=True, compile_only=True)


@needs_nvrtc
def test_corrupt_or_foreign_cache_files_are_discarded_and_recompiled(tmp_path):
    """the disk cache verifies key, length and payload hash: a truncated or foreign file is deleted and the
    kernel recompiled instead of being handed to the driver; a cache directory others can write to is not used"""
    d = os.environ["TFK_CACHE_DIR"]
    m = capi.Model(W.urdf_text("kuka.urdf"), host_libm=0)
    j, l = list(m.get_joint_ids(W.KUKA_JOINTS)), list(m.get_link_ids(["lbr_iiwa_link_7"]))
    a = m.specialize(l, j, 2, 0, True, 0, wait=True, compile_only=True)
    files = [f for f in os.listdir(d) if f.endswith(".cubin")]
    assert a["from_cache"] == 0 and len(files) == 1
    path = os.path.join(d, files[0])
    good = open(path, "rb").read()
    assert good[:8] == b"TFKCUBN1" and len(good) == a["cubin_bytes"] + 40
    assert (os.stat(d).st_mode & 0o077) == 0
    for bad in (good[:len(good) // 2], good[:40] + bytes(len(good) - 40), b"\x7fELF" + good[4:], good[:-1] + bytes([good[-1] ^ 1])):
        open(path, "wb").write(bad)
        b = capi.Model(W.urdf_text("kuka.urdf"), host_libm=0).specialize(l, j, 2, 0, True, 0, wait=True, compile_only=True)
        assert b["from_cache"] == 0 and b["cubin_bytes"] == a["cubin_bytes"]
        assert open(path, "rb").read() == good  # rewritten
    b = capi.Model(W.urdf_text("kuka.urdf"), host_libm=0).specialize(l, j, 2, 0, True, 0, wait=True, compile_only=True)
    assert b["from_cache"] == 1
    os.chmod(d, 0o777)  # writable by others: not trusted, neither read nor written
    try:
        b = capi.Model(W.urdf_text("kuka.urdf"), host_libm=0).specialize(l, j, 2, 0, True, 0, wait=True, compile_only=True)
        assert b["from_cache"] == 0
    finally:
        os.chmod(d, 0o700)


MULTI = [  # workload (BASELINE config), base override, jac, precision
    ("cfg4", None, True, 0), ("cfg4", 2, True, 0), ("cfg4", 1, False, 0), ("cfg5", None, True, 1), ("cfg5", 1, True, 1),
    ("cfg4", None, True, 1),
]


@pytest.mark.gpu
@needs_nvrtc
@pytest.mark.parametrize("cfg,base_o,jac,prec", MULTI)
def test_gpu_specialised_two_phase_kernel_bit_for_bit(cfg, base_o, jac, prec):
    """several links (Fetch spheres, PR2 50 links): the plan-specialised two-phase kernel against the generic one and
    against the oracle; 333 configurations = ragged tail tile"""
    import parity
    pair = parity.Pair(W.CONFIGS[cfg][0], strict=True)
    jids, lids, rot, base = pair.setup(cfg)
    base = base if base_o is None else base_o
    m = pair.prod
    q = W.sample_q(m.get_joint_limits(jids), base, 333, seed=9, wide_pitch=True)
    m.set_jit_mode(0)
    P0, J0 = m.solve_fk(q, lids, jids, rot, base, jac, precision=prec)
    assert m.last_launch_info()["kernel"].startswith("fk_twophase_kernel")
    m.set_jit_mode(1)
    info = m.specialize(lids, jids, rot, base, jac, prec, wait=True)
    assert info["state"] == 3 and info["block"] == 128
    P1, J1 = m.solve_fk(q, lids, jids, rot, base, jac, precision=prec)
    Pr, Jr = pair.oracle_solve(q, lids, jids, rot, base, jac)
    assert m.last_launch_info()["kernel"] == "tfk_spec_kernel (plan-specialised fk_twophase_kernel, NVRTC)"
    if prec == 0:
        assert np.array_equal(P0, P1) and (not jac or np.array_equal(J0, J1))
        parity.assert_bitwise(P1, J1 if jac else P1, Pr, Jr if jac else Pr, cfg + " specialised two-phase")
    else:  # FP32 mode: positions within 1e-5 of the oracle; against the generic FP32 kernel only contraction differs
        # (wide-pitch inputs: the RPY rows of J amplify a last-bit difference by 1 / cos(pitch)^2; the accuracy gate of the
        # mode is tests/test_gpu_parity.py::test_fp32_mode_within_1e5 against the FP64 reference)
        assert np.abs(P0 - P1).max() < 1e-5 and (not jac or np.abs(J0 - J1).max() < 1e-3)
        assert np.abs(P1[:, :3] - Pr[:, :3]).max() < 1e-5


@pytest.mark.gpu
@needs_nvrtc
@pytest.mark.parametrize("urdf,joints,link,rot,base,jac,prec", CASES)
def test_gpu_specialised_kernel_is_the_generic_kernel_bit_for_bit(urdf, joints, link, rot, base, jac, prec):
    m = capi.Model(W.urdf_text(urdf), host_libm=0)
    j, l = list(m.get_joint_ids(joints)), list(m.get_link_ids([link]))
    q = W.sample_q(m.get_joint_limits(j), base, 1000, seed=5, wide_pitch=True)  # ragged tail tile
    m.set_jit_mode(0)
    P0, J0 = m.solve_fk(q, l, j, rot, base, jac, precision=prec)
    assert "spec" not in m.last_launch_info()["kernel"]
    m.set_jit_mode(1)
    info = m.specialize(l, j, rot, base, jac, prec, wait=True)
    assert info["state"] == 3
    P1, J1 = m.solve_fk(q, l, j, rot, base, jac, precision=prec)
    li = m.last_launch_info()
    assert li["kernel"].startswith("tfk_spec_kernel") and li["block"] == info["block"]
    if prec == 0:  # strict FP64: no contraction, so constant folding cannot change a bit
        assert np.array_equal(P0, P1)
        if jac:
            assert np.array_equal(J0, J1)
    else:  # FP32 mode contracts multiply-adds; folding constants changes which ones -> the mode's own tolerance
        assert np.abs(P0 - P1).max() < 1e-5
        if jac:
            assert np.abs(J0 - J1).max() < 1e-4


@pytest.mark.gpu
@needs_nvrtc
@pytest.mark.parametrize("urdf,joints,link,rot,base", [c[:5] for c in CASES[:4]])
def test_gpu_specialised_strict_kernel_on_rare_inputs(urdf, joints, link, rot, base):
    """Rows that are rare for the portable libm and the shared-reciprocal divisions (huge angles, inf, nan, angles within
    2^-40 of a multiple of pi/2, denormals) scattered over a batch -- alone in a warp, many in a warp, in the ragged tail --
    next to rows of exact zeros (a robot at its zero configuration): the specialised kernel's out-of-line rare paths must
    give the generic kernel's result bit for bit (NaN where it has NaN)."""
    m = capi.Model(W.urdf_text(urdf), host_libm=0)
    j, l = list(m.get_joint_ids(joints)), list(m.get_link_ids([link]))
    n = 3000 + 17
    q = W.sample_q(m.get_joint_limits(j), base, n, seed=11, wide_pitch=True)
    rng = np.random.default_rng(12)
    rare = [1e7, -3.3e9, np.inf, -np.inf, np.nan, np.pi / 2, -np.pi, 3 * np.pi / 2 * (1 + 2.0 ** -45), 1e-310, 5e-324, 2.0 ** 40, 1647099.5]
    rows = np.concatenate([rng.choice(n, 60, replace=False), np.arange(1024, 1040), [n - 1, n - 2, 0]])
    for r in rows:
        q[r, rng.integers(0, q.shape[1])] = rare[rng.integers(0, len(rare))]
    q[200:232] = 0.0      # a whole warp at the zero configuration
    q[500] = 0.0
    q[777, :len(j)] = 0.0  # zero joints, random base
    m.set_jit_mode(0)
    P0, J0 = m.solve_fk(q, l, j, rot, base, True)
    assert "spec" not in m.last_launch_info()["kernel"]
    m.set_jit_mode(1)
    assert m.specialize(l, j, rot, base, True, 0, wait=True)["state"] == 3
    P1, J1 = m.solve_fk(q, l, j, rot, base, True)
    assert m.last_launch_info()["kernel"].startswith("tfk_spec_kernel")
    assert np.array_equal(P0, P1, equal_nan=True) and np.array_equal(J0, J1, equal_nan=True)
    clean = np.setdiff1d(np.arange(n), rows)
    assert np.isfinite(P1[clean]).all()


@pytest.mark.gpu
@needs_nvrtc
def test_gpu_auto_mode_compiles_in_the_background_and_switches_over():
    import time
    m = capi.Model(W.urdf_text("kuka.urdf"), host_libm=0)
    j, l = list(m.get_joint_ids(W.KUKA_JOINTS)), list(m.get_link_ids(["lbr_iiwa_link_7"]))
    assert m.get_jit_mode() == 1
    q = W.sample_q(m.get_joint_limits(j), 0, 1 << 16, seed=1)
    ref = None
    kernels = []
    t0 = time.time()
    for it in range(4000):  # 2^22 configurations start the compile; it finishes within seconds
        P, J = m.solve_fk(q, l, j, 2, 0, True)
        k = m.last_launch_info()["kernel"]
        if ref is None:
            ref = (P.copy(), J.copy())
        if not kernels or kernels[-1] != k:
            kernels.append(k)
            assert np.array_equal(P, ref[0]) and np.array_equal(J, ref[1])
        if k.startswith("tfk_spec_kernel") or time.time() - t0 > 120:
            break
    assert len(kernels) == 2 and kernels[0].startswith("fk_fused_kernel") and kernels[1].startswith("tfk_spec_kernel"), kernels


@pytest.mark.gpu
@needs_nvrtc
@pytest.mark.parametrize("urdf", ["pr2.urdf", "fetch.urdf"])
def test_gpu_specialised_kernel_random_queries_vs_oracle(urdf):
    """random single-link queries (link anywhere in the tree, random joint subsets in random order incl. joints that do
    not move the link, every rot/base type, non-zero stored state) through TFK_JIT_ON, bit for bit against the oracle"""
    import parity
    pair = parity.Pair(urdf, strict=True)
    m = pair.prod
    rng = np.random.default_rng(7)
    n_links, n_joints = len(m.get_link_names()), len(m.get_joint_names())
    all_j = list(range(n_joints))
    pair.set_state(all_j, rng.uniform(-0.3, 0.3, n_joints), [0.3, -0.2, 0.1, 0.2, -0.1, 0.4])
    m.set_jit_mode(2)
    done = 0
    for trial in range(40):
        lid = int(rng.integers(0, n_links))
        jids = [int(x) for x in rng.permutation(n_joints)[: int(rng.integers(1, min(n_joints, 12)))]]
        rot, base, jac = int(rng.integers(0, 3)), int(rng.integers(0, 3)), bool(rng.integers(0, 2))
        lim = m.get_joint_limits(jids)
        q = W.sample_q(lim, base, 300, seed=trial, wide_pitch=True)
        P, J = pair.prod_solve(q, [lid], jids, rot, base, jac)
        if not m.last_launch_info()["kernel"].startswith("tfk_spec_kernel"):
            continue  # not eligible (chain too long for the unrolled build): served by the generic kernel
        Pr, Jr = pair.oracle_solve(q, [lid], jids, rot, base, jac)
        parity.assert_bitwise(P, J if jac else P, Pr, Jr if jac else Pr, "trial %d" % trial)
        done += 1
        if done >= 10:
            break
    assert done >= 6
