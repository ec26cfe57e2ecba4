"""Every kernel variant the engine can pick (tinyfk_b200/csrc/fk_launch.h KernelVariant) is exercised on purpose and
compared with the oracle bit for bit.  A synthetic 20-joint "snake" (axes and origins of every sparsity class, one
prismatic joint) reaches the variants the bundled robots cannot: joint records beyond shared memory."""
import numpy as np
import pytest

import parity
from emuapi import EmuModel
from oracle import refapi
from tinyfk_b200 import capi
from tinyfk_b200 import workloads as W

needs_ref = pytest.mark.skipif(not refapi.have_ref(True), reason="oracle/_ref not built")


def snake_urdf(n=20):
    axes = ["0 0 1", "0 1 0", "1 0 0", "0 0 -1", "0.6 0 0.8", "0 -1 0"]
    xyz = ["0.1 0 0", "0 0 0.2", "0 0.05 0", "0.03 -0.02 0.07", "0 0 0"]
    rpy = ["0 0 0", "0 0 0", "0.3 0 0", "0 0 0", "0.1 -0.2 0.3"]
    s = ['<robot name="snake">', '<link name="l00"><inertial><origin xyz="0 0 0.01"/><mass value="1.5"/>'
         '<inertia ixx="0.01" ixy="0" ixz="0" iyy="0.02" iyz="0" izz="0.03"/></inertial></link>']
    for i in range(1, n + 1):
        s.append('<link name="l%02d"><inertial><origin xyz="0.01 0 0"/><mass value="%g"/>'
                 '<inertia ixx="0.001" ixy="0.0001" ixz="0" iyy="0.002" iyz="0" izz="0.003"/></inertial></link>' % (i, 0.5 + 0.1 * i))
        kind = "prismatic" if i == 7 else ("continuous" if i % 5 == 0 else "revolute")
        lim = "" if kind == "continuous" else '<limit lower="-1.2" upper="1.3" effort="1" velocity="1"/>'
        s.append('<joint name="j%02d" type="%s"><parent link="l%02d"/><child link="l%02d"/><origin xyz="%s" rpy="%s"/>'
                 '<axis xyz="%s"/>%s</joint>' % (i, kind, i - 1, i, xyz[i % 5], rpy[i % 5], axes[i % 6], lim))
    s.append("</robot>")
    return "\n".join(s)


@needs_ref
@pytest.mark.parametrize("rot,base", [(1, 2), (2, 1), (0, 0)])
def test_snake_core_equals_pinned_reference_on_cpu(rot, base):
    xml = snake_urdf()
    e, o = EmuModel(xml), refapi.RefModel(xml, strict=True)
    jids = list(e.get_joint_ids(["j%02d" % i for i in range(1, 21)]))
    lid = e.get_link_ids(["l20"])
    q = W.sample_q(o.get_joint_limits(jids), base, 12, seed=rot + 3 * base, wide_pitch=True)
    Pr, Jr = o.solve_fk(q, lid, jids, rot, base, True)
    P, J = e.solve_fk(q, lid, jids, rot, base, True)
    parity.assert_bitwise(P, J, Pr, Jr, "snake")


class _Snake(parity.Pair):
    def __init__(self):
        xml = snake_urdf()
        self.strict = True
        self.prod = capi.Model(xml, host_libm=0)
        self._ref = refapi.RefModel(xml, strict=True)
        self._port = None


@pytest.mark.gpu
@needs_ref
def test_gpu_every_kernel_variant_against_the_oracle():
    seen = {}

    def check(pair, q, lids, jids, rot, base, jac, what):
        P, J = pair.prod_solve(q, lids, jids, rot, base, jac)
        k = pair.prod.last_launch_info()["kernel"]
        Pr, Jr = pair.oracle_solve(q, lids, jids, rot, base, jac)
        parity.assert_bitwise(P, J if jac else P, Pr, Jr if jac else Pr, what + " [" + k + "]")
        seen[k] = seen.get(k, 0) + 1
        return k

    snake = _Snake()
    snake.prod.set_jit_mode(0)
    sj = list(snake.prod.get_joint_ids(["j%02d" % i for i in range(1, 21)]))
    lim = snake.prod.get_joint_limits(sj)
    # 20 joints, one link, FLOATING + RPY: 127 record reals per configuration -> block-private global records
    q = W.sample_q(lim, 2, 700, seed=1, wide_pitch=True)
    assert check(snake, q, snake.prod.get_link_ids(["l20"]), sj, 1, 2, True, "snake tip") == "fk_fused_kernel<global records>"
    # all 21 links: two-phase, joint records too large for shared memory
    allsl = snake.prod.get_link_ids(["l%02d" % i for i in range(21)])
    assert check(snake, q[:200], allsl, sj, 2, 2, True, "snake all links") == "fk_twophase_kernel"
    # the first 6 joints only, 5 links: two-phase with the joint records in shared memory
    q6 = W.sample_q(lim[:6], 1, 500, seed=2)
    k = check(snake, q6, snake.prod.get_link_ids(["l02", "l04", "l06", "l09", "l20"]), sj[:6], 1, 1, True, "snake 5 links")
    assert k == "fk_twophase_kernel<joint records in shared memory>"

    pr2 = parity.Pair("pr2.urdf", strict=True)
    pr2.prod.set_jit_mode(0)
    j = list(pr2.prod.get_joint_ids(W.PR2_RARM))
    l = pr2.prod.get_link_ids(["r_gripper_tool_frame"])
    qp = W.sample_q(pr2.prod.get_joint_limits(j), 1, 900, seed=3, wide_pitch=True)
    assert check(pr2, qp, l, j, 1, 1, True, "cfg2") == "fk_fused_kernel<smem records>"
    assert check(pr2, qp[:, :7], l, j, 2, 0, False, "pose only") == "fk_fused_kernel<smem records, 5 blocks/SM>"
    pr2.prod.set_jit_mode(2)
    assert check(pr2, qp, l, j, 1, 1, True, "cfg2 specialised").startswith("tfk_spec_kernel")
    # the specialised build of the long chain: records fit only once -> 128-thread blocks or not eligible; either way exact
    snake.prod.set_jit_mode(2)
    q8 = W.sample_q(lim[:8], 0, 300, seed=4)
    check(snake, q8, snake.prod.get_link_ids(["l08"]), sj[:8], 2, 0, True, "snake 8 joints specialised")
    assert len(seen) >= 6, seen
