#!/usr/bin/env python3
"""Generates tests/golden/ref_<cfg>_{stock,pinned}.npz from the reference's OWN object code
(oracle/_ref, built from /root/reference by oracle/Makefile) -- run in the build container, where the
reference exists; the fixtures are committed because /root/reference does not travel to the GPU box.

  stock  : reference + the container's glibc libm       (tolerance-tier checks)
  pinned : reference + tinyfk_b200/csrc/fkmath.h libm    (bit-exact checks, finite-difference columns included)

Each file holds: q (inputs), P, J (reference outputs, as Python sees them), joint_names, link_names,
added links, rot/base type, and -- for the non-trivial stored state cases -- the stored angles / base.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
from oracle.refapi import RefModel  # noqa: E402
from tinyfk_b200 import workloads as W  # noqa: E402

N = {"cfg1": 1, "cfg2": 64, "cfg3": 64, "cfg4": 16, "cfg5": 8}


def main():
    for cfg in ("cfg1", "cfg2", "cfg3", "cfg4", "cfg5"):
        urdf, jn, ln, rot, base, _ = W.CONFIGS[cfg]
        for strict in (False, True):
            m = RefModel(W.urdf_text(urdf), strict=strict)
            jids, lids, rot, base = W.setup(m, cfg)
            q = np.array([W.README_Q]) if cfg == "cfg1" else W.sample_q(m.get_joint_limits(jids), base, N[cfg], seed=42)
            P, J = m.solve_fk(q, lids, jids, rot, base, True)
            names = m.get_link_names()
            out = os.path.join(HERE, "ref_%s_%s.npz" % (cfg, "pinned" if strict else "stock"))
            np.savez_compressed(out, q=q, P=P, J=J, joint_ids=np.array(jids), link_ids=np.array(lids),
                                joint_names=np.array(jn), link_names=np.array([names[i] for i in lids]),
                                rot=rot, base=base)
            print(out, P.shape, J.shape)
    # the reference's own golden case (test/data/test_data.json): PR2 + "mylink", floating base, with Jacobian
    import json
    td = json.load(open(os.path.join(HERE, "reference_test_data.json")))
    for strict in (False, True):
        m = RefModel(W.urdf_text("pr2.urdf"), strict=strict)
        pid = m.get_link_ids(["r_upper_arm_link"])[0]
        m.add_new_link("mylink", pid, [0.1, 0.1, 0.1], [0.3, 0.2, 0.1])
        jids, lids = m.get_joint_ids(td["joint_names"]), m.get_link_ids(td["link_names"])
        for rot in (0, 1, 2):
            P, J = m.solve_fk([td["angle_vector"]], lids, jids, rot, 2, True)
            out = os.path.join(HERE, "ref_testdata_rot%d_%s.npz" % (rot, "pinned" if strict else "stock"))
            np.savez_compressed(out, q=np.array([td["angle_vector"]]), P=P, J=J, joint_ids=np.array(jids),
                                link_ids=np.array(lids), rot=rot, base=2)
            print(out, P.shape, J.shape)


if __name__ == "__main__":
    main()
