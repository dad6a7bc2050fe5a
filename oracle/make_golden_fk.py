"""The Panda chain of the reference (robot_resources/panda/urdf/panda.urdf, link0 -> panda_hand, the joints of
``PandaRobot.target_joint_names``, src/models/robot/robot_simulator.py:315-349) as numbers, so that the GPU box
(no /root/reference) can test the forward-kinematics kernels on it, plus float64 oracle positions:

    python -m oracle.make_golden_fk        # from the repo root; needs /root/reference
"""
from __future__ import annotations

import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
OUT = os.path.join(ROOT, "tests", "golden", "panda_chain.npz")
URDF = "/root/reference/robot_resources/panda/urdf/panda.urdf"
JOINTS = [f"panda_joint{i}" for i in range(1, 8)]
LINKS = [f"panda_link{i}" for i in range(1, 9)] + ["panda_hand"]
LIMITS_LO = [-2.9671, -1.8326, -2.9671, -3.1416, -2.9671, -0.0873, -2.9671]
LIMITS_HI = [2.9671, 1.8326, 2.9671, 0.0873, 2.9671, 3.8223, 2.9671]


def main():
    sys.path.insert(0, ROOT)
    import xml.etree.ElementTree as ET

    from oracle import fk_oracle

    root = ET.parse(URDF).getroot()
    by_child = {j.find("child").get("link"): j for j in root.findall("joint")}
    path, link = [], "panda_hand"
    while link in by_child:
        path.append(by_child[link])
        link = by_child[link].find("parent").get("link")
    path.reverse()
    xyz = np.array([[float(v) for v in j.find("origin").get("xyz").split()] for j in path])
    rpy = np.array([[float(v) for v in j.find("origin").get("rpy").split()] for j in path])
    axis = np.array([[float(v) for v in j.find("axis").get("xyz").split()] if j.find("axis") is not None else [0, 0, 1]
                     for j in path])
    dof = np.array([JOINTS.index(j.get("name")) if j.get("name") in JOINTS else -1 for j in path])
    names = [j.find("child").get("link") for j in path]
    assert names == LINKS, names
    rng = np.random.default_rng(0)
    q = rng.uniform(LIMITS_LO, LIMITS_HI, size=(64, 7))
    q[0] = 0.0
    xs = fk_oracle.chain_fk(xyz, rpy, axis, dof, q)
    np.savez(OUT, xyz=xyz, rpy=rpy, axis=axis, dof=dof, link_names=np.array(names), q=q, xs=xs,
             limits_lo=np.array(LIMITS_LO), limits_hi=np.array(LIMITS_HI))
    print(names, "flange at q = 0:", xs[names.index("panda_link8"), 0], "->", OUT)


if __name__ == "__main__":
    main()
