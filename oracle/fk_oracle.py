"""CPU restatement (numpy float64, 4x4 homogeneous transforms) of the serial-chain forward kinematics the
robot cost uses.  TEST INFRASTRUCTURE ONLY -- nothing under ``sigsvgd_b200/`` imports this.

Follows ``Robot.qs_to_joints_xs`` (src/models/robot/robot_simulator.py:138-170): positions of the child-link
frames of ``compute_forward_kinematics_all_links``.  That routine lives in the third-party package
``differentiable_robot_model`` (reference ``setup.py``; absent here, no network): **parity unpinned** against
it.  What is restated is the URDF convention itself (joint origin ``xyz`` / ``rpy`` = fixed-axis roll-pitch-yaw,
then the rotation about ``axis`` by the joint angle), pinned by known answers (tests/test_oracle_fk.py): the
Franka Emika flange at the zero configuration, x = 0.088 m, z = 0.333 + 0.316 + 0.384 - 0.107 = 0.926 m; the
position of a link does not depend on its own joint; distances between consecutive frames are the URDF
offsets' lengths in every configuration; the torch-autograd gradient of the same recursion."""
from __future__ import annotations

import numpy as np


def rpy_matrix(r, p, y):
    rx = np.array([[1, 0, 0], [0, np.cos(r), -np.sin(r)], [0, np.sin(r), np.cos(r)]])
    ry = np.array([[np.cos(p), 0, np.sin(p)], [0, 1, 0], [-np.sin(p), 0, np.cos(p)]])
    rz = np.array([[np.cos(y), -np.sin(y), 0], [np.sin(y), np.cos(y), 0], [0, 0, 1]])
    return rz @ ry @ rx


def axis_rotation(axis, q):
    a = np.asarray(axis, np.float64)
    a = a / np.linalg.norm(a)
    K = np.array([[0, -a[2], a[1]], [a[2], 0, -a[0]], [-a[1], a[0], 0]])
    return np.eye(3) + np.sin(q) * K + (1 - np.cos(q)) * (K @ K)


def chain_fk(xyz, rpy, axis, dof, q):
    """positions [J, B, 3] of the child-link frames for configurations q [B, n_dof]."""
    q = np.asarray(q, np.float64)
    J, B = len(xyz), q.shape[0]
    out = np.zeros((J, B, 3))
    for b in range(B):
        T = np.eye(4)
        for j in range(J):
            O = np.eye(4)
            O[:3, :3], O[:3, 3] = rpy_matrix(*rpy[j]), xyz[j]
            T = T @ O
            out[j, b] = T[:3, 3]
            if dof[j] >= 0:
                Rq = np.eye(4)
                Rq[:3, :3] = axis_rotation(axis[j], q[b, dof[j]])
                T = T @ Rq
    return out


def chain_fk_torch(xyz, rpy, axis, dof, q):
    """the same recursion in torch float64 (autograd-able), batched."""
    import torch

    B = q.shape[0]
    R = torch.eye(3, dtype=q.dtype).expand(B, 3, 3)
    p = torch.zeros(B, 3, dtype=q.dtype)
    out = []
    for j in range(len(xyz)):
        p = p + R @ torch.tensor(xyz[j], dtype=q.dtype)
        R = R @ torch.tensor(rpy_matrix(*rpy[j]), dtype=q.dtype)
        out.append(p)
        if dof[j] >= 0:
            a = torch.tensor(np.asarray(axis[j], np.float64) / np.linalg.norm(axis[j]), dtype=q.dtype)
            K = torch.tensor([[0, -a[2], a[1]], [a[2], 0, -a[0]], [-a[1], a[0], 0]], dtype=q.dtype)
            ang = q[:, dof[j]][:, None, None]
            R = R @ (torch.eye(3, dtype=q.dtype) + torch.sin(ang) * K + (1 - torch.cos(ang)) * (K @ K))
    return torch.stack(out)
