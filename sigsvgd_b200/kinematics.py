"""Forward kinematics of a serial chain on the device.

The reference's robot cost turns joint configurations into link positions with
``Robot.qs_to_joints_xs`` (src/models/robot/robot_simulator.py:138-170) = ``differentiable_robot_model``'s
``compute_forward_kinematics_all_links`` on the Panda URDF (robot_resources/panda/urdf/panda.urdf), keeping the
positions of the target links, and differentiates through it (examples/script_planning_robot.py:107).
``SerialChain`` holds the URDF's joint origins / axes along base -> end link and answers from one CUDA launch
(one more for the gradient w.r.t. the joint angles).  There is no CPU path."""
from __future__ import annotations

import ctypes
import math
import xml.etree.ElementTree as ET
from typing import List, Optional, Sequence

import numpy as np
import torch

from . import ops


def _rpy_matrix(r, p, y):
    cr, sr, cp, sp, cy, sy = math.cos(r), math.sin(r), math.cos(p), math.sin(p), math.cos(y), math.sin(y)
    return np.array([[cy * cp, cy * sp * sr - sy * cr, cy * sp * cr + sy * sr],
                     [sy * cp, sy * sp * sr + cy * cr, sy * sp * cr - cy * sr],
                     [-sp, cp * sr, cp * cr]])


class _ChainFK(torch.autograd.Function):
    @staticmethod
    def forward(ctx, q2d, chain):
        q = q2d.detach().to(torch.float32).contiguous()
        B = q.shape[0]
        xs = torch.empty(chain.n_joints, B, 3, dtype=torch.float32, device=q.device)
        ops.call("sigsvgd_chain_fk", *chain._host_ptrs(), chain.n_joints, chain.n_dof, ops._p(q), B, ops._p(xs), ops._stream())
        ctx.chain, ctx.q, ctx.dtype = chain, q, q2d.dtype
        return xs.to(q2d.dtype)

    @staticmethod
    def backward(ctx, gxs):
        chain, q = ctx.chain, ctx.q
        g = gxs.to(torch.float32).contiguous()
        gq = torch.empty_like(q)
        ops.call("sigsvgd_chain_fk_bwd", *chain._host_ptrs(), chain.n_joints, chain.n_dof, ops._p(q), q.shape[0], ops._p(g),
                 ops._p(gq), ops._stream())
        return gq.to(ctx.dtype), None


class SerialChain:
    """Joints of one base -> end-link path: ``xyz [J,3]``, ``rpy [J,3]`` (URDF joint origins), ``axis [J,3]``,
    ``dof [J]`` (column of the configuration vector that drives the joint, -1 for fixed joints)."""

    def __init__(self, xyz, rpy, axis, dof, link_names: Optional[Sequence[str]] = None):
        self.xyz = np.ascontiguousarray(xyz, np.float32).reshape(-1, 3)
        J = self.xyz.shape[0]
        rpy = np.asarray(rpy, np.float64).reshape(J, 3)
        self.rot = np.ascontiguousarray(np.stack([_rpy_matrix(*r) for r in rpy]).reshape(J, 9), np.float32)
        ax = np.asarray(axis, np.float64).reshape(J, 3)
        nrm = np.linalg.norm(ax, axis=1, keepdims=True)
        self.axis = np.ascontiguousarray(np.where(nrm > 0, ax / np.maximum(nrm, 1e-30), 0.0), np.float32)
        self.dof = np.ascontiguousarray(dof, np.int32).reshape(J)
        self.n_joints, self.n_dof = J, int(self.dof.max()) + 1
        self.link_names = list(link_names) if link_names is not None else [f"link{j + 1}" for j in range(J)]
        if self.n_dof < 1:
            raise ValueError("SerialChain: no movable joint")

    @classmethod
    def from_urdf(cls, urdf_path: str, end_link: str, joint_names: Sequence[str], base_link: Optional[str] = None):
        """The path base -> ``end_link`` of a URDF; ``joint_names[i]`` is driven by column i of the configuration
        (the reference's ``target_joint_names``); every other joint on the path is held at zero."""
        root = ET.parse(urdf_path).getroot()
        by_child = {j.find("child").get("link"): j for j in root.findall("joint")}
        path, link = [], end_link
        while link in by_child and link != base_link:
            path.append(by_child[link])
            link = by_child[link].find("parent").get("link")
        path.reverse()
        if not path:
            raise ValueError(f"SerialChain: no joint leads to link {end_link!r}")
        xyz, rpy, axis, dof, names = [], [], [], [], []
        for j in path:
            o = j.find("origin")
            xyz.append([float(v) for v in (o.get("xyz", "0 0 0") if o is not None else "0 0 0").split()])
            rpy.append([float(v) for v in (o.get("rpy", "0 0 0") if o is not None else "0 0 0").split()])
            a = j.find("axis")
            axis.append([float(v) for v in (a.get("xyz") if a is not None else "0 0 1").split()])
            kind, name = j.get("type"), j.get("name")
            if kind in ("revolute", "continuous") and name in joint_names:
                dof.append(list(joint_names).index(name))
            elif kind in ("revolute", "continuous", "fixed"):
                dof.append(-1)
            else:
                raise NotImplementedError(f"SerialChain: joint {name!r} of type {kind!r} on the path")
            names.append(j.find("child").get("link"))
        return cls(xyz, rpy, axis, dof, names)

    def _host_ptrs(self):
        return tuple(a.ctypes.data_as(ctypes.c_void_p) for a in (self.xyz, self.rot, self.axis, self.dof))

    def positions(self, q: torch.Tensor, links: Optional[Sequence[str]] = None) -> torch.Tensor:
        """``q [..., n_dof] -> [n_links, ..., 3]`` (the layout of ``qs_to_joints_xs``: links first)."""
        if not q.is_cuda:
            raise RuntimeError("SerialChain: CUDA tensors only; sigsvgd_b200 has no CPU fallback")
        if q.shape[-1] != self.n_dof:
            raise ValueError(f"SerialChain: configurations [..., {self.n_dof}] expected, got {tuple(q.shape)}")
        lead = q.shape[:-1]
        xs = _ChainFK.apply(q.reshape(-1, self.n_dof), self)
        if links is not None:
            xs = xs[[self.link_names.index(n) for n in links]]
        return xs.reshape(xs.shape[0], *lead, 3)

    __call__ = positions


__all__ = ["SerialChain"]
