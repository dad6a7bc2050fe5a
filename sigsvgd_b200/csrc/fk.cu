// Forward kinematics of a serial chain (robot-planning cost): link-frame positions of every
// configuration and the gradient w.r.t. the joint angles.  Replaces Robot.qs_to_joints_xs
// (src/models/robot/robot_simulator.py:138-170), which asks differentiable_robot_model for
// compute_forward_kinematics_all_links (a chain of batched 4x4 products, one group of torch kernels per joint)
// and keeps the positions, and autograd through it (examples/script_planning_robot.py:107).
//
// URDF semantics, joint j of the chain (parent -> child):  T_child = T_parent . [R_rpy(j) | xyz(j)] . Rot(axis_j, q)
// (fixed joints: no rotation).  One thread per configuration walks the chain with a 3x3 rotation and a position
// in registers; the backward kernel re-walks it, keeps the joint origins o_j and world axes z_j, and forms
//   d/dq_j sum_l g_l . p_l = z_j . sum_{l > j} (p_l - o_j) x g_l          (geometric Jacobian, suffix sums)
// so nothing is saved between the two launches.  HBM: 4 dof bytes in, 12 J bytes out per configuration.
#include "common.cuh"

namespace sigsvgd {
namespace {

constexpr int kMaxJoints = 16;

struct ChainArgs {
  float xyz[kMaxJoints][3];   // origin translation of joint j in its parent frame
  float rot[kMaxJoints][9];   // origin rotation R_rpy(j), row-major
  float axis[kMaxJoints][3];  // unit rotation axis (child frame); unused for fixed joints
  int dof[kMaxJoints];        // column of q that drives joint j, or -1 (fixed)
  int J, n_dof;
};

struct Mat3 {
  float m[9];
};
__device__ __forceinline__ Mat3 mul(const Mat3& a, const float* b) {
  Mat3 c;
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int k = 0; k < 3; ++k) c.m[3 * i + k] = a.m[3 * i] * b[k] + a.m[3 * i + 1] * b[3 + k] + a.m[3 * i + 2] * b[6 + k];
  return c;
}
// Rodrigues: rotation by angle q about the unit axis a
__device__ __forceinline__ void axis_angle(const float* a, float q, float* r) {
  float s, c;
  sincosf(q, &s, &c);
  const float t = 1.f - c;
  r[0] = c + t * a[0] * a[0], r[1] = t * a[0] * a[1] - s * a[2], r[2] = t * a[0] * a[2] + s * a[1];
  r[3] = t * a[0] * a[1] + s * a[2], r[4] = c + t * a[1] * a[1], r[5] = t * a[1] * a[2] - s * a[0];
  r[6] = t * a[0] * a[2] - s * a[1], r[7] = t * a[1] * a[2] + s * a[0], r[8] = c + t * a[2] * a[2];
}

__global__ void __launch_bounds__(128)
chain_fk_kernel(const __grid_constant__ ChainArgs ch, const float* __restrict__ q, int64_t B, float* __restrict__ xs) {
  const int64_t b = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (b >= B) return;
  Mat3 R = {{1.f, 0.f, 0.f, 0.f, 1.f, 0.f, 0.f, 0.f, 1.f}};
  float p[3] = {0.f, 0.f, 0.f};
  for (int j = 0; j < ch.J; ++j) {
#pragma unroll
    for (int i = 0; i < 3; ++i)
      p[i] += R.m[3 * i] * ch.xyz[j][0] + R.m[3 * i + 1] * ch.xyz[j][1] + R.m[3 * i + 2] * ch.xyz[j][2];
    R = mul(R, ch.rot[j]);
    float* o = xs + ((int64_t)j * B + b) * 3;
    o[0] = p[0], o[1] = p[1], o[2] = p[2];
    if (ch.dof[j] >= 0) {
      float r[9];
      axis_angle(ch.axis[j], __ldg(q + b * ch.n_dof + ch.dof[j]), r);
      R = mul(R, r);
    }
  }
}

__global__ void __launch_bounds__(128)
chain_fk_bwd_kernel(const __grid_constant__ ChainArgs ch, const float* __restrict__ q, int64_t B,
                    const float* __restrict__ gxs, float* __restrict__ gq) {
  const int64_t b = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (b >= B) return;
  Mat3 R = {{1.f, 0.f, 0.f, 0.f, 1.f, 0.f, 0.f, 0.f, 1.f}};
  float p[3] = {0.f, 0.f, 0.f};
  float org[kMaxJoints][3], zw[kMaxJoints][3];  // joint origins and world axes (local memory)
  for (int j = 0; j < ch.J; ++j) {
#pragma unroll
    for (int i = 0; i < 3; ++i)
      p[i] += R.m[3 * i] * ch.xyz[j][0] + R.m[3 * i + 1] * ch.xyz[j][1] + R.m[3 * i + 2] * ch.xyz[j][2];
    R = mul(R, ch.rot[j]);
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      org[j][i] = p[i];
      zw[j][i] = R.m[3 * i] * ch.axis[j][0] + R.m[3 * i + 1] * ch.axis[j][1] + R.m[3 * i + 2] * ch.axis[j][2];
    }
    if (ch.dof[j] >= 0) {
      float r[9];
      axis_angle(ch.axis[j], __ldg(q + b * ch.n_dof + ch.dof[j]), r);
      R = mul(R, r);
    }
  }
  for (int i = 0; i < ch.n_dof; ++i) gq[b * ch.n_dof + i] = 0.f;
  // suffix sums over the links after joint j:  G = sum g_l,  M = sum p_l x g_l
  float G[3] = {0.f, 0.f, 0.f}, M[3] = {0.f, 0.f, 0.f};
  for (int j = ch.J - 1; j >= 0; --j) {
    if (ch.dof[j] >= 0) {
      // sum (p_l - o_j) x g_l = M - o_j x G
      const float* o = org[j];
      const float v0 = M[0] - (o[1] * G[2] - o[2] * G[1]);
      const float v1 = M[1] - (o[2] * G[0] - o[0] * G[2]);
      const float v2 = M[2] - (o[0] * G[1] - o[1] * G[0]);
      gq[b * ch.n_dof + ch.dof[j]] += zw[j][0] * v0 + zw[j][1] * v1 + zw[j][2] * v2;
    }
    // link j's own position does not depend on q_j: it joins the sums for the joints before it
    const float* g = gxs + ((int64_t)j * B + b) * 3;
    const float g0 = __ldg(g), g1 = __ldg(g + 1), g2 = __ldg(g + 2);
    const float* pl = org[j];
    G[0] += g0, G[1] += g1, G[2] += g2;
    M[0] += pl[1] * g2 - pl[2] * g1, M[1] += pl[2] * g0 - pl[0] * g2, M[2] += pl[0] * g1 - pl[1] * g0;
  }
}

int fill_chain(ChainArgs* ch, const float* xyz, const float* rot, const float* axis, const int* dof, int J, int n_dof) {
  SIG_CHECK_ARG(J >= 1 && J <= kMaxJoints, "chain_fk: 1 .. %d joints (got %d)", kMaxJoints, J);
  SIG_CHECK_ARG(n_dof >= 1 && n_dof <= kMaxJoints, "chain_fk: 1 .. %d degrees of freedom (got %d)", kMaxJoints, n_dof);
  SIG_CHECK_ARG(xyz && rot && axis && dof, "chain_fk: NULL chain description");
  ch->J = J, ch->n_dof = n_dof;
  for (int j = 0; j < J; ++j) {
    SIG_CHECK_ARG(dof[j] >= -1 && dof[j] < n_dof, "chain_fk: joint %d drives column %d of %d", j, dof[j], n_dof);
    for (int i = 0; i < 3; ++i) ch->xyz[j][i] = xyz[3 * j + i], ch->axis[j][i] = axis[3 * j + i];
    for (int i = 0; i < 9; ++i) ch->rot[j][i] = rot[9 * j + i];
    ch->dof[j] = dof[j];
  }
  return 0;
}

}  // namespace

int launch_chain_fk(const float* xyz, const float* rot, const float* axis, const int* dof, int J, int n_dof,
                    const float* q, int64_t B, float* xs, cudaStream_t st) {
  ChainArgs ch;
  SIG_PROPAGATE(fill_chain(&ch, xyz, rot, axis, dof, J, n_dof));
  SIG_CHECK_ARG(B >= 0 && B < (1ll << 31) * 128, "chain_fk: bad number of configurations");
  if (B == 0) return 0;
  SIG_CHECK_ARG(q && xs, "chain_fk: NULL pointer");
  chain_fk_kernel<<<(unsigned)ceil_div(B, 128), 128, 0, st>>>(ch, q, B, xs);
  SIG_LAUNCH_OK("chain_fk_kernel");
  return 0;
}

int launch_chain_fk_bwd(const float* xyz, const float* rot, const float* axis, const int* dof, int J, int n_dof,
                        const float* q, int64_t B, const float* gxs, float* gq, cudaStream_t st) {
  ChainArgs ch;
  SIG_PROPAGATE(fill_chain(&ch, xyz, rot, axis, dof, J, n_dof));
  SIG_CHECK_ARG(B >= 0 && B < (1ll << 31) * 128, "chain_fk_bwd: bad number of configurations");
  if (B == 0) return 0;
  SIG_CHECK_ARG(q && gxs && gq, "chain_fk_bwd: NULL pointer");
  chain_fk_bwd_kernel<<<(unsigned)ceil_div(B, 128), 128, 0, st>>>(ch, q, B, gxs, gq);
  SIG_LAUNCH_OK("chain_fk_bwd_kernel");
  return 0;
}

}  // namespace sigsvgd
