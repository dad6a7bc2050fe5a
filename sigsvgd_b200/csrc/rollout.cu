// Stein-MPC rollout of the particle model on the device.
//
// Replaces the Python loop of DuSt._rollout (src/controllers/dust.py:293-300), which calls
// ParticleModel.step (src/models/particle.py:123-173, control_type "acceleration", deterministic) once
// per horizon step -- ~10 torch launches per step, 30 steps per rollout batch -- and the autograd graph
// torch records through it (the kernel terms of TrajectorySVGD differentiate the rollouts w.r.t. the
// actions, src/inference/trajectory_svgd.py:55-69):
//
//   a      = clamp(act / m, -max_acc, max_acc)
//   mask   = can_crash ? map[floor(x / cell + cx), floor(y / cell + cy)] (indices clamped to the grid) : 0
//            (ObstacleMap.get_collisions, src/utils/obstacle_map.py:68-97)
//   s'     = s + ((vx, vy, ax, ay) * dt) * (1 - mask)
//   s'.v   = clamp(s'.v, -max_speed, max_speed)
//
// One thread per rollout, the whole horizon in registers; the fp32 operations are issued in the
// reference's order WITHOUT fused multiply-adds, so the states are bit-identical to the reference's CPU
// result.  The forward also records, per step, what the backward needs (crash factor, which clamps were
// active); the backward is the reverse recursion of the same linear map.
//
// Bound: latency (R threads, H dependent steps); 16 R (H + 1) + 8 R H bytes of HBM traffic.
#include "common.cuh"

namespace sigsvgd {
namespace {

struct RolloutArgs {
  const float* init;   // [4] x, y, vx, vy
  const float* mass;   // [R] or NULL (mass_scalar)
  float mass_scalar, dt, max_acc, max_speed;
  const uint8_t* grid;  // [gx, gy] occupancy (0 / 1) or NULL
  int gx, gy;
  float inv_cell, off_x, off_y;
  int can_crash;
  int64_t R;
  int H;
};

constexpr uint8_t kCrash = 1, kClampVx = 2, kClampVy = 4, kClampAx = 8, kClampAy = 16;

__global__ void __launch_bounds__(128)
particle_rollout_kernel(RolloutArgs p, const float* __restrict__ actions, float* __restrict__ states,
                        uint8_t* __restrict__ flags) {
  const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= p.R) return;
  const float m = p.mass ? __ldg(p.mass + r) : p.mass_scalar;
  float x = __ldg(p.init), y = __ldg(p.init + 1), vx = __ldg(p.init + 2), vy = __ldg(p.init + 3);
  float4* out = reinterpret_cast<float4*>(states + r * (int64_t)(p.H + 1) * 4);
  const float2* act = reinterpret_cast<const float2*>(actions + r * (int64_t)p.H * 2);
  out[0] = make_float4(x, y, vx, vy);
  for (int t = 0; t < p.H; ++t) {
    const float2 u = __ldg(act + t);
    uint8_t fl = 0;
    float ax = __fdiv_rn(u.x, m), ay = __fdiv_rn(u.y, m);
    if (ax < -p.max_acc || ax > p.max_acc) fl |= kClampAx;
    if (ay < -p.max_acc || ay > p.max_acc) fl |= kClampAy;
    ax = fminf(fmaxf(ax, -p.max_acc), p.max_acc);
    ay = fminf(fmaxf(ay, -p.max_acc), p.max_acc);
    float f = 1.f;
    if (p.can_crash && p.grid != nullptr) {
      // X * (1 / cell) + c_offset, floor, clamp to the grid (obstacle_map.py:81-88)
      long long ix = (long long)floorf(__fadd_rn(__fmul_rn(x, p.inv_cell), p.off_x));
      long long iy = (long long)floorf(__fadd_rn(__fmul_rn(y, p.inv_cell), p.off_y));
      ix = min(max(ix, 0ll), (long long)p.gx - 1);
      iy = min(max(iy, 0ll), (long long)p.gy - 1);
      if (__ldg(p.grid + ix * p.gy + iy)) {
        fl |= kCrash;
        f = 0.f;
      }
    }
    // states + x_dot * dt * (1 - mask), in that order, no FMA contraction
    x = __fadd_rn(x, __fmul_rn(__fmul_rn(vx, p.dt), f));
    y = __fadd_rn(y, __fmul_rn(__fmul_rn(vy, p.dt), f));
    float nvx = __fadd_rn(vx, __fmul_rn(__fmul_rn(ax, p.dt), f));
    float nvy = __fadd_rn(vy, __fmul_rn(__fmul_rn(ay, p.dt), f));
    if (nvx < -p.max_speed || nvx > p.max_speed) fl |= kClampVx;
    if (nvy < -p.max_speed || nvy > p.max_speed) fl |= kClampVy;
    vx = fminf(fmaxf(nvx, -p.max_speed), p.max_speed);
    vy = fminf(fmaxf(nvy, -p.max_speed), p.max_speed);
    out[t + 1] = make_float4(x, y, vx, vy);
    flags[r * (int64_t)p.H + t] = fl;
  }
}

// grad_actions[r, t, :] for the cotangent grad_states [R, H+1, 4]
__global__ void __launch_bounds__(128)
particle_rollout_bwd_kernel(RolloutArgs p, const float* __restrict__ grad_states, const uint8_t* __restrict__ flags,
                            float* __restrict__ grad_actions) {
  const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= p.R) return;
  const float m = p.mass ? __ldg(p.mass + r) : p.mass_scalar;
  const float4* g = reinterpret_cast<const float4*>(grad_states + r * (int64_t)(p.H + 1) * 4);
  float2* ga = reinterpret_cast<float2*>(grad_actions + r * (int64_t)p.H * 2);
  float4 a = __ldg(g + p.H);  // adjoint of s_{t+1}
  for (int t = p.H - 1; t >= 0; --t) {
    const uint8_t fl = flags[r * (int64_t)p.H + t];
    const float k = (fl & kCrash) ? 0.f : p.dt;  // dt * (1 - mask)
    const float gvx = (fl & kClampVx) ? 0.f : a.z, gvy = (fl & kClampVy) ? 0.f : a.w;  // through the speed clamp
    ga[t] = make_float2((fl & kClampAx) ? 0.f : gvx * k / m, (fl & kClampAy) ? 0.f : gvy * k / m);
    const float4 gt = __ldg(g + t);
    a = make_float4(a.x + gt.x, a.y + gt.y, fmaf(a.x, k, gvx) + gt.z, fmaf(a.y, k, gvy) + gt.w);
  }
}

int check(const RolloutArgs& p, const char* what) {
  SIG_CHECK_ARG(p.init && p.R >= 0 && p.H >= 1, "%s: bad shape R=%lld H=%d", what, (long long)p.R, p.H);
  SIG_CHECK_ARG(p.dt > 0.f && (p.mass != nullptr || p.mass_scalar > 0.f), "%s: dt and mass must be > 0", what);
  SIG_CHECK_ARG(p.max_acc > 0.f && p.max_speed > 0.f, "%s: bounds must be > 0 (inf = unbounded)", what);
  SIG_CHECK_ARG(!p.can_crash || p.grid == nullptr || (p.gx >= 1 && p.gy >= 1 && p.inv_cell > 0.f),
                "%s: bad occupancy grid", what);
  return 0;
}

}  // namespace

static RolloutArgs make_rollout_args(const float* init, const float* mass, float mass_scalar, int64_t R, int64_t H,
                                     float dt, float max_acc, float max_speed, const uint8_t* grid, int64_t gx,
                                     int64_t gy, float inv_cell, float off_x, float off_y, int can_crash) {
  RolloutArgs p{};
  p.init = init, p.mass = mass, p.mass_scalar = mass_scalar, p.dt = dt, p.max_acc = max_acc, p.max_speed = max_speed;
  p.grid = grid, p.gx = (int)gx, p.gy = (int)gy, p.inv_cell = inv_cell, p.off_x = off_x, p.off_y = off_y;
  p.can_crash = can_crash, p.R = R, p.H = (int)H;
  return p;
}

int launch_particle_rollout(const float* init, const float* actions, const float* mass, float mass_scalar, int64_t R,
                            int64_t H, float dt, float max_acc, float max_speed, const uint8_t* grid, int64_t gx,
                            int64_t gy, float inv_cell, float off_x, float off_y, int can_crash, float* states,
                            uint8_t* flags, cudaStream_t st) {
  const RolloutArgs p = make_rollout_args(init, mass, mass_scalar, R, H, dt, max_acc, max_speed, grid, gx, gy,
                                          inv_cell, off_x, off_y, can_crash);
  SIG_PROPAGATE(check(p, "particle_rollout"));
  SIG_CHECK_ARG(R == 0 || (actions && states && flags), "particle_rollout: NULL pointer");
  SIG_CHECK_ARG(H < (1 << 24) && gx < (1 << 30) && gy < (1 << 30), "particle_rollout: shape too large");
  if (R == 0) return 0;
  particle_rollout_kernel<<<(unsigned)ceil_div(R, 128), 128, 0, st>>>(p, actions, states, flags);
  SIG_LAUNCH_OK("particle_rollout_kernel");
  return 0;
}

int launch_particle_rollout_bwd(const float* mass, float mass_scalar, int64_t R, int64_t H, float dt,
                                const float* grad_states, const uint8_t* flags, float* grad_actions,
                                cudaStream_t st) {
  float one = 1.f;
  RolloutArgs p = make_rollout_args(&one, mass, mass_scalar, R, H, dt, 1.f, 1.f, nullptr, 0, 0, 0.f, 0.f, 0.f, 0);
  SIG_PROPAGATE(check(p, "particle_rollout_bwd"));
  SIG_CHECK_ARG(R == 0 || (grad_states && flags && grad_actions), "particle_rollout_bwd: NULL pointer");
  if (R == 0) return 0;
  particle_rollout_bwd_kernel<<<(unsigned)ceil_div(R, 128), 128, 0, st>>>(p, grad_states, flags, grad_actions);
  SIG_LAUNCH_OK("particle_rollout_bwd_kernel");
  return 0;
}

}  // namespace sigsvgd
