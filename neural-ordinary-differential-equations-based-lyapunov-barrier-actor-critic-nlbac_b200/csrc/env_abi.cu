// Vectorised environment stepping: E independent copies of the reference's UnicycleEnv
// (U envs/unicycle_env.py) advanced in ONE launch, one thread per environment, float64 like the reference's numpy
// state.  HBM-streaming: 24 B of state + 16 B of action in, 7 + 7 doubles out per environment.
//   nlbac_env_cars_reset / nlbac_env_cars_step   SimulatedCarsEnv.reset / step / _get_obs
//                              (C envs/simulated_cars_env.py:42-146, 158-177, 197-213)
//   nlbac_env_pvtol_reset / nlbac_env_pvtol_step PvtolEnv.reset / step / _step / get_obs / obs_compass
//                              (P envs/pvtol_env.py:66-216, 226-246, 248-277, 361-408)
//   nlbac_env_unicycle_reset   UnicycleEnv.reset            (U envs/unicycle_env.py:166-184)
//   nlbac_env_unicycle_step    UnicycleEnv.step / _step     (U envs/unicycle_env.py:52-155), get_obs (:263-279),
//                              obs_compass (:302-318), goal_met (:157-164)
#include "../../include/nlbac.h"
#include "common.cuh"

namespace nlbac {

struct UniConst {
  static constexpr double dt = 0.02, l_p = 0.03, des_v = 2.5, goal_x = 2.5, goal_y = 2.5, goal_size = 0.3;
  static constexpr double reward_goal = 500.0, hazards_radius = 0.5, safety_cost_coef = 1.0;
  static constexpr int max_episode_steps = 1200, n_hazards = 7;
};
__constant__ double c_uni_hazards[7][2] = {{0.0, 0.0},  {0.0, 1.5},  {0.0, -1.5}, {-1.5, 1.5},
                                          {-1.5, -1.5}, {1.5, -1.5}, {1.5, 1.5}};   // (U :30) * 1.5

// get_obs (U :263-279) with obs_compass (U :302-318): vec = (goal - p) R(theta), normalised by |vec| + 1e-3
__device__ __forceinline__ void uni_obs(double x, double y, double th, double* o) {
  const double c = cos(th), s = sin(th);
  const double rx = UniConst::goal_x - x, ry = UniConst::goal_y - y;
  const double dist = sqrt(rx * rx + ry * ry);
  double v0 = rx * c + ry * s, v1 = rx * (-s) + ry * c;
  const double nrm = sqrt(v0 * v0 + v1 * v1) + 0.001;
  v0 /= nrm; v1 /= nrm;
  o[0] = x; o[1] = y; o[2] = c; o[3] = s; o[4] = v0; o[5] = v1; o[6] = exp(-dist);
}

__global__ void env_unicycle_reset_kernel(double* __restrict__ state, double* __restrict__ last_goal_dist,
                                          int* __restrict__ episode_step, const unsigned char* __restrict__ mask,
                                          long long E, double* __restrict__ obs) {
  for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < E; e += (long long)gridDim.x * blockDim.x) {
    if (mask != nullptr && mask[e] == 0) continue;
    state[3 * e] = -2.5; state[3 * e + 1] = -2.5; state[3 * e + 2] = 0.0;
    const double dx = UniConst::goal_x - (-2.47), dy = UniConst::goal_y - (-2.5);   // next_center after reset (U :179)
    last_goal_dist[e] = sqrt(dx * dx + dy * dy);
    episode_step[e] = 0;
    if (obs != nullptr) uni_obs(-2.5, -2.5, 0.0, obs + 7 * e);
  }
}

__global__ void env_unicycle_step_kernel(double* __restrict__ state, double* __restrict__ last_goal_dist,
                                         int* __restrict__ episode_step, const double* __restrict__ action, long long E,
                                         double* __restrict__ obs, double* __restrict__ reward,
                                         double* __restrict__ constraint, double* __restrict__ center,
                                         double* __restrict__ next_center, unsigned char* __restrict__ done,
                                         unsigned char* __restrict__ goal_met, int* __restrict__ n_violation,
                                         double* __restrict__ safety_cost) {
  using K = UniConst;
  for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < E; e += (long long)gridDim.x * blockDim.x) {
    double x = state[3 * e], y = state[3 * e + 1], th = state[3 * e + 2];
    const double a0 = action[2 * e], a1 = action[2 * e + 1];
    // center before the step (U :96-100)
    double c = cos(th), s = sin(th);
    center[2 * e] = x + K::l_p * c;
    center[2 * e + 1] = y + K::l_p * s;
    // state += dt (f + g a), f = 0, g = [[cos, 0], [sin, 0], [0, 1]] (U :104, :296-299)
    x += K::dt * (c * a0);
    y += K::dt * (s * a0);
    th += K::dt * a1;
    // state -= (dt * 0.1 * g(state)) @ [cos(theta), 0] (U :105; `*` and `@` associate left to right)
    c = cos(th); s = sin(th);
    const double k = K::dt * 0.1;
    x -= (k * c) * c;
    y -= (k * s) * c;
    state[3 * e] = x; state[3 * e + 1] = y; state[3 * e + 2] = th;
    const double ncx = x + K::l_p * c, ncy = y + K::l_p * s;
    next_center[2 * e] = ncx; next_center[2 * e + 1] = ncy;
    const int step = episode_step[e] + 1;
    episode_step[e] = step;
    const double gx = K::goal_x - ncx, gy = K::goal_y - ncy;
    const double dist_goal = sqrt(gx * gx + gy * gy);
    constraint[e] = dist_goal;
    const double dv = a0 - K::des_v;
    double r = -(dv * dv) * 0.1 + (last_goal_dist[e] - dist_goal) * 30.0;
    last_goal_dist[e] = dist_goal;
    const bool met = dist_goal <= K::goal_size;     // goal_met (U :157-164)
    if (met) r += K::reward_goal;
    reward[e] = r;
    done[e] = (met || step >= K::max_episode_steps) ? 1 : 0;
    goal_met[e] = met ? 1 : 0;
    int nv = 0;
    double cost = 0.0;
#pragma unroll
    for (int i = 0; i < K::n_hazards; ++i) {          // U :131-153
      const double hx = ncx - c_uni_hazards[i][0], hy = ncy - c_uni_hazards[i][1];
      const double d2 = hx * hx + hy * hy;
      if (d2 < K::hazards_radius * K::hazards_radius) {
        ++nv;
        cost += ((K::hazards_radius - sqrt(d2)) / K::hazards_radius) * K::safety_cost_coef;
      }
    }
    n_violation[e] = nv;
    safety_cost[e] = cost;
    uni_obs(x, y, th, obs + 7 * e);
  }
}

// ------------------------------------------------------------------------------------------------
// SimulatedCarsEnv (C envs/simulated_cars_env.py): state [x_1 v_1 ... x_5 v_5], car 4 controlled
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void cars_obs(const double* st, double* o) {       // _get_obs (C :197-213)
#pragma unroll
  for (int i = 0; i < 5; ++i) { o[2 * i] = st[2 * i] / 100.0; o[2 * i + 1] = st[2 * i + 1] / 30.0; }
}

__global__ void env_cars_reset_kernel(double* __restrict__ state, double* __restrict__ t, int* __restrict__ episode_step,
                                      const double* __restrict__ vel_noise, const unsigned char* __restrict__ mask,
                                      long long E, double* __restrict__ obs) {
  for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < E; e += (long long)gridDim.x * blockDim.x) {
    if (mask != nullptr && mask[e] == 0) continue;
    const double p0[5] = {42.0, 34.0, 26.0, 18.0, 10.0};
    const double v = 3.0 + vel_noise[e];            // ONE draw per reset, shared by the cars (C :170)
    double st[10];
#pragma unroll
    for (int i = 0; i < 5; ++i) { st[2 * i] = p0[i]; st[2 * i + 1] = v; }
    st[7] = 3.0;
#pragma unroll
    for (int i = 0; i < 10; ++i) state[10 * e + i] = st[i];
    t[e] = 0.0;
    episode_step[e] = 0;
    if (obs != nullptr) cars_obs(st, obs + 10 * e);
  }
}

__global__ void env_cars_step_kernel(double* __restrict__ state, double* __restrict__ tt, int* __restrict__ episode_step,
                                     const double* __restrict__ action, long long E, double* __restrict__ obs,
                                     double* __restrict__ reward, double* __restrict__ constraint,
                                     double* __restrict__ prev_pos, double* __restrict__ next_pos,
                                     unsigned char* __restrict__ done, unsigned char* __restrict__ reached,
                                     int* __restrict__ n_violation, double* __restrict__ safety_cost) {
  constexpr double dt = 0.02, kp = 4.0, k_brake = 20.0, should_keep = 9.5, thre = 0.5, reward_goal = 2.0;
  constexpr int max_steps = 300;
  for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < E; e += (long long)gridDim.x * blockDim.x) {
    double st[10];
#pragma unroll
    for (int i = 0; i < 10; ++i) st[i] = state[10 * e + i];
    const double t = tt[e], a = action[e];
    double acc[5];
#pragma unroll
    for (int i = 0; i < 5; ++i) acc[i] = kp * ((i == 0 ? 3.0 - 4.0 * sin(t) : 3.0) - st[2 * i + 1]);    // C :73-77
    const double d01 = st[0] - st[2], d12 = st[2] - st[4], d24 = st[4] - st[8];
    acc[1] += -k_brake * d01 * (d01 < 6.5 ? 1.0 : 0.0);
    acc[2] += -k_brake * d12 * (d12 < 6.5 ? 1.0 : 0.0);
    acc[3] = 0.0;
    acc[4] += -k_brake * d24 * (d24 < 13.0 ? 1.0 : 0.0);
#pragma unroll
    for (int i = 0; i < 5; ++i) acc[i] *= 1.1;                                                           // C :87
#pragma unroll
    for (int i = 0; i < 4; ++i) prev_pos[4 * e + i] = st[4 + i];
    double nst[10];
#pragma unroll
    for (int i = 0; i < 5; ++i) {
      nst[2 * i] = st[2 * i] + dt * st[2 * i + 1];
      nst[2 * i + 1] = st[2 * i + 1] + dt * (i == 3 ? a : acc[i]);        // f_x[7] = 0, g_x[7] = 1 (C :95-100)
    }
#pragma unroll
    for (int i = 0; i < 10; ++i) state[10 * e + i] = nst[i];
    tt[e] = t + dt;
    const int step = episode_step[e] + 1;
    episode_step[e] = step;
#pragma unroll
    for (int i = 0; i < 4; ++i) next_pos[4 * e + i] = nst[4 + i];
    const double d34 = nst[4] - nst[6], d45 = nst[6] - nst[8];
    double r = -0.5 * fabs(a * a) / (double)max_steps;                     // _get_reward (C :148-155)
    const bool in = fabs(d34 - should_keep) < thre;
    if (in) r = r + reward_goal;
    reward[e] = r;
    reached[e] = in ? 1 : 0;
    done[e] = step >= max_steps ? 1 : 0;
    int nv = 0;
    double cost = 0.0;
    if (d34 < 2.5) { ++nv; cost = cost + fabs(d34 - 2.5); }
    if (d45 < 2.5) { ++nv; cost = cost + fabs(d45 - 2.5); }
    n_violation[e] = nv;
    safety_cost[e] = cost;
    constraint[e] = fabs(d34 - should_keep);
    cars_obs(nst, obs + 10 * e);
  }
}

// ------------------------------------------------------------------------------------------------
// PvtolEnv (P envs/pvtol_env.py): state [x y theta vx vy thrust operator_x]
// ------------------------------------------------------------------------------------------------
__constant__ double c_pvtol_hazards[5][2] = {{-2.5, -2.5}, {-2.5, 2.5}, {0.0, -3.5}, {0.0, 3.5}, {-4.5, 0.0}};

__device__ __forceinline__ void pvtol_obs(const double* st, double* o) {      // get_obs (P :361-377), obs_compass (:379-405)
  const double c = cos(st[2]), s = sin(st[2]);
  const double rx = 4.5 - st[0], ry = 4.5 - st[1];
  const double dist = sqrt(rx * rx + ry * ry);
  double v0 = rx * c + ry * s, v1 = rx * (-s) + ry * c;
  const double nrm = sqrt(v0 * v0 + v1 * v1) + 0.001;
  o[0] = st[0]; o[1] = st[1]; o[2] = c; o[3] = s; o[4] = st[3]; o[5] = st[4]; o[6] = st[5]; o[7] = st[6];
  o[8] = v0 / nrm; o[9] = v1 / nrm; o[10] = exp(-dist);
}

__global__ void env_pvtol_reset_kernel(double* __restrict__ state, int* __restrict__ episode_step,
                                       const unsigned char* __restrict__ mask, long long E, double* __restrict__ obs) {
  for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < E; e += (long long)gridDim.x * blockDim.x) {
    if (mask != nullptr && mask[e] == 0) continue;
    const double st[7] = {-4.5, -4.5, 0.0, 0.0, 0.0, 1.0, -4.5};              // P :236-239
#pragma unroll
    for (int i = 0; i < 7; ++i) state[7 * e + i] = st[i];
    episode_step[e] = 0;
    if (obs != nullptr) pvtol_obs(st, obs + 11 * e);
  }
}

__global__ void env_pvtol_step_kernel(double* __restrict__ state, int* __restrict__ episode_step,
                                      const double* __restrict__ action, long long E, double* __restrict__ obs,
                                      double* __restrict__ reward, double* __restrict__ constraint,
                                      double* __restrict__ lya_pre, unsigned char* __restrict__ done,
                                      unsigned char* __restrict__ goal_met, int* __restrict__ n_violation,
                                      double* __restrict__ safety_cost) {
  constexpr double dt = 0.02, follow = 0.7, goal_size = 3.5, reward_goal = 1500.0, haz_r = 0.25, op_dist = 1.0;
  constexpr double y_max = 100.0, y_min = -100.0;
  constexpr int max_steps = 2000;
  for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < E; e += (long long)gridDim.x * blockDim.x) {
    double st[7];
#pragma unroll
    for (int i = 0; i < 7; ++i) st[i] = state[7 * e + i];
    pvtol_obs(st, lya_pre + 11 * e);                                            // P :96
    const double a0 = action[2 * e], a1 = action[2 * e + 1];
    // x' = f(x) + g(x) u on the first six entries (P :99-101, :250-274): g u = [0, 0, u1, 0, 0, u0]
    const double c = cos(st[2]), s = sin(st[2]);
    const double f3 = -s * st[5], f4 = c * st[5] - 1.0;
    const double n0 = st[0] + dt * st[3], n1 = st[1] + dt * st[4], n2 = st[2] + dt * a1;
    const double n3 = st[3] + dt * f3, n4 = st[4] + dt * f4, n5 = st[5] + dt * a0;
    st[0] = n0; st[1] = n1; st[2] = n2; st[3] = n3; st[4] = n4; st[5] = n5;
    st[6] = st[6] + follow * (st[0] - st[6]);                                    // safety operator (P :112-114)
#pragma unroll
    for (int i = 0; i < 7; ++i) state[7 * e + i] = st[i];
    const int step = episode_step[e] + 1;
    episode_step[e] = step;
    const double gx = 4.5 - st[0], gy = 4.5 - st[1];
    const double dist = sqrt(gx * gx + gy * gy);
    constraint[e] = dist;
    double r = -1e-3 * dist;
    const bool met = dist <= goal_size;
    if (met) r += reward_goal;
    reward[e] = r;
    done[e] = (met || step >= max_steps) ? 1 : 0;
    goal_met[e] = met ? 1 : 0;
    int nv[4] = {0, 0, 0, 0};
    double cost[4] = {0.0, 0.0, 0.0, 0.0};
#pragma unroll
    for (int i = 0; i < 5; ++i) {                                                // P :135-153
      const double hx = st[0] - c_pvtol_hazards[i][0], hy = st[1] - c_pvtol_hazards[i][1];
      const double d2 = hx * hx + hy * hy;
      if (d2 < haz_r * haz_r) { ++nv[0]; cost[0] += ((haz_r - sqrt(d2)) / haz_r) * 1.0; }
    }
    const double dop = st[0] - st[6];
    if (fabs(dop) >= op_dist) {                                                  // P :155-180
      nv[1] = 1;
      cost[1] = (dop <= -op_dist) ? (-(dop) - op_dist) * 1.0 : (dop - op_dist) * 1.0;
    }
    if (st[1] > y_max) { nv[2] = 1; cost[2] = (st[1] - y_max) * 1.0; }         // P :182-193
    if (st[1] < y_min) { nv[3] = 1; cost[3] = (y_min - st[1]) * 1.0; }         // P :195-206
#pragma unroll
    for (int i = 0; i < 4; ++i) { n_violation[4 * e + i] = nv[i]; safety_cost[4 * e + i] = cost[i]; }
    pvtol_obs(st, obs + 11 * e);
  }
}

static int env_grid(long long n, int threads) {
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  long long g = (n + threads - 1) / threads;
  const long long cap = (long long)sms * 8;
  return (int)(g < 1 ? 1 : (g > cap ? cap : g));
}

}  // namespace nlbac

using namespace nlbac;

extern "C" {

int nlbac_env_unicycle_reset(double* state, double* last_goal_dist, int32_t* episode_step, const uint8_t* mask,
                             int64_t E, double* obs, void* stream) {
  ::nlbac::NvtxRange nvtx_range_("nlbac:env_unicycle_reset");
  NLBAC_REQUIRE(state && last_goal_dist && episode_step && E >= 0, "bad arguments");
  if (E == 0) return 0;
  env_unicycle_reset_kernel<<<env_grid(E, 256), 256, 0, (cudaStream_t)stream>>>(state, last_goal_dist, episode_step,
                                                                               mask, E, obs);
  NLBAC_CUDA(cudaGetLastError());
  return 0;
}

int nlbac_env_unicycle_step(double* state, double* last_goal_dist, int32_t* episode_step, const double* action,
                            int64_t E, double* obs, double* reward, double* constraint, double* center,
                            double* next_center, uint8_t* done, uint8_t* goal_met, int32_t* num_safety_violation,
                            double* safety_cost, void* stream) {
  ::nlbac::NvtxRange nvtx_range_("nlbac:env_unicycle_step");
  NLBAC_REQUIRE(state && last_goal_dist && episode_step && action && obs && reward && constraint && center &&
                    next_center && done && goal_met && num_safety_violation && safety_cost && E >= 0,
                "bad arguments");
  if (E == 0) return 0;
  env_unicycle_step_kernel<<<env_grid(E, 256), 256, 0, (cudaStream_t)stream>>>(
      state, last_goal_dist, episode_step, action, E, obs, reward, constraint, center, next_center, done, goal_met,
      num_safety_violation, safety_cost);
  NLBAC_CUDA(cudaGetLastError());
  return 0;
}

int nlbac_env_cars_reset(double* state, double* t, int32_t* episode_step, const double* vel_noise, const uint8_t* mask,
                         int64_t E, double* obs, void* stream) {
  ::nlbac::NvtxRange nvtx_range_("nlbac:env_cars_reset");
  NLBAC_REQUIRE(state && t && episode_step && vel_noise && E >= 0, "bad arguments");
  if (E == 0) return 0;
  env_cars_reset_kernel<<<env_grid(E, 256), 256, 0, (cudaStream_t)stream>>>(state, t, episode_step, vel_noise, mask, E,
                                                                           obs);
  NLBAC_CUDA(cudaGetLastError());
  return 0;
}

int nlbac_env_cars_step(double* state, double* t, int32_t* episode_step, const double* action, int64_t E, double* obs,
                        double* reward, double* constraint, double* previous_positions, double* next_positions,
                        uint8_t* done, uint8_t* reached, int32_t* num_safety_violation, double* safety_cost,
                        void* stream) {
  ::nlbac::NvtxRange nvtx_range_("nlbac:env_cars_step");
  NLBAC_REQUIRE(state && t && episode_step && action && obs && reward && constraint && previous_positions &&
                    next_positions && done && reached && num_safety_violation && safety_cost && E >= 0,
                "bad arguments");
  if (E == 0) return 0;
  env_cars_step_kernel<<<env_grid(E, 256), 256, 0, (cudaStream_t)stream>>>(
      state, t, episode_step, action, E, obs, reward, constraint, previous_positions, next_positions, done, reached,
      num_safety_violation, safety_cost);
  NLBAC_CUDA(cudaGetLastError());
  return 0;
}

int nlbac_env_pvtol_reset(double* state, int32_t* episode_step, const uint8_t* mask, int64_t E, double* obs,
                          void* stream) {
  ::nlbac::NvtxRange nvtx_range_("nlbac:env_pvtol_reset");
  NLBAC_REQUIRE(state && episode_step && E >= 0, "bad arguments");
  if (E == 0) return 0;
  env_pvtol_reset_kernel<<<env_grid(E, 256), 256, 0, (cudaStream_t)stream>>>(state, episode_step, mask, E, obs);
  NLBAC_CUDA(cudaGetLastError());
  return 0;
}

int nlbac_env_pvtol_step(double* state, int32_t* episode_step, const double* action, int64_t E, double* obs,
                         double* reward, double* constraint, double* lya_pre_term, uint8_t* done, uint8_t* goal_met,
                         int32_t* num_safety_violation, double* safety_cost, void* stream) {
  ::nlbac::NvtxRange nvtx_range_("nlbac:env_pvtol_step");
  NLBAC_REQUIRE(state && episode_step && action && obs && reward && constraint && lya_pre_term && done && goal_met &&
                    num_safety_violation && safety_cost && E >= 0,
                "bad arguments");
  if (E == 0) return 0;
  env_pvtol_step_kernel<<<env_grid(E, 256), 256, 0, (cudaStream_t)stream>>>(
      state, episode_step, action, E, obs, reward, constraint, lya_pre_term, done, goal_met, num_safety_violation,
      safety_cost);
  NLBAC_CUDA(cudaGetLastError());
  return 0;
}

}  // extern "C"
