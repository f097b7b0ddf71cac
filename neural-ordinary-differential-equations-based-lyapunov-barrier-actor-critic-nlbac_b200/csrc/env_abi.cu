// Vectorised environment stepping: E independent copies of the reference's UnicycleEnv
// (U envs/unicycle_env.py) advanced in ONE launch, one thread per environment, float64 like the reference's numpy
// state.  HBM-streaming: 24 B of state + 16 B of action in, 7 + 7 doubles out per environment.
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

}  // extern "C"
