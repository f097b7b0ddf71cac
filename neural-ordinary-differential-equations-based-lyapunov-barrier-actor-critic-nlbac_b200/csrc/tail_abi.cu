// C ABI of the element-wise / reduction tail of the update step (include/nlbac.h):
//   nlbac_constraints_*  K5  CBF/CLF constraint rows -> filter -> batch sums   (U sac_cbf_clf.py:472-504,
//                            C :474-523, BU :424-460)
//   nlbac_lookahead_*    K6  p(x) = [x, y] + l_p [cos th, sin th]              (U sac_cbf_clf.py:429-437,455-469)
//   nlbac_get_state / nlbac_get_obs_*  K6  DynamicsModel.get_state / get_obs   (U dynamics.py:27-69, C :25-96,
//                            P :25-153, BU :27-198)
//   nlbac_adam_step / nlbac_soft_update  torch.optim.Adam.step (U model.py:257) / utils.soft_update (U utils.py:75-79)
// All are HBM-streaming kernels: one pass over the rows, coalesced, grid sized to the SM count.
#include "../../include/nlbac.h"
#include "common.cuh"

namespace nlbac {

static int tail_grid(long long n, int threads) {
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  long long g = (n + threads - 1) / threads;
  const long long cap = (long long)sms * 8;
  return (int)(g < 1 ? 1 : (g > cap ? cap : g));
}

// ------------------------------------------------------------------------------------------------
// constraint rows.  MAXM rows per sample, kept in registers; block reduction; one atomicAdd per
// (CTA, row).  For B <= 256 (the real minibatches) the grid is one CTA and the sums are deterministic.
// ------------------------------------------------------------------------------------------------
constexpr int MAXM = 16;
constexpr int CNT = 256;

template <bool BWD>
__device__ __forceinline__ int constraint_rows(const nlbac_constraint_desc& d, const float* a, const float* b,
                                               const float* c, const float* e, const float* V, const float* Vn,
                                               long long i, float (&row)[MAXM], const float* g /*[m] scaled*/,
                                               float* gb, float* gc, float* ge, float* gVn) {
  const float v = V[i], vn = Vn[i];
  if (d.kind == NLBAC_CONSTRAINT_PVTOL) {
    // a = s_t (data), b, c, e = s_{t+1..3} [B,7]: column 0 = x, 1 = y, 6 = safety operator's x  (P sac_cbf_clf.py:444-697)
    const float* sp[4] = {a + 7 * i, b + 7 * i, c + 7 * i, e + 7 * i};
    float X[4], Y[4], O[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) { X[j] = sp[j][0]; Y[j] = sp[j][1]; O[j] = sp[j][6]; }
    const float gm = d.gamma_b, am = d.gamma_b - 1.f;
    // third-order discrete CBF expression, evaluated in the reference's order (P :581-587)
    auto cbf3 = [gm](float h0, float h1, float h2, float h3) {
      const float t1 = h3 - h2 + gm * h2, t2 = h2 - h1 + gm * h1, t3 = h1 - h0 + gm * h0;
      const float y = t2 - t3 + gm * t3;
      return -(t1 - t2 + gm * t2 - y + gm * y);
    };
    // d row / d h_j, j = 1..3:  row = -(h3 + 3a h2 + 3a^2 h1 + a^3 h0), a = gamma_b - 1
    const float coef[4] = {0.f, -3.f * am * am, -3.f * am, -1.f};
    float gx[4] = {0.f, 0.f, 0.f, 0.f}, gy[4] = {0.f, 0.f, 0.f, 0.f}, go[4] = {0.f, 0.f, 0.f, 0.f};
    const float rc2 = d.collision_radius * d.collision_radius;
    int k = 0;
    for (int h = 0; h < d.n_hazards; ++h, ++k) {
      const float cx = d.hazards[2 * h], cy = d.hazards[2 * h + 1];
      float hv[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) hv[j] = 0.5f * ((X[j] - cx) * (X[j] - cx) + (Y[j] - cy) * (Y[j] - cy) - rc2);
      row[k] = cbf3(hv[0], hv[1], hv[2], hv[3]);
      if (BWD && row[k] > 0.f) {
#pragma unroll
        for (int j = 1; j < 4; ++j) {
          const float s = g[k] * coef[j];
          gx[j] = fmaf(s, X[j] - cx, gx[j]);
          gy[j] = fmaf(s, Y[j] - cy, gy[j]);
        }
      }
    }
    const float od = d.operator_dist;
    row[k] = cbf3(X[0] - O[0] + od, X[1] - O[1] + od, X[2] - O[2] + od, X[3] - O[3] + od);
    if (BWD && row[k] > 0.f) {
#pragma unroll
      for (int j = 1; j < 4; ++j) { gx[j] += g[k] * coef[j]; go[j] -= g[k] * coef[j]; }
    }
    ++k;
    row[k] = cbf3(-X[0] + O[0] + od, -X[1] + O[1] + od, -X[2] + O[2] + od, -X[3] + O[3] + od);
    if (BWD && row[k] > 0.f) {
#pragma unroll
      for (int j = 1; j < 4; ++j) { gx[j] -= g[k] * coef[j]; go[j] += g[k] * coef[j]; }
    }
    ++k;
    row[k] = cbf3(-Y[0] + d.y_max - d.delta_y, -Y[1] + d.y_max - d.delta_y, -Y[2] + d.y_max - d.delta_y,
                  -Y[3] + d.y_max - d.delta_y);
    if (BWD && row[k] > 0.f) {
#pragma unroll
      for (int j = 1; j < 4; ++j) gy[j] -= g[k] * coef[j];
    }
    ++k;
    row[k] = cbf3(Y[0] - d.y_min - d.delta_y, Y[1] - d.y_min - d.delta_y, Y[2] - d.y_min - d.delta_y,
                  Y[3] - d.y_min - d.delta_y);
    if (BWD && row[k] > 0.f) {
#pragma unroll
      for (int j = 1; j < 4; ++j) gy[j] += g[k] * coef[j];
    }
    ++k;
    row[k] = (vn - v) * d.lya_inv_dt + d.gamma_l * v;
    if (BWD) {
      float* gp[4] = {nullptr, gb + 7 * i, gc + 7 * i, ge + 7 * i};
#pragma unroll
      for (int j = 1; j < 4; ++j) {
        gp[j][0] = gx[j]; gp[j][1] = gy[j]; gp[j][2] = 0.f; gp[j][3] = 0.f; gp[j][4] = 0.f; gp[j][5] = 0.f;
        gp[j][6] = go[j];
      }
      gVn[i] = row[k] > 0.f ? g[k] * d.lya_inv_dt : 0.f;
    }
    return k + 1;
  }
  if (d.kind == NLBAC_CONSTRAINT_UNICYCLE) {
    // a = p(x_t) [B,2], b = p(x_{t+1}) [B,2]
    const float px = a[2 * i], py = a[2 * i + 1], qx = b[2 * i], qy = b[2 * i + 1];
    const float rc2 = d.collision_radius * d.collision_radius;
    float gqx = 0.f, gqy = 0.f;
    for (int h = 0; h < d.n_hazards; ++h) {
      const float cx = d.hazards[2 * h], cy = d.hazards[2 * h + 1];
      const float hs = 0.5f * ((px - cx) * (px - cx) + (py - cy) * (py - cy) - rc2);
      const float hn = 0.5f * ((qx - cx) * (qx - cx) + (qy - cy) * (qy - cy) - rc2);
      row[h] = -((hn - hs) / d.dt) - d.gamma_b * hs;
      if (BWD && row[h] > 0.f) {   // d row / d hn = -1/dt ; d hn / d q = (q - c)
        const float s = -g[h] / d.dt;
        gqx = fmaf(s, qx - cx, gqx);
        gqy = fmaf(s, qy - cy, gqy);
      }
    }
    const int m = d.n_hazards;
    row[m] = ((vn - v) * d.lya_inv_dt) + d.gamma_l * v;
    if (BWD) {
      gb[2 * i] = gqx; gb[2 * i + 1] = gqy;
      gVn[i] = row[m] > 0.f ? g[m] * d.lya_inv_dt : 0.f;
    }
    return m + 1;
  }
  if (d.kind == NLBAC_CONSTRAINT_CARS) {
    // a = x_t [B,10], b = x_{t+1} [B,10], c = x_{t+2} [B,10];  h23 = x3 - x4 - r, h34 = x4 - x5 - r (positions at 4,6,8)
    const float r = d.collision_radius, gb_ = d.gamma_b;
    const float* x0 = a + 10 * i; const float* x1 = b + 10 * i; const float* x2 = c + 10 * i;
    float g1[10], g2[10];
#pragma unroll
    for (int k = 0; k < 10; ++k) g1[k] = g2[k] = 0.f;
#pragma unroll
    for (int p = 0; p < 2; ++p) {
      const int ia = 4 + 2 * p, ib = 6 + 2 * p;
      const float h0 = x0[ia] - x0[ib] - r, h1 = x1[ia] - x1[ib] - r, h2 = x2[ia] - x2[ib] - r;
      const float l1 = h1 - h0 + gb_ * h0;
      const float l2 = h2 - h1 + gb_ * h1;
      row[p] = -(l2 - l1) - gb_ * l1;
      if (BWD && row[p] > 0.f) {
        // d row/d h2 = -1 ; d row/d h1 = -(-1 + gb) + (1 - gb) = 2 - 2 gb   [l2: -1+gb, l1: +1 then *(1-gb)]
        const float dh2 = -g[p];
        const float dh1 = g[p] * ((1.f - gb_) + (1.f - gb_));
        g2[ia] += dh2; g2[ib] -= dh2;
        g1[ia] += dh1; g1[ib] -= dh1;
      }
    }
    row[2] = (vn - v) * d.lya_inv_dt + d.gamma_l * v;
    if (BWD) {
#pragma unroll
      for (int k = 0; k < 10; ++k) { gb[10 * i + k] = g1[k]; gc[10 * i + k] = g2[k]; }
      gVn[i] = row[2] > 0.f ? g[2] * d.lya_inv_dt : 0.f;
    }
    return 3;
  }
  // NLBAC_CONSTRAINT_BARRIER: a = B(x_t,u_t) (detached) [B], b = B(x_{t+1},u_{t+1}) [B]
  const float bv = a[i], bn = b[i];
  row[0] = -(bn - bv) - d.gamma_b * bv;
  row[1] = (vn - v) * d.lya_inv_dt + d.gamma_l * v;
  if (BWD) {
    gb[i] = row[0] > 0.f ? -g[0] : 0.f;
    gVn[i] = row[1] > 0.f ? g[1] * d.lya_inv_dt : 0.f;
  }
  return 2;
}

__global__ void __launch_bounds__(CNT) constraints_fwd_kernel(nlbac_constraint_desc d, const float* a, const float* b,
                                                              const float* c, const float* e, const float* V,
                                                              const float* Vn, long long B, float* req) {
  __shared__ float red[CNT / 32][MAXM];
  float acc[MAXM];
#pragma unroll
  for (int k = 0; k < MAXM; ++k) acc[k] = 0.f;
  int m = 0;
  for (long long i = (long long)blockIdx.x * CNT + threadIdx.x; i < B; i += (long long)gridDim.x * CNT) {
    float row[MAXM];
    m = constraint_rows<false>(d, a, b, c, e, V, Vn, i, row, nullptr, nullptr, nullptr, nullptr, nullptr);
#pragma unroll
    for (int k = 0; k < MAXM; ++k)
      if (k < m && row[k] > 0.f) acc[k] += row[k];   // torch.where(matr > 0, matr, 0)
  }
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
#pragma unroll
  for (int k = 0; k < MAXM; ++k) {
    const float s = warp_sum(acc[k]);
    if (lane == 0) red[warp][k] = s;
  }
  __syncthreads();
  if (threadIdx.x < d.m) {
    float s = 0.f;
#pragma unroll
    for (int w = 0; w < CNT / 32; ++w) s += red[w][threadIdx.x];
    atomicAdd(req + threadIdx.x, s / d.batch_size);
  }
}

__global__ void __launch_bounds__(CNT) constraints_bwd_kernel(nlbac_constraint_desc d, const float* a, const float* b,
                                                              const float* c, const float* e, const float* V,
                                                              const float* Vn, long long B, const float* grad_req,
                                                              float* gb, float* gc, float* ge, float* gVn) {
  __shared__ float g[MAXM];
  if (threadIdx.x < MAXM) g[threadIdx.x] = threadIdx.x < d.m ? grad_req[threadIdx.x] / d.batch_size : 0.f;
  __syncthreads();
  for (long long i = (long long)blockIdx.x * CNT + threadIdx.x; i < B; i += (long long)gridDim.x * CNT) {
    float row[MAXM];
    constraint_rows<true>(d, a, b, c, e, V, Vn, i, row, g, gb, gc, ge, gVn);
  }
}

// ------------------------------------------------------------------------------------------------
// lookahead point and obs/state maps
// ------------------------------------------------------------------------------------------------
__global__ void lookahead_fwd_kernel(const float* state, int ld, float l_p, long long B, float* p) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < B; i += (long long)gridDim.x * blockDim.x) {
    const float th = state[i * ld + 2];
    float s, c;
    sincosf(th, &s, &c);
    p[2 * i] = fmaf(l_p, c, state[i * ld]);
    p[2 * i + 1] = fmaf(l_p, s, state[i * ld + 1]);
  }
}
__global__ void lookahead_bwd_kernel(const float* state, int ld, float l_p, long long B, const float* gp, float* gs) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < B; i += (long long)gridDim.x * blockDim.x) {
    const float th = state[i * ld + 2];
    float s, c;
    sincosf(th, &s, &c);
    const float gx = gp[2 * i], gy = gp[2 * i + 1];
    gs[i * ld] = gx;
    gs[i * ld + 1] = gy;
    gs[i * ld + 2] = l_p * (gy * c - gx * s);
    for (int k = 3; k < ld; ++k) gs[i * ld + k] = 0.f;
  }
}

// get_state: the reference does the arithmetic in float64 numpy and rounds back
__global__ void get_state_kernel(int kind, const float* obs, long long B, float* state, float* state_dyn) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < B; i += (long long)gridDim.x * blockDim.x) {
    if (kind == NLBAC_ENV_UNICYCLE) {
      const float* o = obs + 7 * i;
      state[3 * i] = o[0]; state[3 * i + 1] = o[1];
      state[3 * i + 2] = (float)atan2((double)o[3], (double)o[2]);
    } else if (kind == NLBAC_ENV_CARS) {
      for (int k = 0; k < 10; ++k) state[10 * i + k] = (float)((double)obs[10 * i + k] * ((k & 1) ? 30.0 : 100.0));
    } else {  // PVTOL: obs[11] -> state7 and its first 6
      const float* o = obs + 11 * i;
      const float th = (float)atan2((double)o[3], (double)o[2]);
      const float s7[7] = {o[0], o[1], th, o[4], o[5], o[6], o[7]};
      for (int k = 0; k < 7; ++k) state[7 * i + k] = s7[k];
      if (state_dyn)
        for (int k = 0; k < 6; ++k) state_dyn[6 * i + k] = s7[k];
    }
  }
}

struct ObsGeom { int n_s, n_o; float gx, gy; };
__device__ __forceinline__ ObsGeom obs_geom(int kind) {
  if (kind == NLBAC_ENV_UNICYCLE) return {3, 7, 2.5f, 2.5f};   // BU dynamics.py:106-107
  if (kind == NLBAC_ENV_CARS) return {10, 10, 0.f, 0.f};
  return {7, 11, 4.5f, 4.5f};                                    // P dynamics.py:115-116
}
// obs = [x, y, cos, sin, (pass-through), compass_x, compass_y, exp(-dist)]
template <bool BWD>
__global__ void get_obs_kernel(int kind, const float* state, long long B, float* obs, const float* gobs, float* gstate) {
  const ObsGeom G = obs_geom(kind);
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < B; i += (long long)gridDim.x * blockDim.x) {
    const float* s = state + (long long)G.n_s * i;
    if (kind == NLBAC_ENV_CARS) {
      for (int k = 0; k < 10; ++k) {
        const float div = (k & 1) ? 30.0f : 100.0f;
        if (!BWD) obs[10 * i + k] = s[k] / div;
        else gstate[10 * i + k] = gobs[10 * i + k] / div;
      }
      continue;
    }
    const int np = G.n_s - 3;   // pass-through columns after theta
    float st, ct;
    sincosf(s[2], &st, &ct);
    const float rx = G.gx - s[0], ry = G.gy - s[1];
    const float dist = sqrtf(rx * rx + ry * ry);
    const float vx = ct * rx + st * ry, vy = -st * rx + ct * ry;   // Rs = [[c, s], [-s, c]]
    const float nv = sqrtf(vx * vx + vy * vy);
    const float den = nv + 0.001f;
    const float e = expf(-dist);
    if (!BWD) {
      float* o = obs + (long long)G.n_o * i;
      o[0] = s[0]; o[1] = s[1]; o[2] = ct; o[3] = st;
      for (int k = 0; k < np; ++k) o[4 + k] = s[3 + k];
      o[4 + np] = vx / den; o[5 + np] = vy / den; o[6 + np] = e;
    } else {
      const float* g = gobs + (long long)G.n_o * i;
      float* gs = gstate + (long long)G.n_s * i;
      const float gcx = g[4 + np], gcy = g[5 + np], ge = g[6 + np];
      // compass = v / (|v| + 1e-3):  dL/dv = g/den - v (g.v) / (|v| den^2)
      const float gv_dot = gcx * vx + gcy * vy;
      const float k2 = nv > 0.f ? gv_dot / (nv * den * den) : 0.f;
      const float gvx = gcx / den - vx * k2, gvy = gcy / den - vy * k2;
      // v = R r:  dL/dr = R^T gv ;  dL/dtheta = gv . (dR/dtheta r) = gvx * vy - gvy * vx
      float grx = ct * gvx - st * gvy, gry = st * gvx + ct * gvy;
      // exp(-dist): d/dr = -e * r / dist
      if (dist > 0.f) { grx -= ge * e * rx / dist; gry -= ge * e * ry / dist; }
      gs[0] = g[0] - grx;   // r = goal - p
      gs[1] = g[1] - gry;
      gs[2] = -g[2] * st + g[3] * ct + gvx * vy - gvy * vx;
      for (int k = 0; k < np; ++k) gs[3 + k] = g[4 + k];
    }
  }
}

// ------------------------------------------------------------------------------------------------
// Adam (torch.optim.Adam defaults: no weight decay, no amsgrad) and Polyak averaging, flat tensors
// ------------------------------------------------------------------------------------------------
__global__ void adam_kernel(float* p, const float* g, float* m, float* v, long long n, float lr, float b1, float b2,
                            float eps, float bc1, float bc2_sqrt) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const float gi = g[i];
    const float mi = m[i] + (gi - m[i]) * (1.f - b1);      // exp_avg.lerp_(grad, 1 - beta1)
    const float vi = fmaf(v[i], b2, (1.f - b2) * gi * gi);  // exp_avg_sq.mul_(b2).addcmul_(g, g, 1 - b2)
    m[i] = mi; v[i] = vi;
    const float denom = sqrtf(vi) / bc2_sqrt + eps;
    p[i] = p[i] - (lr / bc1) * (mi / denom);
  }
}
__global__ void soft_update_kernel(float* target, const float* src, long long n, float tau) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    target[i] = target[i] * (1.0f - tau) + src[i] * tau;
}

// multi-tensor forms: one launch for every parameter tensor of an optimiser / of a target network
// (blockIdx.y = tensor; the pointer table travels in the kernel parameters)
struct AdamTable {
  float* p[NLBAC_MAX_TENSORS]; const float* g[NLBAC_MAX_TENSORS]; float* m[NLBAC_MAX_TENSORS];
  float* v[NLBAC_MAX_TENSORS]; long long n[NLBAC_MAX_TENSORS];
};
struct PolyakTable { float* t[NLBAC_MAX_TENSORS]; const float* s[NLBAC_MAX_TENSORS]; long long n[NLBAC_MAX_TENSORS]; };
__global__ void adam_multi_kernel(const __grid_constant__ AdamTable tb, float lr, float b1, float b2, float eps,
                                  float bc1, float bc2_sqrt) {
  const int k = blockIdx.y;
  float* p = tb.p[k]; const float* g = tb.g[k]; float* m = tb.m[k]; float* v = tb.v[k];
  const long long n = tb.n[k];
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const float gi = g[i];
    const float mi = m[i] + (gi - m[i]) * (1.f - b1);
    const float vi = fmaf(v[i], b2, (1.f - b2) * gi * gi);
    m[i] = mi; v[i] = vi;
    const float denom = sqrtf(vi) / bc2_sqrt + eps;
    p[i] = p[i] - (lr / bc1) * (mi / denom);
  }
}
// Adam with the step counter ON THE DEVICE (graph-capturable: nothing about the update depends on host state).
// step_ctl[0] = step counter, step_ctl[1] = ticket word (both zero-initialised by the caller).  With `bump` the launch
// works with step = step_ctl[0] + 1 and the LAST block to finish (ticket) writes it back: every other block has read the
// counter by then, so no separate increment launch is needed.
__global__ void adam_multi_dev_kernel(const __grid_constant__ AdamTable tb, int* step_ctl, int bump, float lr,
                                      float b1, float b2, float eps) {
  const int k = blockIdx.y;
  float* p = tb.p[k]; const float* g = tb.g[k]; float* m = tb.m[k]; float* v = tb.v[k];
  const long long n = tb.n[k];
  // the two bias corrections are the same for the whole launch: ONE thread per block forms them (two double-precision
  // pow() per THREAD were most of this kernel's 8-9 us at 70-140 k parameters)
  __shared__ float s_bc[2];
  __shared__ int s_step;
  const bool work = (long long)blockIdx.x * blockDim.x < n;   // small tensors of the group: most blocks have nothing to do
  if (threadIdx.x == 0) {
    const int step = *reinterpret_cast<volatile int*>(step_ctl) + (bump ? 1 : 0);
    s_step = step;
    if (work) {
      s_bc[0] = (float)(1.0 - pow((double)b1, (double)step));
      s_bc[1] = (float)sqrt(1.0 - pow((double)b2, (double)step));
    }
  }
  __syncthreads();
  if (work) {
  const float bc1 = s_bc[0], bc2_sqrt = s_bc[1];
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const float gi = g[i], m0 = m[i], v0 = v[i], p0 = p[i];   // all loads before the first store
    const float mi = m0 + (gi - m0) * (1.f - b1);
    const float vi = fmaf(v0, b2, (1.f - b2) * gi * gi);
    m[i] = mi; v[i] = vi;
    const float denom = sqrtf(vi) / bc2_sqrt + eps;
    p[i] = p0 - (lr / bc1) * (mi / denom);
  }
  }
  if (bump && threadIdx.x == 0) {
    const unsigned total = gridDim.x * gridDim.y;
    const unsigned t = atomicAdd(reinterpret_cast<unsigned*>(step_ctl + 1), 1u);
    if (t == total - 1) {            // every block has read step_ctl[0] (it took its ticket after doing so)
      step_ctl[0] = s_step;
      step_ctl[1] = 0;
    }
  }
}
__global__ void soft_update_multi_kernel(const __grid_constant__ PolyakTable tb, float tau) {
  const int k = blockIdx.y;
  float* t = tb.t[k]; const float* s = tb.s[k];
  const long long n = tb.n[k];
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    t[i] = t[i] * (1.0f - tau) + s[i] * tau;
}

// replay ring gather: out[i][:] = ring[idx[i]][:]  (one thread per float; a warp reads / writes whole rows)
__global__ void replay_gather_kernel(const float* __restrict__ ring, int width, const long long* __restrict__ idx,
                                     long long B, float* __restrict__ out) {
  const long long n = B * width;
  for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < n; e += (long long)gridDim.x * blockDim.x) {
    const long long i = e / width;
    const int c = (int)(e - i * width);
    out[e] = ring[idx[i] * width + c];
  }
}

// the same gather, every field written to its OWN contiguous buffer (the static input buffers of a captured update):
// saves the strided -> contiguous copy per field that splitting a [B, width] row matrix costs
struct GatherSplit {
  float* out[NLBAC_REPLAY_MAX_FIELDS];
  int off[NLBAC_REPLAY_MAX_FIELDS], w[NLBAC_REPLAY_MAX_FIELDS];
  int n_fields;
};
__global__ void replay_gather_split_kernel(const float* __restrict__ ring, int width, const long long* __restrict__ idx,
                                           long long B, const __grid_constant__ GatherSplit g) {
  const long long n = B * width;
  for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < n; e += (long long)gridDim.x * blockDim.x) {
    const long long i = e / width;
    const int c = (int)(e - i * width);
    int f = 0;
    while (f + 1 < g.n_fields && c >= g.off[f + 1]) ++f;
    if (g.out[f] != nullptr && c < g.off[f] + g.w[f]) g.out[f][i * g.w[f] + (c - g.off[f])] = ring[idx[i] * width + c];
  }
}

static int check_constraint(const nlbac_constraint_desc* d) {
  NLBAC_REQUIRE(d != nullptr, "null constraint descriptor");
  NLBAC_REQUIRE(d->kind >= NLBAC_CONSTRAINT_UNICYCLE && d->kind <= NLBAC_CONSTRAINT_PVTOL, "unknown constraint kind %d", d->kind);
  const int want = d->kind == NLBAC_CONSTRAINT_UNICYCLE ? d->n_hazards + 1
                   : (d->kind == NLBAC_CONSTRAINT_CARS ? 3 : (d->kind == NLBAC_CONSTRAINT_PVTOL ? d->n_hazards + 5 : 2));
  NLBAC_REQUIRE(d->m == want && d->m <= MAXM, "constraint rows m=%d, expected %d (<= %d)", d->m, want, MAXM);
  NLBAC_REQUIRE(d->n_hazards >= 0 && d->n_hazards <= NLBAC_MAX_HAZARDS, "n_hazards=%d", d->n_hazards);
  NLBAC_REQUIRE(d->batch_size > 0.f && d->dt > 0.f, "batch_size and dt must be positive");
  return 0;
}

// ------------------------------------------------------------------------------------------------
// Augmented Lagrangian of get_policy_loss_2 / backup_get_policy_loss_2 on device-resident state:
// the multiplier update, the rho schedule, the CBF/CLF ratio and the loss in ONE single-thread launch (m <= 32
// constraint rows) instead of ~25 scalar float64 torch ops; same arithmetic and the same float64 / float32 types as
// constraints.augmented_lagrangian_device (which follows U sac_cbf_clf.py:506-530, :611-640 statement by statement).
// ------------------------------------------------------------------------------------------------
struct LagArgs {
  const float* req;   // [m] required_matrix (sums over the batch)
  double* lambda;     // [m] multipliers (updated when *due > 0)
  double* rho;        // augmented term (advanced every call)
  const double* due;
  const long long* updates;   // device-resident update counter: due = (*updates % interval == 0); null -> *due
  float* loss;        // [1]
  float* coef;        // [2m]: lin_i, quad_i  (d loss / d req_i = lin_i + 2 quad_i c_i)
  double cost_limit, lambda_max, aug_ratio, ratio_floor;
  int interval;
  int m, use_ratio, has_floor, clf_last;   // clf_last: the last row is the CLF row (primary loss); 0 for the backup loss
};
__global__ void lagrangian_fwd_kernel(LagArgs a) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  const int m = a.m;
  const float cl = (float)a.cost_limit;
  double ratio = 1.0;
  if (a.clf_last && a.use_ratio) {
    double s = 0.0;
    for (int i = 0; i < m - 1; ++i) s += (double)(a.req[i] - cl);
    ratio = fabs(s / (double)(m - 1)) / fabs((double)(a.req[m - 1] - cl));
    if (a.has_floor) ratio = fmax(ratio, a.ratio_floor);
  }
  const double rho0 = *a.rho;
  const double due = a.updates != nullptr ? ((*a.updates % a.interval == 0) ? 1.0 : 0.0) : *a.due;
  const double rho = fmin(rho0 * a.aug_ratio, 200.0);
  *a.rho = rho;
  float tot = 0.f;
  for (int i = 0; i < m; ++i) {
    double lam = a.lambda[i];
    if (due > 0.0) {
      lam = fmin(fmax(lam + rho0 * (double)a.req[i], 0.01), a.lambda_max);
      a.lambda[i] = lam;
    }
    const double w = (a.clf_last && i == m - 1) ? ratio : 1.0;
    const float lin = (float)(lam * w), quad = (float)(rho / 2.0 * w * w);
    const float c = a.req[i] - cl;
    a.coef[i] = lin;
    a.coef[m + i] = quad;
    tot += lin * c + quad * c * c;
  }
  a.loss[0] = tot;
}
__global__ void lagrangian_bwd_kernel(const float* req, const float* coef, int m, float cl, const float* gloss,
                                      float* greq) {
  const int i = threadIdx.x;
  if (i >= m) return;
  const float c = req[i] - cl;
  greq[i] = gloss[0] * (coef[i] + 2.f * coef[m + i] * c);
}

// ------------------------------------------------------------------------------------------------
// Scalar glue of update_parameters as single launches: the TD target (U sac_cbf_clf.py:234-238),
// F.mse_loss with mean reduction (:244-246) and the SAC term of the actor loss (:270).  Each replaces 4-10 one-op
// torch launches per call inside the captured update; same float32 operation order as the reference's expressions
// (no FMA contraction across its separate ops).
// ------------------------------------------------------------------------------------------------
__global__ void td_target_kernel(const float* q1, const float* q2, const float* logp, const float* alpha,
                                 const float* reward, const float* mask, float gamma, long long B, float* out) {
  const long long b = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  float v = q1[b];
  if (q2 != nullptr) v = fminf(v, q2[b]);
  if (logp != nullptr) v = __fsub_rn(v, __fmul_rn(alpha[0], logp[b]));
  out[b] = __fadd_rn(reward[b], __fmul_rn(__fmul_rn(mask[b], gamma), v));
}

__device__ __forceinline__ float block_sum_256(float v, float* red) {
  v = warp_sum(v);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  float t = 0.f;
  if (threadIdx.x < 32) {
    t = threadIdx.x < (blockDim.x >> 5) ? red[threadIdx.x] : 0.f;
    t = warp_sum(t);
  }
  return t;   // valid in warp 0
}
__global__ void __launch_bounds__(256) mse_mean_fwd_kernel(const float* pred, const float* target, long long n,
                                                           float* loss) {
  __shared__ float red[8];
  float s = 0.f;
  for (long long i = threadIdx.x; i < n; i += 256) {
    const float d = __fsub_rn(pred[i], target[i]);
    s = __fadd_rn(s, __fmul_rn(d, d));
  }
  s = block_sum_256(s, red);
  if (threadIdx.x == 0) loss[0] = s / (float)n;
}
__global__ void mse_mean_bwd_kernel(const float* pred, const float* target, long long n, const float* gloss,
                                    float* gpred) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  gpred[i] = gloss[0] * (2.f / (float)n) * (pred[i] - target[i]);
}
// loss = mean(alpha * logp - min(q1, q2))
__global__ void __launch_bounds__(256) sac_actor_fwd_kernel(const float* logp, const float* q1, const float* q2,
                                                            const float* alpha, long long B, float* loss) {
  __shared__ float red[8];
  const float al = alpha[0];
  float s = 0.f;
  for (long long i = threadIdx.x; i < B; i += 256)
    s = __fadd_rn(s, __fsub_rn(__fmul_rn(al, logp[i]), fminf(q1[i], q2[i])));
  s = block_sum_256(s, red);
  if (threadIdx.x == 0) loss[0] = s / (float)B;
}
// torch.min(a, b) backward: the smaller input takes the gradient, a tie splits it in halves
__global__ void sac_actor_bwd_kernel(const float* q1, const float* q2, const float* alpha, long long B,
                                     const float* gloss, float* glogp, float* gq1, float* gq2) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B) return;
  const float g = gloss[0] / (float)B;
  glogp[i] = g * alpha[0];
  const float a = q1[i], b = q2[i];
  gq1[i] = a < b ? -g : (a == b ? -0.5f * g : 0.f);
  gq2[i] = b < a ? -g : (a == b ? -0.5f * g : 0.f);
}

}  // namespace nlbac

using namespace nlbac;

extern "C" {

int nlbac_constraints4_fwd(const nlbac_constraint_desc* desc, const float* a, const float* b, const float* c,
                           const float* e, const float* V, const float* Vnext, int64_t B, float* req, void* stream) {
  ::nlbac::NvtxRange nvtx_range_("nlbac:constraints4_fwd");
  if (int rc = check_constraint(desc)) return rc;
  NLBAC_REQUIRE(a && b && V && Vnext && req, "null pointer");
  NLBAC_REQUIRE((desc->kind != NLBAC_CONSTRAINT_CARS && desc->kind != NLBAC_CONSTRAINT_PVTOL) || c,
                "cars / pvtol constraints need x_{t+2}");
  NLBAC_REQUIRE(desc->kind != NLBAC_CONSTRAINT_PVTOL || e, "pvtol constraints need x_{t+3}");
  NLBAC_REQUIRE(B >= 0, "negative batch");
  cudaStream_t st = (cudaStream_t)stream;
  NLBAC_CUDA(cudaMemsetAsync(req, 0, desc->m * 4, st));
  if (B == 0) return 0;
  constraints_fwd_kernel<<<tail_grid(B, CNT), CNT, 0, st>>>(*desc, a, b, c, e, V, Vnext, B, req);
  NLBAC_CUDA(cudaGetLastError());
  return 0;
}

int nlbac_constraints4_bwd(const nlbac_constraint_desc* desc, const float* a, const float* b, const float* c,
                           const float* e, const float* V, const float* Vnext, int64_t B, const float* grad_req,
                           float* grad_b, float* grad_c, float* grad_e, float* grad_Vnext, void* stream) {
  ::nlbac::NvtxRange nvtx_range_("nlbac:constraints4_bwd");
  if (int rc = check_constraint(desc)) return rc;
  NLBAC_REQUIRE(a && b && V && Vnext && grad_req && grad_b && grad_Vnext, "null pointer");
  NLBAC_REQUIRE((desc->kind != NLBAC_CONSTRAINT_CARS && desc->kind != NLBAC_CONSTRAINT_PVTOL) || (c && grad_c),
                "cars / pvtol constraints need x_{t+2} and its gradient");
  NLBAC_REQUIRE(desc->kind != NLBAC_CONSTRAINT_PVTOL || (e && grad_e), "pvtol constraints need x_{t+3} and its gradient");
  if (B <= 0) return 0;
  constraints_bwd_kernel<<<tail_grid(B, CNT), CNT, 0, (cudaStream_t)stream>>>(*desc, a, b, c, e, V, Vnext, B, grad_req,
                                                                              grad_b, grad_c, grad_e, grad_Vnext);
  NLBAC_CUDA(cudaGetLastError());
  return 0;
}

int nlbac_constraints_fwd(const nlbac_constraint_desc* desc, const float* a, const float* b, const float* c,
                          const float* V, const float* Vnext, int64_t B, float* req, void* stream) {
  ::nlbac::NvtxRange nvtx_range_("nlbac:constraints_fwd");
  NLBAC_REQUIRE(desc == nullptr || desc->kind != NLBAC_CONSTRAINT_PVTOL, "pvtol constraints take four states: use nlbac_constraints4_fwd");
  return nlbac_constraints4_fwd(desc, a, b, c, nullptr, V, Vnext, B, req, stream);
}

int nlbac_constraints_bwd(const nlbac_constraint_desc* desc, const float* a, const float* b, const float* c,
                          const float* V, const float* Vnext, int64_t B, const float* grad_req, float* grad_b,
                          float* grad_c, float* grad_Vnext, void* stream) {
  ::nlbac::NvtxRange nvtx_range_("nlbac:constraints_bwd");
  NLBAC_REQUIRE(desc == nullptr || desc->kind != NLBAC_CONSTRAINT_PVTOL, "pvtol constraints take four states: use nlbac_constraints4_bwd");
  return nlbac_constraints4_bwd(desc, a, b, c, nullptr, V, Vnext, B, grad_req, grad_b, grad_c, nullptr, grad_Vnext, stream);
}

int nlbac_lookahead_fwd(const float* state, int n_s, float l_p, int64_t B, float* p, void* stream) {
  ::nlbac::NvtxRange nvtx_range_("nlbac:lookahead_fwd");
  NLBAC_REQUIRE(state && p && n_s >= 3 && B >= 0, "bad arguments");
  if (B == 0) return 0;
  lookahead_fwd_kernel<<<tail_grid(B, 256), 256, 0, (cudaStream_t)stream>>>(state, n_s, l_p, B, p);
  NLBAC_CUDA(cudaGetLastError());
  return 0;
}
int nlbac_lookahead_bwd(const float* state, int n_s, float l_p, int64_t B, const float* grad_p, float* grad_state,
                        void* stream) {
  ::nlbac::NvtxRange nvtx_range_("nlbac:lookahead_bwd");
  NLBAC_REQUIRE(state && grad_p && grad_state && n_s >= 3 && B >= 0, "bad arguments");
  if (B == 0) return 0;
  lookahead_bwd_kernel<<<tail_grid(B, 256), 256, 0, (cudaStream_t)stream>>>(state, n_s, l_p, B, grad_p, grad_state);
  NLBAC_CUDA(cudaGetLastError());
  return 0;
}

int nlbac_get_state(int env, const float* obs, int64_t B, float* state, float* state_dyn, void* stream) {
  ::nlbac::NvtxRange nvtx_range_("nlbac:get_state");
  NLBAC_REQUIRE(env >= NLBAC_ENV_UNICYCLE && env <= NLBAC_ENV_PVTOL, "unknown env %d", env);
  NLBAC_REQUIRE(obs && state && B >= 0, "bad arguments");
  if (B == 0) return 0;
  get_state_kernel<<<tail_grid(B, 256), 256, 0, (cudaStream_t)stream>>>(env, obs, B, state, state_dyn);
  NLBAC_CUDA(cudaGetLastError());
  return 0;
}
int nlbac_get_obs_fwd(int env, const float* state, int64_t B, float* obs, void* stream) {
  ::nlbac::NvtxRange nvtx_range_("nlbac:get_obs_fwd");
  NLBAC_REQUIRE(env >= NLBAC_ENV_UNICYCLE && env <= NLBAC_ENV_PVTOL, "unknown env %d", env);
  NLBAC_REQUIRE(state && obs && B >= 0, "bad arguments");
  if (B == 0) return 0;
  get_obs_kernel<false><<<tail_grid(B, 256), 256, 0, (cudaStream_t)stream>>>(env, state, B, obs, nullptr, nullptr);
  NLBAC_CUDA(cudaGetLastError());
  return 0;
}
int nlbac_get_obs_bwd(int env, const float* state, int64_t B, const float* grad_obs, float* grad_state, void* stream) {
  ::nlbac::NvtxRange nvtx_range_("nlbac:get_obs_bwd");
  NLBAC_REQUIRE(env >= NLBAC_ENV_UNICYCLE && env <= NLBAC_ENV_PVTOL, "unknown env %d", env);
  NLBAC_REQUIRE(state && grad_obs && grad_state && B >= 0, "bad arguments");
  if (B == 0) return 0;
  get_obs_kernel<true><<<tail_grid(B, 256), 256, 0, (cudaStream_t)stream>>>(env, state, B, nullptr, grad_obs, grad_state);
  NLBAC_CUDA(cudaGetLastError());
  return 0;
}

int nlbac_adam_step(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, int64_t n, int step, float lr,
                    float beta1, float beta2, float eps, void* stream) {
  ::nlbac::NvtxRange nvtx_range_("nlbac:adam_step");
  NLBAC_REQUIRE(param && grad && exp_avg && exp_avg_sq && n >= 0 && step >= 1, "bad arguments");
  if (n == 0) return 0;
  const double bc1 = 1.0 - pow((double)beta1, step), bc2 = 1.0 - pow((double)beta2, step);
  adam_kernel<<<tail_grid(n, 256), 256, 0, (cudaStream_t)stream>>>(param, grad, exp_avg, exp_avg_sq, n, lr, beta1, beta2,
                                                                   eps, (float)bc1, (float)sqrt(bc2));
  NLBAC_CUDA(cudaGetLastError());
  return 0;
}
int nlbac_replay_gather(const float* ring, int width, const int64_t* idx, int64_t B, float* out, void* stream) {
  ::nlbac::NvtxRange nvtx_range_("nlbac:replay_gather");
  NLBAC_REQUIRE(width >= 1 && B >= 0, "bad sizes");
  if (B == 0) return 0;
  NLBAC_REQUIRE(ring && idx && out, "null pointer");
  replay_gather_kernel<<<tail_grid(B * width, 256), 256, 0, (cudaStream_t)stream>>>(
      ring, width, reinterpret_cast<const long long*>(idx), B, out);
  NLBAC_CUDA(cudaGetLastError());
  return 0;
}

int nlbac_replay_gather_split(const float* ring, int width, const int64_t* idx, int64_t B, int n_fields,
                              const int* offsets, const int* widths, float* const* outs, void* stream) {
  ::nlbac::NvtxRange nvtx_range_("nlbac:replay_gather_split");
  NLBAC_REQUIRE(ring && idx && offsets && widths && outs, "null pointer");
  NLBAC_REQUIRE(B >= 0 && width >= 1 && n_fields >= 1 && n_fields <= NLBAC_REPLAY_MAX_FIELDS, "bad sizes");
  if (B == 0) return 0;
  GatherSplit g;
  g.n_fields = n_fields;
  int prev_end = 0;
  for (int f = 0; f < n_fields; ++f) {
    NLBAC_REQUIRE(offsets[f] >= prev_end && widths[f] >= 0 && offsets[f] + widths[f] <= width,
                  "field %d: columns [%d, %d) out of order or outside the row", f, offsets[f], offsets[f] + widths[f]);
    g.out[f] = outs[f]; g.off[f] = offsets[f]; g.w[f] = widths[f];
    prev_end = offsets[f] + widths[f];
  }
  replay_gather_split_kernel<<<tail_grid(B * width, 256), 256, 0, (cudaStream_t)stream>>>(
      ring, width, reinterpret_cast<const long long*>(idx), B, g);
  NLBAC_CUDA(cudaGetLastError());
  return 0;
}

int nlbac_adam_step_multi(int n_tensors, float* const* params, const float* const* grads, float* const* exp_avg,
                          float* const* exp_avg_sq, const int64_t* numels, int step, float lr, float beta1,
                          float beta2, float eps, void* stream) {
  ::nlbac::NvtxRange nvtx_range_("nlbac:adam_step_multi");
  NLBAC_REQUIRE(n_tensors >= 0 && n_tensors <= NLBAC_MAX_TENSORS, "n_tensors=%d (max %d)", n_tensors, NLBAC_MAX_TENSORS);
  NLBAC_REQUIRE(params && grads && exp_avg && exp_avg_sq && numels && step >= 1, "bad arguments");
  if (n_tensors == 0) return 0;
  AdamTable tb;
  long long nmax = 0;
  for (int k = 0; k < n_tensors; ++k) {
    NLBAC_REQUIRE(numels[k] >= 0 && (numels[k] == 0 || (params[k] && grads[k] && exp_avg[k] && exp_avg_sq[k])),
                  "tensor %d: null pointer", k);
    tb.p[k] = params[k]; tb.g[k] = grads[k]; tb.m[k] = exp_avg[k]; tb.v[k] = exp_avg_sq[k]; tb.n[k] = numels[k];
    if (numels[k] > nmax) nmax = numels[k];
  }
  if (nmax == 0) return 0;
  const double bc1 = 1.0 - pow((double)beta1, step), bc2 = 1.0 - pow((double)beta2, step);
  adam_multi_kernel<<<dim3(tail_grid(nmax, 256), n_tensors), 256, 0, (cudaStream_t)stream>>>(tb, lr, beta1, beta2, eps,
                                                                                          (float)bc1, (float)sqrt(bc2));
  NLBAC_CUDA(cudaGetLastError());
  return 0;
}
int nlbac_adam_step_multi_dev(int n_tensors, float* const* params, const float* const* grads, float* const* exp_avg,
                              float* const* exp_avg_sq, const int64_t* numels, int* step_dev, int bump, float lr,
                              float beta1, float beta2, float eps, void* stream) {
  ::nlbac::NvtxRange nvtx_range_("nlbac:adam_step_multi_dev");
  NLBAC_REQUIRE(n_tensors >= 0 && n_tensors <= NLBAC_MAX_TENSORS, "n_tensors=%d (max %d)", n_tensors, NLBAC_MAX_TENSORS);
  NLBAC_REQUIRE(params && grads && exp_avg && exp_avg_sq && numels && step_dev, "bad arguments");
  cudaStream_t st = (cudaStream_t)stream;
  NLBAC_REQUIRE(n_tensors > 0 || !bump, "bump needs at least one tensor (the update kernel advances the counter)");
  if (n_tensors == 0) return 0;
  AdamTable tb;
  long long nmax = 0;
  for (int k = 0; k < n_tensors; ++k) {
    NLBAC_REQUIRE(numels[k] >= 0 && (numels[k] == 0 || (params[k] && grads[k] && exp_avg[k] && exp_avg_sq[k])),
                  "tensor %d: null pointer", k);
    tb.p[k] = params[k]; tb.g[k] = grads[k]; tb.m[k] = exp_avg[k]; tb.v[k] = exp_avg_sq[k]; tb.n[k] = numels[k];
    if (numels[k] > nmax) nmax = numels[k];
  }
  if (nmax == 0) nmax = 1;   // the launch still advances the counter
  adam_multi_dev_kernel<<<dim3(tail_grid(nmax, 256), n_tensors), 256, 0, st>>>(tb, step_dev, bump, lr, beta1, beta2, eps);
  NLBAC_CUDA(cudaGetLastError());
  return 0;
}
int nlbac_soft_update_multi(int n_tensors, float* const* targets, const float* const* sources, const int64_t* numels,
                            float tau, void* stream) {
  ::nlbac::NvtxRange nvtx_range_("nlbac:soft_update_multi");
  NLBAC_REQUIRE(n_tensors >= 0 && n_tensors <= NLBAC_MAX_TENSORS, "n_tensors=%d (max %d)", n_tensors, NLBAC_MAX_TENSORS);
  NLBAC_REQUIRE(targets && sources && numels, "bad arguments");
  if (n_tensors == 0) return 0;
  PolyakTable tb;
  long long nmax = 0;
  for (int k = 0; k < n_tensors; ++k) {
    NLBAC_REQUIRE(numels[k] >= 0 && (numels[k] == 0 || (targets[k] && sources[k])), "tensor %d: null pointer", k);
    tb.t[k] = targets[k]; tb.s[k] = sources[k]; tb.n[k] = numels[k];
    if (numels[k] > nmax) nmax = numels[k];
  }
  if (nmax == 0) return 0;
  soft_update_multi_kernel<<<dim3(tail_grid(nmax, 256), n_tensors), 256, 0, (cudaStream_t)stream>>>(tb, tau);
  NLBAC_CUDA(cudaGetLastError());
  return 0;
}
int nlbac_td_target(const float* q1, const float* q2, const float* logp, const float* alpha, const float* reward,
                    const float* mask, float gamma, int64_t B, float* out, void* stream) {
  NLBAC_REQUIRE(q1 && reward && mask && out && (logp == nullptr || alpha != nullptr), "null pointer");
  NLBAC_REQUIRE(B >= 0, "negative batch");
  if (B == 0) return 0;
  td_target_kernel<<<(unsigned)((B + 255) / 256), 256, 0, (cudaStream_t)stream>>>(q1, q2, logp, alpha, reward, mask,
                                                                                  gamma, B, out);
  NLBAC_CUDA(cudaGetLastError());
  return 0;
}

int nlbac_mse_mean_fwd(const float* pred, const float* target, int64_t n, float* loss, void* stream) {
  NLBAC_REQUIRE(pred && target && loss && n >= 1, "bad arguments");
  mse_mean_fwd_kernel<<<1, 256, 0, (cudaStream_t)stream>>>(pred, target, n, loss);
  NLBAC_CUDA(cudaGetLastError());
  return 0;
}

int nlbac_mse_mean_bwd(const float* pred, const float* target, int64_t n, const float* grad_loss, float* grad_pred,
                       void* stream) {
  NLBAC_REQUIRE(pred && target && grad_loss && grad_pred && n >= 1, "bad arguments");
  mse_mean_bwd_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(pred, target, n, grad_loss,
                                                                                     grad_pred);
  NLBAC_CUDA(cudaGetLastError());
  return 0;
}

int nlbac_sac_actor_term_fwd(const float* logp, const float* q1, const float* q2, const float* alpha, int64_t B,
                             float* loss, void* stream) {
  NLBAC_REQUIRE(logp && q1 && q2 && alpha && loss && B >= 1, "bad arguments");
  sac_actor_fwd_kernel<<<1, 256, 0, (cudaStream_t)stream>>>(logp, q1, q2, alpha, B, loss);
  NLBAC_CUDA(cudaGetLastError());
  return 0;
}

int nlbac_sac_actor_term_bwd(const float* q1, const float* q2, const float* alpha, int64_t B, const float* grad_loss,
                             float* grad_logp, float* grad_q1, float* grad_q2, void* stream) {
  NLBAC_REQUIRE(q1 && q2 && alpha && grad_loss && grad_logp && grad_q1 && grad_q2 && B >= 1, "bad arguments");
  sac_actor_bwd_kernel<<<(unsigned)((B + 255) / 256), 256, 0, (cudaStream_t)stream>>>(q1, q2, alpha, B, grad_loss,
                                                                                      grad_logp, grad_q1, grad_q2);
  NLBAC_CUDA(cudaGetLastError());
  return 0;
}

int nlbac_lagrangian_fwd(const float* req, int m, int clf_last, double* lambda, double* rho, const double* due,
                         const int64_t* updates, int interval, double cost_limit, double lambda_max, double aug_ratio, int use_ratio, int has_floor,
                         double ratio_floor, float* loss, float* coef, void* stream) {
  ::nlbac::NvtxRange nvtx_range_("nlbac:lagrangian_fwd");
  NLBAC_REQUIRE(req && lambda && rho && (due || updates) && loss && coef, "null pointer");
  NLBAC_REQUIRE(updates == nullptr || interval >= 1, "bad multiplier interval %d", interval);
  NLBAC_REQUIRE(m >= 1 && m <= 32 && (!clf_last || m >= 2), "bad number of constraint rows %d", m);
  LagArgs a{req, lambda, rho, due, reinterpret_cast<const long long*>(updates), loss, coef, cost_limit, lambda_max,
            aug_ratio, ratio_floor, interval, m, use_ratio, has_floor, clf_last};
  lagrangian_fwd_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(a);
  NLBAC_CUDA(cudaGetLastError());
  return 0;
}

int nlbac_lagrangian_bwd(const float* req, const float* coef, int m, double cost_limit, const float* grad_loss,
                         float* grad_req, void* stream) {
  NLBAC_REQUIRE(req && coef && grad_loss && grad_req && m >= 1 && m <= 32, "bad arguments");
  lagrangian_bwd_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(req, coef, m, (float)cost_limit, grad_loss, grad_req);
  NLBAC_CUDA(cudaGetLastError());
  return 0;
}

int nlbac_soft_update(float* target, const float* source, int64_t n, float tau, void* stream) {
  ::nlbac::NvtxRange nvtx_range_("nlbac:soft_update");
  NLBAC_REQUIRE(target && source && n >= 0, "bad arguments");
  if (n == 0) return 0;
  soft_update_kernel<<<tail_grid(n, 256), 256, 0, (cudaStream_t)stream>>>(target, source, n, tau);
  NLBAC_CUDA(cudaGetLastError());
  return 0;
}

}  // extern "C"
