// Compile-time description of a NODE right-hand side and of its packed parameter image.
//
// A model is 1 or 2 ReLU MLP "chains" that share the input:
//   affine    (U/P/BU/BP  model.py NeuralODEModel(in, out1, out2)):  x' = f_net(x) + g_net(x).reshape(n_s,n_u) u
//   black-box (C          model.py NeuralODEModel(in, out)):         x' = net([x, u, const])
// Every chain is  IN0 -> H -> ... -> H -> OUT  with NL Linear layers (NL-2 hidden HxH matrices).
//
// Packed image (floats), produced by pack_params_kernel, consumed by the solver kernels:
//   [ RES_FWD | RES_BWD | HID_T[NHID] | HID_N[NHID] ]
//   RES_FWD per chain: Wt0 [INP0][HP] (first layer, transposed, zero padded), b0 [HP],
//                      hidden biases [NHC][HP], WtL [H][OUTP] (last layer transposed), bL [OUTP]
//   RES_BWD per chain: Wn0 [H][INP0] (first layer, natural), WnL [OUTP][HP] (last layer, natural)
//   HID_T[h]: hidden matrix h transposed  [k_in][HP]   (forward:  out[n] = sum_k a[k] Wt[k][n])
//   HID_N[h]: hidden matrix h natural     [j_out][HP]  (backward: dx[k] = sum_j d[j] Wn[j][k])
// HID_* are the packets streamed through the shared-memory ring by 1-D TMA bulk copies.
#pragma once
#include "common.cuh"

namespace nlbac {

constexpr int KIND_AFFINE = 0;
constexpr int KIND_BLACKBOX = 1;

template <int KIND_, int NS_, int NU_, int NC_, int H_, int NF_, int NG_>
struct NodeCfg {
  static constexpr int KIND = KIND_, NS = NS_, NU = NU_, NC = NC_, H = H_, NF = NF_, NG = NG_;
  static constexpr int D = NS + NU + NC;
  static constexpr int NCONST = NU + NC;
  static constexpr int NCH = (KIND == KIND_AFFINE) ? 2 : 1;
  static constexpr int IN0 = (KIND == KIND_AFFINE) ? NS : D;
  static constexpr int INP0 = round_up(IN0, 4);
  static constexpr int HP = round_up(H, 8);
  static constexpr int RS = HP + 4;  // activation row stride: RS/4 odd -> conflict-free LDS.128 over rows
  static constexpr int HID_FLOATS = H * HP;
  static constexpr int HID_BYTES = HID_FLOATS * 4;

  __host__ __device__ static constexpr int nl(int c) { return c == 0 ? NF : NG; }
  __host__ __device__ static constexpr int nhc(int c) { return nl(c) - 2; }
  __host__ __device__ static constexpr int out(int c) { return c == 0 ? NS : NS * NU; }
  __host__ __device__ static constexpr int outp(int c) { return round_up(out(c), 4); }
  static constexpr int NHID = nhc(0) + (NCH == 2 ? nhc(1) : 0);
  static constexpr int OUTP_MAX = (NCH == 2 && outp(1) > outp(0)) ? outp(1) : outp(0);
  static constexpr int NL_MAX = (NCH == 2 && NG > NF) ? NG : NF;
  __host__ __device__ static constexpr int hid_base(int c) { return c == 0 ? 0 : nhc(0); }

  // ---- RES_FWD ----
  __host__ __device__ static constexpr int res_fwd_chain(int c) {
    return INP0 * HP + HP + nhc(c) * HP + H * outp(c) + outp(c);
  }
  __host__ __device__ static constexpr int off_chain_fwd(int c) { return c == 0 ? 0 : res_fwd_chain(0); }
  __host__ __device__ static constexpr int off_wt0(int c) { return off_chain_fwd(c); }
  __host__ __device__ static constexpr int off_b0(int c) { return off_wt0(c) + INP0 * HP; }
  __host__ __device__ static constexpr int off_bh(int c, int l /*1..nhc*/) { return off_b0(c) + HP + (l - 1) * HP; }
  __host__ __device__ static constexpr int off_wtl(int c) { return off_b0(c) + HP + nhc(c) * HP; }
  __host__ __device__ static constexpr int off_bl(int c) { return off_wtl(c) + H * outp(c); }
  static constexpr int RES_FWD = res_fwd_chain(0) + (NCH == 2 ? res_fwd_chain(1) : 0);
  // ---- RES_BWD ----
  __host__ __device__ static constexpr int res_bwd_chain(int c) { return H * INP0 + outp(c) * HP; }
  __host__ __device__ static constexpr int off_wn0(int c) { return RES_FWD + (c == 0 ? 0 : res_bwd_chain(0)); }
  __host__ __device__ static constexpr int off_wnl(int c) { return off_wn0(c) + H * INP0; }
  static constexpr int RES_BWD = res_bwd_chain(0) + (NCH == 2 ? res_bwd_chain(1) : 0);
  static constexpr int RES_ALL = RES_FWD + RES_BWD;
  // ---- streamed packets ----
  __host__ __device__ static constexpr int off_hid_t(int h) { return RES_ALL + h * HID_FLOATS; }
  __host__ __device__ static constexpr int off_hid_n(int h) { return RES_ALL + (NHID + h) * HID_FLOATS; }
  static constexpr int PACKED = RES_ALL + 2 * NHID * HID_FLOATS;

  // ---- flat parameter vector, state-dict order ----
  __host__ __device__ static constexpr int lin(int c, int l) { return l == 0 ? IN0 : H; }
  __host__ __device__ static constexpr int lout(int c, int l) { return l == nl(c) - 1 ? out(c) : H; }
  __host__ __device__ static constexpr int chain_params(int c) {
    int n = 0;
    for (int l = 0; l < nl(c); ++l) n += lin(c, l) * lout(c, l) + lout(c, l);
    return n;
  }
  __host__ __device__ static constexpr int poff_w(int c, int l) {
    int n = (c == 0) ? 0 : chain_params(0);
    for (int i = 0; i < l; ++i) n += lin(c, i) * lout(c, i) + lout(c, i);
    return n;
  }
  __host__ __device__ static constexpr int poff_b(int c, int l) { return poff_w(c, l) + lin(c, l) * lout(c, l); }
  static constexpr int NPARAM = chain_params(0) + (NCH == 2 ? chain_params(1) : 0);

  // ---- gradient accumulation image (same tensors, every W/b start rounded to 16 bytes so the
  //      dW tiles can be reduced into it with red.global.add.v4.f32) ----
  __host__ __device__ static constexpr int goff_w(int c, int l) {
    int n = 0;
    for (int cc = 0; cc <= c; ++cc)
      for (int i = 0; i < (cc == c ? l : nl(cc)); ++i)
        n += round_up(lin(cc, i) * lout(cc, i), 4) + round_up(lout(cc, i), 4);
    return n;
  }
  __host__ __device__ static constexpr int goff_b(int c, int l) { return goff_w(c, l) + round_up(lin(c, l) * lout(c, l), 4); }
  static constexpr int NGRAD = goff_b(NCH - 1, nl(NCH - 1) - 1) + round_up(out(NCH - 1), 4);

  static_assert(NF >= 3 && (NCH == 1 || NG >= 3), "need at least one hidden matrix per chain");
  static_assert(HP % 8 == 0 && (RS / 4) % 2 == 1, "layout assumptions");
  static_assert(RES_FWD % 4 == 0 && RES_ALL % 4 == 0 && HID_FLOATS % 4 == 0, "16-byte alignment of packets");
};

// the shapes the reference ships + the assumed quadrotor shape
using CfgU = NodeCfg<KIND_AFFINE, 3, 2, 0, 100, 5, 4>;     // Unicycle  NeuralODEModel(3,3,6)
using CfgP = NodeCfg<KIND_AFFINE, 6, 2, 0, 100, 5, 4>;     // PVTOL     NeuralODEModel(6,6,12)
using CfgC = NodeCfg<KIND_BLACKBOX, 10, 1, 1, 64, 4, 0>;   // Cars      NeuralODEModel(12,10)
using CfgQ = NodeCfg<KIND_BLACKBOX, 6, 2, 0, 64, 4, 0>;    // Quadrotor 8 -> 6, hidden 64 assumed

static_assert(CfgU::NPARAM == 52209 && CfgP::NPARAM == 53718 && CfgC::NPARAM == 9802, "SURVEY App. C");

}  // namespace nlbac
