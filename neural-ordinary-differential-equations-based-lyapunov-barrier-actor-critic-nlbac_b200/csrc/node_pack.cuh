// Parameter packing: nn.Linear tensors (state-dict order) -> padded / transposed kernel image.
#pragma once
#include "node_cfg.cuh"

namespace nlbac {

struct PackSeg {
  const float* src;
  int dst_off;   // floats into the packed image
  int rows, cols;          // destination extent (row-major, ld = cols)
  int src_rows, src_cols;  // source extent (row-major, ld = src_cols)
  int transpose;           // dst[r][c] = src[c][r]
};

constexpr int MAX_PACK_SEGS = 40;
struct PackTable {
  PackSeg seg[MAX_PACK_SEGS];
  int n;
};

static __global__ void pack_params_kernel(PackTable tab, float* __restrict__ packed) {
  const PackSeg s = tab.seg[blockIdx.y];
  const int total = s.rows * s.cols;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int r = i / s.cols, c = i - r * s.cols;
    float v = 0.f;
    if (s.transpose) {
      if (c < s.src_rows && r < s.src_cols) v = s.src[c * s.src_cols + r];
    } else {
      if (r < s.src_rows && c < s.src_cols) v = s.src[r * s.src_cols + c];
    }
    packed[s.dst_off + i] = v;
  }
}

// params: W0,b0,W1,b1,... chain 0 then chain 1 (device pointers held in a host array)
template <class C>
PackTable build_pack_table(const float* const* params) {
  PackTable t;
  t.n = 0;
  auto add = [&](const float* src, int off, int rows, int cols, int sr, int sc, int tr) {
    PackSeg& s = t.seg[t.n++];
    s.src = src; s.dst_off = off; s.rows = rows; s.cols = cols; s.src_rows = sr; s.src_cols = sc; s.transpose = tr;
  };
  int pi = 0;
  for (int c = 0; c < C::NCH; ++c) {
    const int NL = C::nl(c);
    for (int l = 0; l < NL; ++l) {
      const float* W = params[pi++];
      const float* b = params[pi++];
      if (l == 0) {
        add(W, C::off_wt0(c), C::INP0, C::HP, C::H, C::IN0, 1);
        add(b, C::off_b0(c), 1, C::HP, 1, C::H, 0);
        add(W, C::off_wn0(c), C::H, C::INP0, C::H, C::IN0, 0);
      } else if (l == NL - 1) {
        add(W, C::off_wtl(c), C::H, C::outp(c), C::out(c), C::H, 1);
        add(b, C::off_bl(c), 1, C::outp(c), 1, C::out(c), 0);
        add(W, C::off_wnl(c), C::outp(c), C::HP, C::out(c), C::H, 0);
      } else {
        const int h = C::hid_base(c) + (l - 1);
        add(W, C::off_hid_t(h), C::H, C::HP, C::H, C::H, 1);
        add(W, C::off_hid_n(h), C::H, C::HP, C::H, C::H, 0);
        add(b, C::off_bh(c, l), 1, C::HP, 1, C::H, 0);
      }
    }
  }
  return t;
}

template <class C>
int launch_pack(const float* const* params, float* packed, cudaStream_t st) {
  PackTable t = build_pack_table<C>(params);
  dim3 grid(8, t.n);
  pack_params_kernel<<<grid, 256, 0, st>>>(t, packed);
  NLBAC_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace nlbac
