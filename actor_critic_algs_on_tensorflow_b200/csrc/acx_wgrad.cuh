// Internal interface of the tensor-core weight-gradient kernel (acx_wgrad.cu), called by acx_mlp_tc_train.cu.
#pragma once
#include "acx_common.cuh"

struct AcxWgradArgs {
  const float* x2;      // [rows][128][n]  activation cache, layer 2
  const float* d3;      // [rows][H3][n]   dPre3
  const float* obs;     // [rows][ob][n]   observations (layer 1 is recomputed from them)
  const float* params;  // flat network parameters (W1 at offW1 [ob][H1], b1 at offb1)
  int ob, offW1, offb1, lrelu;
  const float* d2;      // [rows][128][n]  dPre2
  int rows, n, H1, H3;
  float* scratch;       // per-CTA gradient rows [grid][scr_stride]; dW3 at offW3 ([128][H3]), dW2 at offW2 ([H1][128])
  int scr_stride, offW3, offW2, offb3, offb2;   // db3 [H3] at offb3, db2 [128] at offb2
  int grid;             // CTAs = scratch rows; CTA c owns the 128-sample tiles c, c + grid, ...
};

int acx_wgrad_launch(const AcxWgradArgs& a, void* stream);

// chain kernel of the two-kernel form (acx_mlp_tc_chain.cu): head, back-propagation products, dW1/db1/head gradients
// into a.scratch rows; dPre3 -> d3 [rows][H3][n], dPre2 -> d2 [rows][128][n].  head: acxmlp::HEAD_*.
#include "acx_mlp.cuh"
int acx_tc_chain_launch(const acxmlp::Args& a, const float* packed, float* d3, float* d2, int head, int grid,
                        void* stream);
