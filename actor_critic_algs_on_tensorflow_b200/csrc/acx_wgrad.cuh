// Internal interface of the tensor-core weight-gradient kernel (acx_wgrad.cu), called by acx_mlp_tc_train.cu.
#pragma once
#include "acx_common.cuh"

struct AcxWgradArgs {
  const float* x2;      // [rows][128][n]  activation cache, layer 2
  const float* d3;      // [rows][H3][n]   dPre3
  const float* x1;      // [rows][H1][n]   layer-1 activations
  const float* d2;      // [rows][128][n]  dPre2
  int rows, n, H1, H3;
  float* scratch;       // per-CTA gradient rows [grid][scr_stride]; dW3 at offW3 ([128][H3]), dW2 at offW2 ([H1][128])
  int scr_stride, offW3, offW2;
  int grid;             // CTAs = scratch rows; CTA c owns the 128-sample tiles c, c + grid, ...
};

int acx_wgrad_launch(const AcxWgradArgs& a, void* stream);
