// Optimizer math shared by the single-GPU fused update (core.cu) and the in-kernel
// all-reduce + update (comm/collectives.cu).  Semantics: torch.optim.Adam (no amsgrad,
// bias-corrected, eps added after the sqrt, L2 weight decay folded into the gradient) and
// torch.optim.SGD (optional momentum) — src/motion/trainer/base.py:42-43, example_ddp.py:54.
#pragma once
#include <math.h>
#include "common.cuh"

namespace b200rnn {

struct OptScalars {
  float lr, beta1, beta2, eps, wd, momentum, grad_scale;
  float step_size, inv_sqrt_bc2;  // Adam bias corrections, computed in double on the host
  int kind;
};

static inline OptScalars make_opt_scalars(const b200rnn_optim_desc* opt) {
  OptScalars o;
  o.lr = opt->lr; o.beta1 = opt->beta1; o.beta2 = opt->beta2; o.eps = opt->eps;
  o.wd = opt->weight_decay; o.momentum = opt->momentum; o.grad_scale = opt->grad_scale;
  o.kind = opt->kind;
  if (opt->kind == 0) {
    const double bc1 = 1.0 - pow((double)opt->beta1, (double)opt->step);
    const double bc2 = 1.0 - pow((double)opt->beta2, (double)opt->step);
    o.step_size = (float)((double)opt->lr / bc1);
    o.inv_sqrt_bc2 = (float)(1.0 / sqrt(bc2));
  } else {
    o.step_size = opt->lr; o.inv_sqrt_bc2 = 1.0f;
  }
  return o;
}

__device__ __forceinline__ void apply_update(const OptScalars& o, float g, float& p, float& m, float& v) {
  if (o.kind == 0) {
    if (o.wd != 0.0f) g = fmaf(o.wd, p, g);
    m = fmaf(o.beta1, m, (1.0f - o.beta1) * g);
    v = fmaf(o.beta2, v, (1.0f - o.beta2) * g * g);
    const float denom = sqrtf(v) * o.inv_sqrt_bc2 + o.eps;
    p -= o.step_size * (m / denom);
  } else {
    if (o.wd != 0.0f) g = fmaf(o.wd, p, g);
    if (o.momentum != 0.0f) { m = fmaf(o.momentum, m, g); g = m; }
    p -= o.lr * g;
  }
}

}  // namespace b200rnn
