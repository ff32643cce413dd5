/*
 * step_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * Plain-C, single-threaded CPU restatement of the arithmetic on rocket's per-iteration
 * training step.  The reference itself (dsenushkin/rocket) does no arithmetic: its
 * Optimizer capsule calls `self._optimizer.step(); self._optimizer.zero_grad()`
 * (rocket/core/optimizer.py:129-130) and everything numeric happens inside torch, a
 * third-party dependency pinned only as `torch>=1.13.0` (requirements.txt:1); the version
 * restated here is torch 2.11.0 as installed on this image ($TORCH below =
 * site-packages/torch).  Each function cites the torch source it follows.
 *
 * Pinning: tests/test_oracle.py checks every function here BIT-EXACTLY against torch's own
 * CPU kernels (torch.optim.AdamW/Adam/SGD with foreach=False, torch.argmax) and against
 * tests/golden/*.npz, which were produced by running the UNMODIFIED reference Looper step
 * from /root/reference (tests/golden/make_golden.py).  The fused-multiply-add placement
 * below is what torch 2.11's vectorised CPU kernels do; it was found by probing and is
 * locked in by those bit-exact tests.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
 * may load this.  Nothing under rocket_b200/ does.
 *
 * Build: gcc -O2 -ffp-contract=off -shared -fPIC step_oracle.c -lm   (oracle/build.py)
 * -ffp-contract=off: the only fused operations are the explicit fmaf() calls.
 */
#include <math.h>
#include <stdint.h>
#include <string.h>

/* clip_coef_clamped: $TORCH/nn/utils/clip_grad.py:165-169.
 * `max_norm / (total_norm + 1e-6)` is Tensor.__rdiv__ = reciprocal() * max_norm
 * ($TORCH/_tensor.py:1119-1120), all in fp32; clamp(max=1.0) keeps NaN.            */
float oracle_clip_coef(float total_norm, float max_norm) {
  float c = (1.0f / (total_norm + 1e-6f)) * max_norm;
  return c > 1.0f ? 1.0f : c;
}

/* Two-level L2 norm: per-tensor norms, then the norm of those
 * ($TORCH/nn/utils/clip_grad.py:96-107).  Accumulated in double and rounded once, which
 * is what a correctly rounded fp32 result needs; torch's own reduction order is an
 * implementation detail of its CPU/GPU kernels (tests allow 2 ulp against it).       */
float oracle_l2norm_multi(const float* const* g, const int64_t* numel, int n_tensors,
                          float grad_scale, float* tensor_norms /* may be NULL */) {
  double tot = 0.0;
  for (int t = 0; t < n_tensors; ++t) {
    double s = 0.0;
    for (int64_t i = 0; i < numel[t]; ++i) {
      float y = g[t][i] * grad_scale;
      s += (double)y * (double)y;
    }
    float nt = (float)sqrt(s);
    if (tensor_norms) tensor_norms[t] = nt;
    tot += (double)nt * (double)nt;
  }
  return (float)sqrt(tot);
}

static inline float prolog(float g, float grad_scale, int has_clip, float coef) {
  if (grad_scale != 1.0f) g = g * grad_scale; /* DDP's 1/world_size */
  if (has_clip) g = g * coef;                 /* clip_grad.py:182 g.mul_(coef) */
  return g;
}

/* _single_tensor_adam, capturable=False, amsgrad=False, maximize=False
 * ($TORCH/optim/adam.py:396-547).  `step` is the value after `step_t += 1` (:414).   */
void oracle_adam_step(float* p, float* g, float* m, float* v, int64_t n, double lr,
                      double beta1, double beta2, double eps, double weight_decay,
                      int64_t step, int decoupled, float grad_scale, int has_clip,
                      float total_norm, float max_norm, int zero_grads) {
  const float coef = has_clip ? oracle_clip_coef(total_norm, max_norm) : 1.0f;
  const float decay = (float)(1.0 - lr * weight_decay);          /* :419 */
  const float wd = (float)weight_decay;                          /* :430 */
  const float w1 = (float)(1.0 - beta1);                         /* :457 */
  const float b2 = (float)beta2, w2 = (float)(1.0 - beta2);      /* :476 */
  const double bc1 = 1.0 - pow(beta1, (double)step);             /* :531 */
  const double bc2 = 1.0 - pow(beta2, (double)step);             /* :532 */
  const float neg_step = (float)(-(lr / bc1));                   /* :534, :547 */
  const float s2 = (float)pow(bc2, 0.5);                         /* :536 */
  const float epsf = (float)eps;
  for (int64_t i = 0; i < n; ++i) {
    float gi = prolog(g[i], grad_scale, has_clip, coef);
    float pi = p[i];
    if (weight_decay != 0.0) {
      if (decoupled) pi = pi * decay;                            /* param.mul_(1 - lr*wd) */
      else gi = fmaf(wd, pi, gi);                                /* grad.add(param, alpha=wd) */
    }
    /* exp_avg.lerp_(grad, 1-beta1): ATen lerp = fmadd(w, end-start, start) for |w| < 0.5,
     * else fmadd(w-1, end-start, end).                                              */
    const float d = gi - m[i];
    const float mi = fabsf(w1) < 0.5f ? fmaf(w1, d, m[i]) : fmaf(w1 - 1.0f, d, gi);
    /* exp_avg_sq.mul_(beta2).addcmul_(grad, grad, value=1-beta2): self + (value*t1)*t2, fused */
    const float vi = fmaf(w2 * gi, gi, v[i] * b2);
    const float denom = sqrtf(vi) / s2 + epsf;                   /* :545 */
    pi = pi + (neg_step * mi) / denom;                           /* :547 addcdiv_: self + value*t1/t2 */
    p[i] = pi; m[i] = mi; v[i] = vi;
    if (zero_grads) g[i] = 0.0f;                                 /* optimizer.py:130 (zero-fill form) */
  }
}

/* _single_tensor_sgd ($TORCH/optim/sgd.py:343-380), maximize=False.
 * `buf` may be NULL when momentum == 0.  first_step: momentum buffer does not exist yet. */
void oracle_sgd_step(float* p, float* g, float* buf, int64_t n, double lr, double momentum,
                     double dampening, double weight_decay, int nesterov, int first_step,
                     float grad_scale, int has_clip, float total_norm, float max_norm,
                     int zero_grads) {
  const float coef = has_clip ? oracle_clip_coef(total_norm, max_norm) : 1.0f;
  const float wd = (float)weight_decay, mu = (float)momentum;
  const float omd = (float)(1.0 - dampening), neg_lr = (float)(-lr);
  for (int64_t i = 0; i < n; ++i) {
    float gi = prolog(g[i], grad_scale, has_clip, coef);
    float pi = p[i];
    if (weight_decay != 0.0) gi = fmaf(wd, pi, gi);              /* :356 */
    if (momentum != 0.0) {
      float bi;
      if (first_step) bi = gi;                                   /* :362 clone */
      else bi = fmaf(omd, gi, buf[i] * mu);                      /* :365 */
      buf[i] = bi;
      gi = nesterov ? fmaf(mu, bi, gi) : bi;                     /* :368-370 */
    }
    p[i] = fmaf(neg_lr, gi, pi);                                 /* :380 */
    if (zero_grads) g[i] = 0.0f;
  }
}

/* The example metric of the reference, examples/mnist.py:26-34:
 *   pr = logits.argmax(1); total += pr.numel(); positive += sum(gt == pr)
 * torch.argmax: first maximal index; NaN compares as the maximum.                    */
void oracle_argmax_count(const float* logits, const int64_t* labels, int64_t n_rows,
                         int64_t n_cols, int64_t row_stride, int64_t* counters) {
  int64_t hits = 0;
  for (int64_t r = 0; r < n_rows; ++r) {
    const float* row = logits + r * row_stride;
    int64_t best = -1;
    float bv = 0.0f;
    for (int64_t j = 0; j < n_cols; ++j) {
      const float x = row[j];
      if (best < 0) { best = j; bv = x; continue; }
      if (bv != bv) break;                 /* already NaN: first NaN wins */
      if (x != x || x > bv) { best = j; bv = x; }
    }
    if (best >= 0 && best == labels[r]) ++hits;
  }
  counters[0] += hits;
  counters[1] += n_rows;
}

/* Loss bookkeeping, rocket/core/loss.py:95-98:  value += loss.item() / accumulation_steps
 * (Python floats = IEEE double).                                                      */
double oracle_scalar_accumulate(double acc, float value, double divisor) {
  return acc + (double)value / divisor;
}
