// loss.cuh — KGE/loss.py:49-204 per (positive, negative) pair, with dL/dscore
// (SURVEY.md Appendix E "Losses").  n = positives in the batch, k = negatives each.
#pragma once
#include "common.cuh"

namespace kge {

struct LossDev {
  int kind;            // kge_loss_kind
  float margin, temperature;
  float inv_n, inv_nk; // 1/n, 1/(n*k)
};

__device__ __forceinline__ float sigmoidf_(float x) { return 1.0f / (1.0f + expf(-x)); }
// tf.math.log_sigmoid(x) = -softplus(-x), evaluated stably
__device__ __forceinline__ float logsigmoidf_(float x) { return fminf(x, 0.f) - log1pf(expf(-fabsf(x))); }

// softplus(x) = -log_sigmoid(-x) and sigmoid(x) from ONE exponential (the pair the BCE / self-adversarial
// negative terms need; every lane of the warp evaluates it once per negative)
__device__ __forceinline__ void softplus_sigmoid(float x, float& sp, float& sg) {
  // e in (0, 1]: the MUFU-based exp2 / log2 / rcp are accurate to ~2 ulp here (|arg| of ex2 is the
  // only error amplifier and exp(-|x|) underflows harmlessly), far inside the 1e-4 parity tolerance
  const float e = __expf(-fabsf(x));
  const float d = 1.0f + e;
  const float r = __frcp_rn(d);
  sp = fmaxf(x, 0.f) + ((e > 1e-4f) ? __logf(d) : e * (1.0f - 0.5f * e));   // log1p(e) for tiny e: series
  sg = (x >= 0.f) ? r : e * r;
}

// Contribution of negative j: loss term, d/dneg, and the part of d/dpos it carries.
// w = self-adversarial weight softmax_j(alpha * neg) (1 otherwise).
__device__ __forceinline__ void loss_neg(const LossDev& L, float sp, float sn, float w,
                                         float& term, float& dneg, float& dpos_part) {
  switch (L.kind) {
    case KGE_LOSS_HINGE: {           // loss.py:81-82
      float a = L.margin + sn - sp;
      term = fmaxf(a, 0.f) * L.inv_nk;
      dneg = (a >= 0.f) ? L.inv_nk : 0.f;     // clip_by_value passes gradient at the boundary
      dpos_part = -dneg;
      break;
    }
    case KGE_LOSS_LOGISTIC: {        // loss.py:112-113, naive log(1 + exp)
      float e = expf(sn - sp);
      term = logf(1.0f + e);
      dneg = e / (1.0f + e);
      dpos_part = -dneg;
      break;
    }
    case KGE_LOSS_BCE: {             // loss.py:140-143: -log_sigmoid(-neg)
      float sp_, sg_;
      softplus_sigmoid(sn, sp_, sg_);
      term = sp_ * L.inv_n;
      dneg = sg_ * L.inv_n;
      dpos_part = 0.f;
      break;
    }
    case KGE_LOSS_SELF_ADV: {        // loss.py:176-182 (weights are stop_gradient)
      float sp_, sg_;
      softplus_sigmoid(sn + L.margin, sp_, sg_);
      term = w * sp_ * L.inv_n;
      dneg = w * sg_ * L.inv_n;
      dpos_part = 0.f;
      break;
    }
    default:                         // KGE_LOSS_SQUARE, loss.py:201-204
      term = 0.5f * sn * sn * L.inv_n;
      dneg = sn * L.inv_n;
      dpos_part = 0.f;
  }
}

// Terms that depend on the positive alone.
__device__ __forceinline__ void loss_pos(const LossDev& L, float sp, float& term, float& dpos) {
  switch (L.kind) {
    case KGE_LOSS_BCE:
      term = -logsigmoidf_(sp) * L.inv_n;
      dpos = -sigmoidf_(-sp) * L.inv_n;
      break;
    case KGE_LOSS_SELF_ADV:
      term = -logsigmoidf_(sp + L.margin) * L.inv_n;
      dpos = -sigmoidf_(-(sp + L.margin)) * L.inv_n;
      break;
    case KGE_LOSS_SQUARE:
      term = 0.5f * (sp - 1.0f) * (sp - 1.0f) * L.inv_n;
      dpos = (sp - 1.0f) * L.inv_n;
      break;
    default:
      term = 0.f; dpos = 0.f;
  }
}

}  // namespace kge
