// Update rules shared by the linear-learner kernels (PA / PA-I / PA-II, PA regressor, Pegasos SVM).
// Literature formulas (Crammer et al. 2006; Shalev-Shwartz et al. 2007); the reference names the
// learners only (utils/parsers/requestStream/PipelineMap.scala:68).
#pragma once
#include <cuda_runtime.h>

namespace omldm {

// kind: 0 = PA classifier, 1 = PA regressor, 2 = SVM (Pegasos)
// variant (PA family): 0 = PA, 1 = PA-I, 2 = PA-II

__device__ __forceinline__ float pa_tau(float loss, float nrm2, int variant, float C) {
  if (loss <= 0.f || nrm2 <= 0.f) return 0.f;
  if (variant == 0) return loss / nrm2;
  if (variant == 1) return fminf(C, loss / nrm2);
  return loss / (nrm2 + 1.f / (2.f * C));
}

// Returns the coefficient c such that w += c * x, and reports loss / aux metric.
__device__ __forceinline__ float linear_coef(int kind, int variant, float C, float eps, float f,
                                             float y_raw, float nrm2, float& loss, float& aux) {
  if (kind == 0) {
    const float y = y_raw > 0.f ? 1.f : -1.f;
    loss = fmaxf(0.f, 1.f - y * f);
    aux = (y * f <= 0.f) ? 1.f : 0.f;  // mistake indicator
    return y * pa_tau(loss, nrm2, variant, C);
  } else if (kind == 1) {
    const float err = f - y_raw;
    loss = fmaxf(0.f, fabsf(err) - eps);
    aux = err * err;  // squared error
    return (err > 0.f ? -1.f : 1.f) * pa_tau(loss, nrm2, variant, C);
  }
  const float y = y_raw > 0.f ? 1.f : -1.f;
  loss = fmaxf(0.f, 1.f - y * f);
  aux = (y * f <= 0.f) ? 1.f : 0.f;
  return 0.f;  // SVM handled by the caller (needs the step counter)
}

}  // namespace omldm
