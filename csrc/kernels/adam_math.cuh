// Adam arithmetic shared by the local optimizer kernel and the fused reduce+Adam+broadcast
// collective.  Matches torch.optim.Adam (L2-style weight decay folded into the gradient,
// bias-corrected moments, eps added outside the square root).
#pragma once
#include <cuda_runtime.h>
#include <math.h>

namespace dlbb {

struct AdamHyper {
    float lr;
    float beta1, beta2;
    float eps;
    float weight_decay;
    float grad_scale;   // multiplies the raw gradient (1/world_size, loss-scale removal)
    float bc1_inv;      // 1 / (1 - beta1^t)
    float bc2_inv_sqrt; // 1 / sqrt(1 - beta2^t)
};

inline AdamHyper make_adam_hyper(float lr, float beta1, float beta2, float eps, float weight_decay,
                                 float grad_scale, int step) {
    AdamHyper h;
    h.lr = lr; h.beta1 = beta1; h.beta2 = beta2; h.eps = eps;
    h.weight_decay = weight_decay; h.grad_scale = grad_scale;
    h.bc1_inv = static_cast<float>(1.0 / (1.0 - pow(static_cast<double>(beta1), step)));
    h.bc2_inv_sqrt = static_cast<float>(1.0 / sqrt(1.0 - pow(static_cast<double>(beta2), step)));
    return h;
}

__device__ __forceinline__ void adam_update(float& w, float& m, float& v, float g_raw,
                                            const AdamHyper& h) {
    float g = g_raw * h.grad_scale;
    if (h.weight_decay != 0.f) g = fmaf(h.weight_decay, w, g);
    m = fmaf(h.beta1, m, (1.f - h.beta1) * g);
    v = fmaf(h.beta2, v, (1.f - h.beta2) * g * g);
    const float denom = sqrtf(v) * h.bc2_inv_sqrt + h.eps;
    w -= h.lr * (m * h.bc1_inv) / denom;
}

}  // namespace dlbb
