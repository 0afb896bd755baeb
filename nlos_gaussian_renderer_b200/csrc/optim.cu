// optim.cu -- fused multi-tensor Adam step for the Gaussian parameter groups.
//
// The step that follows the backward pass in the reference's fitting loop: GaussianModel.training_setup builds ONE torch.optim.Adam
// over six parameter groups (mu, f_dc, f_rest, opacity, scaling, rotation) with per-group learning rates, eps = 1e-15, default betas,
// no weight decay (/root/reference/gaussian_model/gaussian_model.py:223-249; stepped at main.py:161,213).  torch launches several
// kernels per group; here every group is one segment of ONE launch (HBM-bound: 16 B read + 12 B written per parameter).
// Arithmetic follows torch.optim.Adam's single-tensor path operation by operation:
//   m += (g - m) (1 - b1) ; v = v b2 + (1 - b2) g g ; p -= (lr / (1 - b1^t)) * m / (sqrt(v) / sqrt(1 - b2^t) + eps)
#include "common.cuh"
#include "../../include/nlosb200.h"

using namespace nlos;

#define ADAM_MAX_TENSORS 8

struct AdamArgs {
    float* p[ADAM_MAX_TENSORS];
    const float* g[ADAM_MAX_TENSORS];
    float* m[ADAM_MAX_TENSORS];
    float* v[ADAM_MAX_TENSORS];
    long long end[ADAM_MAX_TENSORS];     // exclusive prefix ends of the segments in the flattened index space
    float step_size[ADAM_MAX_TENSORS];   // lr / (1 - b1^t)
    int n;
    float b1, b2, eps, inv_sqrt_bc2;
};

__global__ void __launch_bounds__(256) adam_multi_kernel(const AdamArgs A)
{
    const long long total = A.end[A.n - 1];
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        int t = 0;
#pragma unroll
        for (int k = 0; k < ADAM_MAX_TENSORS - 1; ++k) t += (k < A.n - 1 && i >= A.end[k]) ? 1 : 0;
        const long long j = i - (t ? A.end[t - 1] : 0);
        const float g = A.g[t][j];
        float m = A.m[t][j], v = A.v[t][j];
        m = fmaf(g - m, 1.0f - A.b1, m);
        v = fmaf(v, A.b2, (1.0f - A.b2) * g * g);
        A.m[t][j] = m;
        A.v[t][j] = v;
        const float denom = sqrtf(v) * A.inv_sqrt_bc2 + A.eps;
        A.p[t][j] = fmaf(-A.step_size[t], m / denom, A.p[t][j]);
    }
}

// params/grads/exp_avg/exp_avg_sq: HOST arrays of n_tensors device pointers; sizes: HOST array of element counts; lrs: HOST array of
// per-tensor learning rates; step = 1, 2, ... (the optimiser step being taken, for the bias corrections).
extern "C" int nlosb200_adam_step(int n_tensors, float* const* params, const float* const* grads, float* const* exp_avg, float* const* exp_avg_sq,
                                  const long long* sizes, const float* lrs, float beta1, float beta2, float eps, int step, void* stream)
{
    NLOS_REQUIRE(n_tensors >= 1 && n_tensors <= ADAM_MAX_TENSORS, "adam_step: n_tensors=%d not in [1, %d]", n_tensors, ADAM_MAX_TENSORS);
    NLOS_REQUIRE(params && grads && exp_avg && exp_avg_sq && sizes && lrs && step >= 1, "adam_step: bad arguments");
    AdamArgs A;
    long long run = 0;
    const double bc1 = 1.0 - pow((double)beta1, step), bc2 = 1.0 - pow((double)beta2, step);
    for (int t = 0; t < ADAM_MAX_TENSORS; ++t) {
        const int s = t < n_tensors ? t : n_tensors - 1;
        NLOS_REQUIRE(sizes[s] >= 0 && (sizes[s] == 0 || (params[s] && grads[s] && exp_avg[s] && exp_avg_sq[s])), "adam_step: tensor %d", s);
        A.p[t] = params[s]; A.g[t] = grads[s]; A.m[t] = exp_avg[s]; A.v[t] = exp_avg_sq[s];
        if (t < n_tensors) run += sizes[t];
        A.end[t] = run;
        A.step_size[t] = (float)((double)lrs[s] / bc1);
    }
    A.n = n_tensors; A.b1 = beta1; A.b2 = beta2; A.eps = eps; A.inv_sqrt_bc2 = (float)(1.0 / sqrt(bc2));
    if (run == 0) return 0;
    const long long nb = (run + 255) / 256;
    adam_multi_kernel<<<(int)(nb < 148 * 16 ? nb : 148 * 16), 256, 0, (cudaStream_t)stream>>>(A);
    NLOS_CHECK_LAUNCH("adam_multi_kernel");
    return 0;
}
