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

// ==== MCMC relocation: opacity / scale of a Gaussian that is replaced by N copies of itself ============================================
// GaussianModel._update_params (/root/reference/gaussian_model/gaussian_model.py:471-483) calls compute_relocation_cuda(opacity_old,
// scale_old, N), a name the reference never defines or imports ("broken"): it is the helper of the public
// 3DGS-MCMC code (Kheradmand et al., "3D Gaussian Splatting as Markov Chain Monte Carlo", NeurIPS 2024, eq. 9), which is not vendored
// under /root/reference.  PARITY UNPINNED: restated from the publication --
//     o_new = 1 - (1 - o_old)^(1/N)
//     s_new = s_old * o_old / sum_{i=1..N} sum_{k=0..i-1} C(i-1, k) (-1)^k o_new^(k+1) / sqrt(k+1)
// with the counts clamped to [1, n_max - 1] (n_max = 51 upstream) and the binomial table C(n, k) = binoms[n * n_max + k] passed in.
// The summation keeps the published order (i outer, k inner, float accumulator) because the series alternates with binomial-sized
// terms: its float rounding IS the function for large N.  The k-dependent factor is formed once per k (double, as an int-argument
// pow / sqrt evaluates upstream) instead of once per (i, k): O(N) transcendental work, O(N^2 / 2) fused multiply-adds.
#define RELOC_MAX_N 64
__global__ void __launch_bounds__(128) relocation_kernel(const float* __restrict__ opacity_old, const float* __restrict__ scale_old,
                                                         const int* __restrict__ counts, const float* __restrict__ binoms, int n_max, int P,
                                                         float* __restrict__ opacity_new, float* __restrict__ scale_new)
{
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= P) return;
    const int N = min(max(counts[idx], 1), n_max - 1);
    const float o_old = opacity_old[idx];
    const float o_new = 1.0f - powf(1.0f - o_old, 1.0f / (float)N);
    opacity_new[idx] = o_new;
    double term[RELOC_MAX_N];
    double pw = (double)o_new;                               // o_new^(k+1)
    for (int k = 0; k < N; ++k) {
        term[k] = ((k & 1) ? -1.0 : 1.0) / sqrt((double)(k + 1)) * pw;
        pw *= (double)o_new;
    }
    float denom = 0.f;
    for (int i = 1; i <= N; ++i)
        for (int k = 0; k < i; ++k) denom = (float)((double)denom + (double)binoms[(size_t)(i - 1) * n_max + k] * term[k]);
    const float coeff = o_old / denom;
#pragma unroll
    for (int a = 0; a < 3; ++a) scale_new[(size_t)idx * 3 + a] = coeff * scale_old[(size_t)idx * 3 + a];
}

extern "C" int nlosb200_compute_relocation(const float* opacity_old, const float* scale_old, const int* counts, const float* binoms, int n_max,
                                           int P, float* opacity_new, float* scale_new, void* stream)
{
    NLOS_REQUIRE(P >= 0 && n_max >= 2 && n_max <= RELOC_MAX_N, "compute_relocation: n_max=%d not in [2, %d]", n_max, RELOC_MAX_N);
    if (P == 0) return 0;
    NLOS_REQUIRE(opacity_old && scale_old && counts && binoms && opacity_new && scale_new, "compute_relocation: null pointer");
    relocation_kernel<<<(P + 127) / 128, 128, 0, (cudaStream_t)stream>>>(opacity_old, scale_old, counts, binoms, n_max, P, opacity_new, scale_new);
    NLOS_CHECK_LAUNCH("relocation_kernel");
    return 0;
}
