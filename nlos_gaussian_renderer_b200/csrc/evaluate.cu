// evaluate.cu -- dense evaluation of the Gaussian mixture at arbitrary points (volume / mesh evaluation sampling).
//
// The reference evaluates the fitted model on the sample lattice of one scan point with the dense PyTorch path
// (/root/reference/nlos_helpers.py:40-69 gaussian2volume -> GaussianModel.estimate_rho_w(..., out_separately=True),
//  /root/reference/gaussian_model/gaussian_model.py:297-364): [Ng, Na] temporaries of pdf, then sums over the Gaussians:
//     density(x) = sum_g pdf_g(x) sigmoid(o_g)          rho_density(x) = sum_g pdf_g(x) sigmoid(o_g) albedo_g
// (estimate_rho_w_no_occlusion, :346-364).  Here: no temporaries, no culling (the dense sum is over ALL Gaussians, as in the reference),
// pdf with the CUDA renderer's arithmetic  exp(-0.5 |M (x - mu)|^2),  M = diag(1/(s+1e-8)) R^T  (include/cuda_utils.cuh:124-151), from the
// 128-byte records nlosb200_preprocess writes.
//
// One thread owns EV_PPT points in registers; the CTA streams the Gaussians of its y-slice through shared memory (16 floats each, read
// back as broadcast LDS.128); per (Gaussian, point): 3 SUB + 12 FMA/MUL + 1 MUFU.EX2 + 2 FMA.  FP32-pipe bound.
#include "common.cuh"
#include "../../include/nlosb200.h"

using namespace nlos;

#define EV_THREADS 256
#define EV_PPT 4
#define EV_CHUNK 256

__global__ void __launch_bounds__(EV_THREADS) evaluate_points_kernel(const float* __restrict__ points, long long n_points, const float* __restrict__ records,
                                                                      int N, int g_per_slice,
                                                                      float* __restrict__ density, float* __restrict__ rho_density)
{
    __shared__ float4 sg[EV_CHUNK][4];
    const long long p0 = ((long long)blockIdx.x * EV_THREADS + threadIdx.x) * EV_PPT;
    float px[EV_PPT], py[EV_PPT], pz[EV_PPT], dn[EV_PPT], rd[EV_PPT];
#pragma unroll
    for (int k = 0; k < EV_PPT; ++k) {
        const long long p = p0 + k < n_points ? p0 + k : n_points - 1;
        px[k] = points[3 * p]; py[k] = points[3 * p + 1]; pz[k] = points[3 * p + 2];
        dn[k] = 0.f; rd[k] = 0.f;
    }
    const int g_begin = blockIdx.y * g_per_slice;
    const int g_end = min(N, g_begin + g_per_slice);
    const float kk = 0.84932180028801904f;   // sqrt(0.5 log2 e): pdf = 2^-(|kk M (x-mu)|^2)
    for (int c0 = g_begin; c0 < g_end; c0 += EV_CHUNK) {
        const int n = min(EV_CHUNK, g_end - c0);
        __syncthreads();
        for (int i = threadIdx.x; i < n; i += EV_THREADS) {
            const float4* r = reinterpret_cast<const float4*>(records + (size_t)(c0 + i) * NLOS_REC_FLOATS);
            const float4 a = r[0], b = r[1], c = r[2], d = r[3], e = r[4];   // mu op | M0..3 | M4..7 | M8 bmin | bmax alb
            sg[i][0] = make_float4(a.x, a.y, a.z, a.w);
            sg[i][1] = make_float4(kk * b.x, kk * b.y, kk * b.z, a.w * e.w);
            sg[i][2] = make_float4(kk * b.w, kk * c.x, kk * c.y, 0.f);
            sg[i][3] = make_float4(kk * c.z, kk * c.w, kk * d.x, 0.f);
        }
        __syncthreads();
#pragma unroll 2
        for (int i = 0; i < n; ++i) {
            const float4 m = sg[i][0], r0 = sg[i][1], r1 = sg[i][2], r2 = sg[i][3];
#pragma unroll
            for (int k = 0; k < EV_PPT; ++k) {
                const float dx = px[k] - m.x, dy = py[k] - m.y, dz = pz[k] - m.z;
                const float l0 = fmaf(r0.z, dz, fmaf(r0.y, dy, r0.x * dx));
                const float l1 = fmaf(r1.z, dz, fmaf(r1.y, dy, r1.x * dx));
                const float l2 = fmaf(r2.z, dz, fmaf(r2.y, dy, r2.x * dx));
                const float pdf = ex2_approx(-fmaf(l2, l2, fmaf(l1, l1, l0 * l0)));
                dn[k] = fmaf(m.w, pdf, dn[k]);
                rd[k] = fmaf(r0.w, pdf, rd[k]);
            }
        }
    }
#pragma unroll
    for (int k = 0; k < EV_PPT; ++k)
        if (p0 + k < n_points) {
            if (gridDim.y == 1) {
                density[p0 + k] = dn[k];
                rho_density[p0 + k] = rd[k];
            } else {
                atomicAdd(density + p0 + k, dn[k]);
                atomicAdd(rho_density + p0 + k, rd[k]);
            }
        }
}

// ---- records in the DENSE PyTorch semantics (SURVEY.md App. A.1 column a), which is what gaussian2volume actually evaluates in the reference:
//      gaussian_model.py:265      scales = exp(get_scaling * scaling_modifier) = exp(exp(_scaling) * mod)   (double activation)
//      gaussian_model.py:281-282  T = R (x - mu)  (R, not R^T), no epsilon on the scales; build_rotation has no zero-norm guard (gaussian_utils.py:189-210)
//      gaussian_model.py:310,359  albedo = clamp_min(C0 * f_dc + 0.5, 0) at degree 0
// Only the fields nlosb200_evaluate_points reads are meaningful (mean, sigmoid(opacity), M = diag(1/s) R, albedo).
__global__ void __launch_bounds__(256) preprocess_dense_kernel(const float* __restrict__ means, const float* __restrict__ log_scales,
                                                               const float* __restrict__ rots, const float* __restrict__ opac,
                                                               const float* __restrict__ feats, int N, int sh_dim, float mod, float* __restrict__ records)
{
    int g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= N) return;
    float w = rots[4 * g], x = rots[4 * g + 1], y = rots[4 * g + 2], z = rots[4 * g + 3];
    const float inv = 1.0f / sqrtf(w * w + x * x + y * y + z * z);
    w *= inv; x *= inv; y *= inv; z *= inv;
    const float R[9] = {1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y),
                        2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x),
                        2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)};
    float is[3];
    for (int j = 0; j < 3; ++j) is[j] = 1.0f / expf(expf(log_scales[3 * g + j]) * mod);
    const float op = 1.0f / (1.0f + expf(-opac[g]));
    const float alb = fmaxf(fmaf(NLOS_SH_C0, feats[(size_t)g * sh_dim], 0.5f), 0.0f);
    float4* rec = reinterpret_cast<float4*>(records + (size_t)g * NLOS_REC_FLOATS);
    rec[0] = make_float4(means[3 * g], means[3 * g + 1], means[3 * g + 2], op);
    rec[1] = make_float4(R[0] * is[0], R[1] * is[0], R[2] * is[0], R[3] * is[1]);       // row j of M = row j of R / s_j
    rec[2] = make_float4(R[4] * is[1], R[5] * is[1], R[6] * is[2], R[7] * is[2]);
    rec[3] = make_float4(R[8] * is[2], 0.f, 0.f, 0.f);
    rec[4] = make_float4(0.f, 0.f, 0.f, alb);
    rec[5] = make_float4(0.f, 0.f, 0.f, 0.f); rec[6] = rec[5]; rec[7] = rec[5];
}

extern "C" int nlosb200_preprocess_dense(const float* means, const float* log_scales, const float* rotations, const float* opacities,
                                         const float* features, int N, int sh_dim, float scaling_modifier, float* records, void* stream)
{
    if (N <= 0) return 0;
    NLOS_REQUIRE(means && log_scales && rotations && opacities && features && records && sh_dim >= 1, "preprocess_dense: null pointer");
    NLOS_REQUIRE(((uintptr_t)records % 16) == 0, "preprocess_dense: records must be 16-byte aligned");
    preprocess_dense_kernel<<<(N + 255) / 256, 256, 0, (cudaStream_t)stream>>>(means, log_scales, rotations, opacities, features, N, sh_dim,
                                                                               scaling_modifier, records);
    NLOS_CHECK_LAUNCH("preprocess_dense_kernel");
    return 0;
}

// points [n_points,3]; records [N,32] from nlosb200_preprocess (degree-0 albedo) optionally followed by nlosb200_compat_albedo (view-dependent
// albedo of one camera position); density, rho_density [n_points] (overwritten).
extern "C" int nlosb200_evaluate_points(const float* points, long long n_points, const float* records, int N,
                                        float* density, float* rho_density, void* stream)
{
    NLOS_REQUIRE(n_points >= 0 && N >= 0, "evaluate_points: n_points=%lld N=%d", n_points, N);
    if (n_points == 0) return 0;
    NLOS_REQUIRE(points && density && rho_density && (N == 0 || records), "evaluate_points: null pointer");
    cudaStream_t s = (cudaStream_t)stream;
    const long long per_cta = (long long)EV_THREADS * EV_PPT;
    const long long nx = (n_points + per_cta - 1) / per_cta;
    NLOS_REQUIRE(nx <= 0x7fffffffLL, "evaluate_points: too many points");
    // split the Gaussians over gridDim.y until the grid covers about two waves of the 148 SMs (slices of >= EV_CHUNK Gaussians)
    long long ny = (2 * 148 + nx - 1) / nx;
    const long long max_ny = (N + EV_CHUNK - 1) / EV_CHUNK;
    if (ny > max_ny) ny = max_ny;
    if (ny > 65535) ny = 65535;
    if (ny < 1) ny = 1;
    const int g_per_slice = (int)(((N + ny - 1) / ny + EV_CHUNK - 1) / EV_CHUNK * EV_CHUNK);
    ny = N > 0 ? (N + g_per_slice - 1) / g_per_slice : 1;
    if (ny > 1) {
        cudaMemsetAsync(density, 0, sizeof(float) * n_points, s);
        cudaMemsetAsync(rho_density, 0, sizeof(float) * n_points, s);
    }
    evaluate_points_kernel<<<dim3((unsigned)nx, (unsigned)ny), EV_THREADS, 0, s>>>(points, n_points, records, N, g_per_slice > 0 ? g_per_slice : EV_CHUNK,
                                                                                   density, rho_density);
    NLOS_CHECK_LAUNCH("evaluate_points_kernel");
    return 0;
}
