// rayquad.cuh -- per-(Gaussian, ray) set-up shared by the one-scan-point operators (compat.cu) and the batched occlusion renderer (occluded.cu)
#pragma once
#include "common.cuh"
#ifndef LOG2E_F
#define LOG2E_F 1.4426950408889634f
#endif
using namespace nlos;

// per-(Gaussian, ray) set-up shared by the forward and backward kernels: centred quadratic of
// m2(t) = |M (o + t d - mu)|^2 = m2min + C (t - rs)^2      (cuda_utils.cuh:124-151 in factored form)
struct RayQuad {
    float rs, gam, al, op, alb;      // gam = -0.5 log2(e) C ; al = -0.5 log2(e) m2min
};
__device__ __forceinline__ void ray_quad(const float* __restrict__ rec, float ox, float oy, float oz, float dx, float dy, float dz,
                                         RayQuad& Q, float* a3 = nullptr, float* b3 = nullptr, float* p3 = nullptr)
{
    const float4* r4 = reinterpret_cast<const float4*>(rec);
    const float4 v0 = r4[0], v1 = r4[1], v2 = r4[2], v3 = r4[3], v4 = r4[4];
    const float cx = ox - v0.x, cy = oy - v0.y, cz = oz - v0.z;
    const float ax = v1.x * cx + v1.y * cy + v1.z * cz, ay = v1.w * cx + v2.x * cy + v2.y * cz, az = v2.z * cx + v2.w * cy + v3.x * cz;
    const float bx = v1.x * dx + v1.y * dy + v1.z * dz, by = v1.w * dx + v2.x * dy + v2.y * dz, bz = v2.z * dx + v2.w * dy + v3.x * dz;
    const float C = bx * bx + by * by + bz * bz;
    const float rs = C > 0.f ? -(ax * bx + ay * by + az * bz) / C : 0.f;
    const float px = fmaf(rs, bx, ax), py = fmaf(rs, by, ay), pz = fmaf(rs, bz, az);
    Q.rs = rs;
    Q.gam = -0.5f * LOG2E_F * C;
    Q.al = -0.5f * LOG2E_F * (px * px + py * py + pz * pz);
    Q.op = v0.w;
    Q.alb = v4.w;
    if (a3) { a3[0] = ax; a3[1] = ay; a3[2] = az; }
    if (b3) { b3[0] = bx; b3[1] = by; b3[2] = bz; }
    if (p3) { p3[0] = px; p3[1] = py; p3[2] = pz; }
}

// 1 - exp(-x) without cancellation for small x (the reference computes 1.0f - expf(-x), volume_renderer.cu:115)
__device__ __forceinline__ float one_minus_exp_neg(float x)
{
    if (x < 0.03f) return x * (1.0f - 0.5f * x * (1.0f - (1.0f / 3.0f) * x * (1.0f - 0.25f * x)));
    return 1.0f - ex2_approx(-LOG2E_F * x);
}

