// compat.cu -- the reference's three pybind entry points, re-designed for sm_100a behind the C ABI.
//
// Replaces (reference, /root/reference/submodules/cuda_renderer, bound in src/bindings.cpp:26-35):
//   filter_gaussians_per_ray   src/ray_aabb.cu:10-102              (1 thread / ray, serial over all Gaussians)
//   render_rays                src/volume_renderer.cu:16-305       (1 thread / ray, re-activates every Gaussian per sample)
//   render_rays_analytic       src/volume_renderer_analytic.cu:23-241 (1 thread / ray, 2 KB local-memory section stack)
// and adds the backward of render_rays that gaussian_model/cuda_autograd.py:139-172 leaves as zeros.
//
// These are the ONE-scan-point-per-call operators that CUDARenderFunction / NLOSGaussianRenderer /
// SectionGaussianRendererCUDA drive; the batched [scan_y, scan_x, time_bin] renderer lives in transient.cu.
//
// B200 design (instead of 1 thread per ray => 4 CTAs for a 32x32 lattice):
//   * one WARP per ray, 8 rays per CTA => 128 CTAs for 1024 rays;
//   * filter: the [N,6] AABB array streams through shared memory in 896-box stages with TMA bulk copies
//     (2-stage mbarrier ring shared by the CTA's 8 rays); each lane slab-tests one box, hits are
//     ballot-ranked so the list is ascending and cut at `cap` exactly like the reference's serial loop;
//   * render: lanes split the ray's list for the per-(Gaussian, ray) set-up (a = M(o-mu), b = M d, the
//     centred quadratic m2(t) = m2min + C (t - r*)^2), broadcast it with shuffles, and then own every 32nd
//     sample: 2 FMA + 1 MUFU.EX2 per sample-eval, accumulated in per-warp shared-memory rows; the
//     transmittance product runs in the reference's sample order; outputs are written coalesced;
//   * analytic: lanes compute ray/ellipsoid sections in parallel, ballot-compact the first 128, rank-sort
//     them (stable, == the reference's insertion sort), evaluate tau in parallel and composite in order.
#include "common.cuh"
#include "rayquad.cuh"
#include "../../include/nlosb200.h"

using namespace nlos;

#define CP_WPB 8                 // rays (warps) per CTA
#define CP_FSTAGE 896            // AABBs per TMA stage of the filter kernel (2 stages = 42 KB static shared memory)
#define CP_MAX_SECTIONS 128      // MAX_SECTIONS_PER_RAY, volume_renderer_analytic.cu:11

// ------------------------------------------------------------------------------------------------
// filter_gaussians_per_ray : out int32 [n_rays, cap+1]  (col 0 = count, then ascending ids, unused = -1)
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(CP_WPB * 32) compat_filter_kernel(const float* __restrict__ ray_o, const float* __restrict__ ray_d,
                                                                     const float* __restrict__ bboxes, int n_rays, int N, int cap,
                                                                     int* __restrict__ out)
{
    __shared__ __align__(128) float s_box[2][CP_FSTAGE * 6];
    __shared__ uint64_t s_bar[2];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int ray = blockIdx.x * CP_WPB + warp;
    const bool live = ray < n_rays;
    if (threadIdx.x == 0) {
        mbar_init(&s_bar[0], 1);
        mbar_init(&s_bar[1], 1);
        mbar_fence_init();
    }
    __syncthreads();
    float ox = 0, oy = 0, oz = 0, dx = 0, dy = 0, dz = 0;
    if (live) {
        ox = ray_o[3 * ray + 0]; oy = ray_o[3 * ray + 1]; oz = ray_o[3 * ray + 2];
        dx = ray_d[3 * ray + 0]; dy = ray_d[3 * ray + 1]; dz = ray_d[3 * ray + 2];
    }
    int* row = out + (size_t)(live ? ray : 0) * (cap + 1);
    int count = 0;
    bool full = !live || cap <= 0;
    const int n_stages = (N + CP_FSTAGE - 1) / CP_FSTAGE;
    auto issue = [&](int st) {
        int cnt = min(CP_FSTAGE, N - st * CP_FSTAGE);
        uint32_t bytes = (uint32_t)cnt * 24u;          // 6 floats per box; 24*cnt is a multiple of 16 unless cnt is odd
        uint32_t bulk = bytes & ~15u;
        mbar_arrive_expect_tx(&s_bar[st & 1], bulk);
        if (bulk) bulk_g2s(s_box[st & 1], bboxes + (size_t)st * CP_FSTAGE * 6, bulk, &s_bar[st & 1]);
    };
    if (threadIdx.x == 0 && n_stages > 0) issue(0);
    int st = 0;
    for (; st < n_stages; ++st) {
        const int cnt = min(CP_FSTAGE, N - st * CP_FSTAGE);
        if (threadIdx.x == 0 && st + 1 < n_stages) {
            fence_proxy_async();                       // generic-proxy reads of that buffer (stage st-1) precede the async write
            issue(st + 1);
        }
        mbar_wait(&s_bar[st & 1], (st >> 1) & 1);
        // tail of an odd-sized last stage (8 bytes) is not 16-byte granular: fetch it with plain loads
        if ((cnt * 24) & 15) {
            if (threadIdx.x < 2) s_box[st & 1][cnt * 6 - 2 + threadIdx.x] = bboxes[((size_t)st * CP_FSTAGE + cnt) * 6 - 2 + threadIdx.x];
            __syncthreads();
        }
        if (!full) {
            const float* sb = s_box[st & 1];
            for (int base = 0; base < cnt && !full; base += 32) {
                int gl = base + lane;
                bool hit = false;
                if (gl < cnt) hit = ray_aabb_x(ox, oy, oz, dx, dy, dz, sb + gl * 6);
                unsigned b = __ballot_sync(0xffffffffu, hit);
                int pos = count + __popc(b & ((1u << lane) - 1u));
                if (hit && pos < cap) row[1 + pos] = st * CP_FSTAGE + gl;
                count += __popc(b);
                if (count >= cap) { count = cap; full = true; }
            }
        }
        // every warp must be done with this stage before it is refilled two iterations later
        if (__syncthreads_and(full)) break;          // all 8 rays of this CTA are at their cap: stop streaming
    }
    // a prefetch may still be in flight after the early exit: never leave the CTA with an outstanding bulk copy
    if (st + 1 < n_stages) mbar_wait(&s_bar[(st + 1) & 1], ((st + 1) >> 1) & 1);
    if (live) {
        for (int i = count + lane; i < cap; i += 32) row[1 + i] = -1;
        if (lane == 0) row[0] = count;
    }
}

extern "C" int nlosb200_filter_gaussians_per_ray(const float* ray_origins, const float* ray_directions, const float* bboxes, int n_rays,
                                                 int N, int cap, int* out, void* stream)
{
    if (n_rays <= 0) return 0;
    NLOS_REQUIRE(cap >= 0 && N >= 0, "filter_gaussians_per_ray: bad sizes");
    NLOS_REQUIRE(ray_origins && ray_directions && out && (bboxes || N == 0), "filter_gaussians_per_ray: null pointer");
    NLOS_REQUIRE(((uintptr_t)bboxes % 16) == 0, "filter_gaussians_per_ray: bboxes must be 16-byte aligned");
    compat_filter_kernel<<<(n_rays + CP_WPB - 1) / CP_WPB, CP_WPB * 32, 0, (cudaStream_t)stream>>>(ray_origins, ray_directions, bboxes, n_rays,
                                                                                                   N, cap, out);
    NLOS_CHECK_LAUNCH("compat_filter_kernel");
    return 0;
}

// ------------------------------------------------------------------------------------------------
// view-dependent albedo of one scan point ("camera"): rho_g = max(SH(dir(mu_g - cam)) + 0.5, 0)
// (volume_renderer.cu:107-111, spherical_harmonics.cuh:20-80 -- the all-positive sign convention of the
// CUDA reference, NOT the PlenOctree signs of gaussian_model/sh_utils.py).  Writes REC_ALB of each record.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) compat_albedo_kernel(const float* __restrict__ means, const float* __restrict__ feats,
                                                            const float* __restrict__ cam, int N, int sh_dim, int degree,
                                                            float* __restrict__ records)
{
    int g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= N) return;
    float vx = means[3 * g + 0] - cam[0], vy = means[3 * g + 1] - cam[1], vz = means[3 * g + 2] - cam[2];
    float il = 1.0f / (sqrtf(vx * vx + vy * vy + vz * vz) + 1e-8f);          // cuda_utils.cuh:48-51
    vx *= il; vy *= il; vz *= il;
    float Y[16];
    sh_basis_all(degree, vx, vy, vz, Y);
    int K = (degree + 1) * (degree + 1);
    float r = 0.f;
    for (int i = 0; i < K; ++i) r = fmaf(feats[(size_t)g * sh_dim + i], Y[i], r);
    records[(size_t)g * NLOS_REC_FLOATS + REC_ALB] = fmaxf(r + 0.5f, 0.0f);
}

// backward of compat_albedo_kernel (the reference has none): dalb = dacc[g][13] = dL/d(albedo_g) -> d_features[g, :] (overwritten) and
// d_means[g] += the dependence of the view direction on the mean
__global__ void __launch_bounds__(256) compat_albedo_backward_kernel(const float* __restrict__ means, const float* __restrict__ feats,
                                                                     const float* __restrict__ cam, int N, int sh_dim, int degree,
                                                                     const float* __restrict__ dacc, float* __restrict__ d_means,
                                                                     float* __restrict__ d_feats)
{
    int g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= N) return;
    const float cx = means[3 * g + 0] - cam[0], cy = means[3 * g + 1] - cam[1], cz = means[3 * g + 2] - cam[2];
    const float D = sqrtf(cx * cx + cy * cy + cz * cz);
    const float* f = feats + (size_t)g * sh_dim;
    const float alb = sh_pair_albedo(f, degree, cx, cy, cz, D);
    const float dpre = alb > 0.f ? dacc[(size_t)g * NLOS_GRAD_FLOATS + 13] : 0.f;
    float dsh[16], gx = 0.f, gy = 0.f, gz = 0.f;
    sh_pair_backward(f, degree, cx, cy, cz, D, dpre, dsh, gx, gy, gz);
    const int K = (degree + 1) * (degree + 1);
    for (int k = 0; k < sh_dim; ++k) d_feats[(size_t)g * sh_dim + k] = k < K ? dsh[k] : 0.f;
    d_means[3 * g + 0] += gx; d_means[3 * g + 1] += gy; d_means[3 * g + 2] += gz;
}

// ------------------------------------------------------------------------------------------------
// render_rays forward: rho_density, density, transmittance  [n_rays, T]   (volume_renderer.cu:16-185)
// dynamic shared memory: [wpb][3][T] floats (density, weighted sum, transmittance) + [T] sample radii
// ------------------------------------------------------------------------------------------------
template <bool OCC>
__global__ void __launch_bounds__(CP_WPB * 32) compat_render_kernel(const float* __restrict__ ray_o, const float* __restrict__ ray_d,
                                                                     const float* __restrict__ t_samples, const float* __restrict__ records,
                                                                     const int* __restrict__ filter, int cap, int n_rays, int T, int N,
                                                                     float c, float deltaT, float* __restrict__ rho,
                                                                     float* __restrict__ density, float* __restrict__ trans)
{
    extern __shared__ __align__(16) float cp_smem[];
    const int wpb = blockDim.x >> 5;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float* s_t = cp_smem;
    float* s_den = cp_smem + T + (size_t)warp * 3 * T;
    float* s_ws = s_den + T;
    float* s_tr = s_ws + T;
    for (int s = threadIdx.x; s < T; s += blockDim.x) s_t[s] = t_samples[s];
    const int ray = blockIdx.x * wpb + warp;
    const bool live = ray < n_rays;
    for (int s = lane; s < T; s += 32) { s_den[s] = 0.f; s_ws[s] = 0.f; }
    __syncthreads();
    if (!live) return;
    const float ox = ray_o[3 * ray + 0], oy = ray_o[3 * ray + 1], oz = ray_o[3 * ray + 2];
    const float dx = ray_d[3 * ray + 0], dy = ray_d[3 * ray + 1], dz = ray_d[3 * ray + 2];
    const int* row = filter + (size_t)ray * (cap + 1);
    const int n = min(row[0], cap);
    const float cdt = c * deltaT;
    for (int base = 0; base < n; base += 32) {
        RayQuad Q = {0.f, 0.f, 0.f, 0.f, 0.f};
        const int m = min(32, n - base);
        if (lane < m) {
            int g = row[1 + base + lane];
            if (g >= 0 && g < N) ray_quad(records + (size_t)g * NLOS_REC_FLOATS, ox, oy, oz, dx, dy, dz, Q);
        }
        for (int q = 0; q < m; ++q) {
            const float rs = __shfl_sync(0xffffffffu, Q.rs, q), gam = __shfl_sync(0xffffffffu, Q.gam, q);
            const float al = __shfl_sync(0xffffffffu, Q.al, q), op = __shfl_sync(0xffffffffu, Q.op, q);
            const float alb = __shfl_sync(0xffffffffu, Q.alb, q);
            if (op == 0.f) continue;
#pragma unroll 4
            for (int s = lane; s < T; s += 32) {
                const float t = s_t[s] - rs;
                const float v = op * ex2_approx(fmaf(gam, t * t, al));        // pdf * sigmoid(opacity)
                s_den[s] += v;
                s_ws[s] += OCC ? one_minus_exp_neg(v * cdt) * alb : v * alb;     // :114-117 / :174-176
            }
        }
    }
    __syncwarp();
    float* o_rho = rho + (size_t)ray * T;
    float* o_den = density + (size_t)ray * T;
    float* o_tr = trans + (size_t)ray * T;
    if (!OCC) {
        for (int s = lane; s < T; s += 32) {
            o_rho[s] = s_ws[s] * c * deltaT;                                   // :182
            o_den[s] = s_den[s];
            o_tr[s] = 1.0f;
        }
        return;
    }
    // occlusion: T_0 = 1 ; T_{s+1} = T_s * expf(-density_s * c * deltaT) in sample order, stop when T < 1e-4 (:120-137)
    for (int s = lane; s < T; s += 32) s_tr[s] = expf(-s_den[s] * c * deltaT);
    __syncwarp();
    int alive = T;
    if (lane == 0) {
        float Tr = 1.0f;
        for (int s = 0; s < T; ++s) {
            const float st = s_tr[s];
            s_tr[s] = Tr;
            Tr *= st;
            if (Tr < 1e-4f) { alive = s + 1; break; }
        }
    }
    alive = __shfl_sync(0xffffffffu, alive, 0);
    __syncwarp();
    for (int s = lane; s < T; s += 32) {
        const bool a = s < alive;
        const float tr = a ? s_tr[s] : 0.f;
        o_rho[s] = a ? tr * s_ws[s] : 0.f;
        o_den[s] = a ? s_den[s] : 0.f;
        o_tr[s] = tr;
    }
}

// ------------------------------------------------------------------------------------------------
// render_rays backward (what gaussian_model/cuda_autograd.py:139-172 returns zeros for).
//   inputs : saved rho / transmittance of the forward, the three output gradients [n_rays, T]
//   output : dacc [N,16] += { dL/dM (9), dL/dmu (3), dL/dop, dL/dalbedo, -, - }   (REDs; caller zeroes)
// Per ray the output gradients are first folded into two per-sample weights
//   gD[s] = dL/d density_s  (incl. the effect of density_s on every later transmittance)
//   gW[s] = dL/d (weighted sum)_s
// then every listed Gaussian is re-evaluated: we_s = dL/dv_s * e_s, three moments S0,S1,S2 of (t - r*)
// reduce over the warp with shuffles, and ONE lane per Gaussian applies the chain rule and flushes.
// dynamic shared memory: [T] radii + [wpb][3][T].
// ------------------------------------------------------------------------------------------------
template <bool OCC>
__global__ void __launch_bounds__(CP_WPB * 32) compat_render_backward_kernel(
    const float* __restrict__ ray_o, const float* __restrict__ ray_d, const float* __restrict__ t_samples,
    const float* __restrict__ records, const int* __restrict__ filter, int cap, int n_rays, int T, int N, float c, float deltaT,
    const float* __restrict__ rho, const float* __restrict__ trans, const float* __restrict__ g_rho, const float* __restrict__ g_den,
    const float* __restrict__ g_tr, float* __restrict__ dacc)
{
    extern __shared__ __align__(16) float cp_smem[];
    const int wpb = blockDim.x >> 5;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float* s_t = cp_smem;
    float* s_gd = cp_smem + T + (size_t)warp * 3 * T;
    float* s_gw = s_gd + T;
    float* s_x = s_gw + T;
    for (int s = threadIdx.x; s < T; s += blockDim.x) s_t[s] = t_samples[s];
    __syncthreads();
    const int ray = blockIdx.x * wpb + warp;
    if (ray >= n_rays) return;
    const float cdt = c * deltaT;
    const size_t ro = (size_t)ray * T;
    if (!OCC) {
        for (int s = lane; s < T; s += 32) {
            s_gd[s] = g_den ? g_den[ro + s] : 0.f;
            s_gw[s] = g_rho ? g_rho[ro + s] * c * deltaT : 0.f;
        }
    } else {
        for (int s = lane; s < T; s += 32) {
            const float tr = trans[ro + s];
            const bool a = tr > 0.f;                      // zero-filled samples after the early stop carry no gradient
            const float gr = g_rho ? g_rho[ro + s] : 0.f;
            s_gd[s] = (a && g_den) ? g_den[ro + s] : 0.f;
            s_gw[s] = a ? gr * tr : 0.f;
            s_x[s] = a ? fmaf(g_tr ? g_tr[ro + s] : 0.f, tr, gr * rho[ro + s]) : 0.f;      // dL/dT_s * T_s
        }
        __syncwarp();
        // dT_s/d density_s' = -c dT T_s for s > s'  =>  gD[s'] -= c dT * sum_{s > s'} dL/dT_s T_s   (warp suffix scan)
        float carry = 0.f;
        for (int ch = (T - 1) / 32; ch >= 0; --ch) {
            const int s = ch * 32 + lane;
            const float own = s < T ? s_x[s] : 0.f;
            float x = own;
#pragma unroll
            for (int off = 1; off < 32; off <<= 1) {
                float y = __shfl_down_sync(0xffffffffu, x, off);
                if (lane + off < 32) x += y;
            }
            if (s < T && trans[ro + s] > 0.f) s_gd[s] -= cdt * (x - own + carry);
            carry += __shfl_sync(0xffffffffu, x, 0);
        }
    }
    __syncwarp();
    const float ox = ray_o[3 * ray + 0], oy = ray_o[3 * ray + 1], oz = ray_o[3 * ray + 2];
    const float dx = ray_d[3 * ray + 0], dy = ray_d[3 * ray + 1], dz = ray_d[3 * ray + 2];
    const int* row = filter + (size_t)ray * (cap + 1);
    const int n = min(row[0], cap);
    for (int base = 0; base < n; base += 32) {
        RayQuad Q = {0.f, 0.f, 0.f, 0.f, 0.f};
        float a3[3] = {0, 0, 0}, b3[3] = {0, 0, 0}, p3[3] = {0, 0, 0};
        const int m = min(32, n - base);
        int g = -1;
        if (lane < m) {
            g = row[1 + base + lane];
            if (g >= 0 && g < N) ray_quad(records + (size_t)g * NLOS_REC_FLOATS, ox, oy, oz, dx, dy, dz, Q, a3, b3, p3);
            else g = -1;
        }
        float m0 = 0.f, m1 = 0.f, m2 = 0.f, ma = 0.f;      // this lane's Gaussian: S0, S1, S2, dL/dalbedo
        for (int q = 0; q < m; ++q) {
            const float rs = __shfl_sync(0xffffffffu, Q.rs, q), gam = __shfl_sync(0xffffffffu, Q.gam, q);
            const float al = __shfl_sync(0xffffffffu, Q.al, q), op = __shfl_sync(0xffffffffu, Q.op, q);
            const float alb = __shfl_sync(0xffffffffu, Q.alb, q);
            if (op == 0.f) continue;
            float S0 = 0.f, S1 = 0.f, S2 = 0.f, Sa = 0.f;
#pragma unroll 2
            for (int s = lane; s < T; s += 32) {
                const float t = s_t[s] - rs;
                const float e = ex2_approx(fmaf(gam, t * t, al));
                const float v = op * e;
                float dv;
                if (OCC) {
                    const float ex = ex2_approx(-LOG2E_F * v * cdt);
                    dv = fmaf(s_gw[s] * alb * cdt, ex, s_gd[s]);
                    Sa = fmaf(s_gw[s], one_minus_exp_neg(v * cdt), Sa);
                } else {
                    dv = fmaf(s_gw[s], alb, s_gd[s]);
                    Sa = fmaf(s_gw[s], v, Sa);
                }
                const float we = dv * e, wet = we * t;
                S0 += we;
                S1 += wet;
                S2 = fmaf(wet, t, S2);
            }
            S0 = warp_sum(S0); S1 = warp_sum(S1); S2 = warp_sum(S2); Sa = warp_sum(Sa);
            if (lane == q) { m0 = S0; m1 = S1; m2 = S2; ma = Sa; }
        }
        if (g >= 0) {
            const float4* r4 = reinterpret_cast<const float4*>(records + (size_t)g * NLOS_REC_FLOATS);
            const float4 v0 = r4[0], v1 = r4[1], v2 = r4[2], v3 = r4[3];
            const float op = Q.op, rs = Q.rs;
            const float ux = m0 * p3[0] + m1 * b3[0], uy = m0 * p3[1] + m1 * b3[1], uz = m0 * p3[2] + m1 * b3[2];
            const float gax = -op * ux, gay = -op * uy, gaz = -op * uz;                                     // dL/da
            const float gbx = -op * fmaf(rs, ux, m1 * p3[0] + m2 * b3[0]), gby = -op * fmaf(rs, uy, m1 * p3[1] + m2 * b3[1]),
                        gbz = -op * fmaf(rs, uz, m1 * p3[2] + m2 * b3[2]);                                  // dL/db
            const float cx = ox - v0.x, cy = oy - v0.y, cz = oz - v0.z;                                     // a = M (o - mu), b = M d
            float o[14];
            o[0] = fmaf(gax, cx, gbx * dx); o[1] = fmaf(gax, cy, gbx * dy); o[2] = fmaf(gax, cz, gbx * dz);
            o[3] = fmaf(gay, cx, gby * dx); o[4] = fmaf(gay, cy, gby * dy); o[5] = fmaf(gay, cz, gby * dz);
            o[6] = fmaf(gaz, cx, gbz * dx); o[7] = fmaf(gaz, cy, gbz * dy); o[8] = fmaf(gaz, cz, gbz * dz);
            o[9] = -(v1.x * gax + v1.w * gay + v2.z * gaz);                                                 // dL/dmu = -M^T dL/da
            o[10] = -(v1.y * gax + v2.x * gay + v2.w * gaz);
            o[11] = -(v1.z * gax + v2.y * gay + v3.x * gaz);
            o[12] = m0;                                                                                     // dL/dop
            o[13] = ma;                                                                                     // dL/dalbedo
            float* dst = dacc + (size_t)g * NLOS_GRAD_FLOATS;
#pragma unroll
            for (int i = 0; i < 14; ++i)
                if (o[i] != 0.f) atomicAdd(dst + i, o[i]);
        }
    }
}

// ------------------------------------------------------------------------------------------------
// render_rays_analytic (reference-faithful "section" renderer): one scalar per ray
//   volume_renderer_analytic.cu:23-173, analytic_integration.cuh:38-104 (section), :123-172 (tau), :178-192 (sort)
// ------------------------------------------------------------------------------------------------
struct SecCoef { float A, B, Cc, G, alb; };     // A = |w/s|^2, B = 2 (v/s).(w/s), Cc = |v/s|^2, G = op * sx*sy*sz

__device__ __forceinline__ void section_setup(int g, const float* __restrict__ means, const float* __restrict__ log_scales,
                                              const float* __restrict__ rots, const float* __restrict__ opac,
                                              const float* __restrict__ feats, const float* __restrict__ cam, int sh_dim, int degree,
                                              float mod, float ox, float oy, float oz, float dx, float dy, float dz, SecCoef& S)
{
    float R[9];
    {
        float w = rots[4 * g + 0], x = rots[4 * g + 1], y = rots[4 * g + 2], z = rots[4 * g + 3];
        float norm = sqrtf(w * w + x * x + y * y + z * z);
        if (norm < 1e-8f) {
            R[0] = 1; R[1] = 0; R[2] = 0; R[3] = 0; R[4] = 1; R[5] = 0; R[6] = 0; R[7] = 0; R[8] = 1;
        } else {
            w /= norm; x /= norm; y /= norm; z /= norm;
            R[0] = 1.f - 2.f * (y * y + z * z); R[1] = 2.f * (x * y - w * z); R[2] = 2.f * (x * z + w * y);
            R[3] = 2.f * (x * y + w * z); R[4] = 1.f - 2.f * (x * x + z * z); R[5] = 2.f * (y * z - w * x);
            R[6] = 2.f * (x * z - w * y); R[7] = 2.f * (y * z + w * x); R[8] = 1.f - 2.f * (x * x + y * y);
        }
    }
    const float sx = expf(log_scales[3 * g + 0]) * mod, sy = expf(log_scales[3 * g + 1]) * mod, sz = expf(log_scales[3 * g + 2]) * mod;
    const float mx = means[3 * g + 0], my = means[3 * g + 1], mz = means[3 * g + 2];
    const float ex = ox - mx, ey = oy - my, ez = oz - mz;
    const float vx = (R[0] * ex + R[3] * ey + R[6] * ez) / sx, vy = (R[1] * ex + R[4] * ey + R[7] * ez) / sy,
                vz = (R[2] * ex + R[5] * ey + R[8] * ez) / sz;
    const float wx = (R[0] * dx + R[3] * dy + R[6] * dz) / sx, wy = (R[1] * dx + R[4] * dy + R[7] * dz) / sy,
                wz = (R[2] * dx + R[5] * dy + R[8] * dz) / sz;
    S.A = wx * wx + wy * wy + wz * wz;
    S.B = 2.0f * (vx * wx + vy * wy + vz * wz);
    S.Cc = vx * vx + vy * vy + vz * vz;
    S.G = (1.0f / (1.0f + expf(-opac[g]))) * sx * sy * sz;
    float ux = mx - cam[0], uy = my - cam[1], uz = mz - cam[2];
    float il = 1.0f / (sqrtf(ux * ux + uy * uy + uz * uz) + 1e-8f);
    float Y[16];
    sh_basis_all(degree, ux * il, uy * il, uz * il, Y);
    float r = 0.f;
    const int K = (degree + 1) * (degree + 1);
    for (int i = 0; i < K; ++i) r = fmaf(feats[(size_t)g * sh_dim + i], Y[i], r);
    S.alb = fmaxf(r + 0.5f, 0.0f);
}

__global__ void __launch_bounds__(CP_WPB * 32) compat_analytic_kernel(
    const float* __restrict__ ray_o, const float* __restrict__ ray_d, float t_min, float t_max, const int* __restrict__ filter, int cap,
    const float* __restrict__ means, const float* __restrict__ log_scales, const float* __restrict__ rots, const float* __restrict__ opac,
    const float* __restrict__ feats, const float* __restrict__ cam, int n_rays, int N, int sh_dim, int degree, float mod, float sigma,
    int max_sections, float* __restrict__ out)
{
    __shared__ float s_te[CP_WPB][CP_MAX_SECTIONS], s_tx[CP_WPB][CP_MAX_SECTIONS];
    __shared__ float s_A[CP_WPB][CP_MAX_SECTIONS], s_B[CP_WPB][CP_MAX_SECTIONS], s_C[CP_WPB][CP_MAX_SECTIONS];
    __shared__ float s_G[CP_WPB][CP_MAX_SECTIONS], s_alb[CP_WPB][CP_MAX_SECTIONS];
    __shared__ float s_st[CP_WPB][CP_MAX_SECTIONS], s_w[CP_WPB][CP_MAX_SECTIONS];
    __shared__ int s_ord[CP_WPB][CP_MAX_SECTIONS];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int ray = blockIdx.x * CP_WPB + warp;
    if (ray >= n_rays) return;
    const float ox = ray_o[3 * ray + 0], oy = ray_o[3 * ray + 1], oz = ray_o[3 * ray + 2];
    const float dx = ray_d[3 * ray + 0], dy = ray_d[3 * ray + 1], dz = ray_d[3 * ray + 2];
    const int* row = filter + (size_t)ray * (cap + 1);
    const int n = min(row[0], cap);
    int ns = 0;
    const unsigned lt = (1u << lane) - 1u;
    for (int base = 0; base < n && ns < max_sections; base += 32) {
        bool ok = false;
        SecCoef S;
        float te = 0.f, tx = 0.f;
        if (base + lane < n) {
            const int g = row[1 + base + lane];
            if (g >= 0 && g < N) {
                section_setup(g, means, log_scales, rots, opac, feats, cam, sh_dim, degree, mod, ox, oy, oz, dx, dy, dz, S);
                // |v + t w|^2 = sigma^2  (analytic_integration.cuh:86-103)
                const float qc = S.Cc - sigma * sigma;
                const float disc = S.B * S.B - 4.0f * S.A * qc;
                if (disc >= 0.f) {
                    const float sq = sqrtf(disc);
                    te = fmaxf((-S.B - sq) / (2.0f * S.A), t_min);
                    tx = fminf((-S.B + sq) / (2.0f * S.A), t_max);
                    ok = te < tx;
                }
            }
        }
        const unsigned b = __ballot_sync(0xffffffffu, ok);
        const int pos = ns + __popc(b & lt);
        if (ok && pos < max_sections) {
            s_te[warp][pos] = te; s_tx[warp][pos] = tx;
            s_A[warp][pos] = S.A; s_B[warp][pos] = S.B; s_C[warp][pos] = S.Cc; s_G[warp][pos] = S.G; s_alb[warp][pos] = S.alb;
        }
        ns = min(max_sections, ns + __popc(b));
    }
    __syncwarp();
    // stable rank sort by t_enter == the reference's insertion sort (analytic_integration.cuh:178-192)
    for (int e = lane; e < ns; e += 32) {
        const float key = s_te[warp][e];
        int rank = 0;
        for (int f = 0; f < ns; ++f) {
            const float kf = s_te[warp][f];
            rank += (kf < key || (kf == key && f < e)) ? 1 : 0;
        }
        s_ord[warp][rank] = e;
    }
    __syncwarp();
    for (int r = lane; r < ns; r += 32) {
        const int e = s_ord[warp][r];
        const float a = s_C[warp][e], bq = s_B[warp][e], cq = s_A[warp][e];
        const float t0 = s_te[warp][e], t1 = s_tx[warp][e];
        const float Gf = s_G[warp][e] * sqrtf(6.283185307179586f / cq);
        const float ef = expf(-0.5f * (a - bq * bq / (4.0f * cq)));
        const float isq = 1.0f / (2.0f * sqrtf(cq));
        const float tau = fmaxf(Gf * ef * (erff((bq + 2.0f * cq * t1) * isq) - erff((bq + 2.0f * cq * t0) * isq)), 0.0f);
        s_st[warp][r] = expf(-tau);
        s_w[warp][r] = one_minus_exp_neg(tau) * s_alb[warp][e];
    }
    __syncwarp();
    if (lane == 0) {
        float Tr = 1.0f, acc = 0.f;
        for (int r = 0; r < ns; ++r) {
            acc = fmaf(Tr, s_w[warp][r], acc);                 // volume_renderer_analytic.cu:150-170
            Tr *= s_st[warp][r];
            if (Tr < 1e-4f) break;
        }
        out[ray] = acc;
    }
}

// ------------------------------------------------------------------------------------------------ host side
static int compat_wpb(int T, size_t* smem_out)
{
    // [T] radii + wpb * 3 * [T]; shrink the CTA until it fits 200 KB
    for (int wpb = CP_WPB; wpb >= 1; wpb >>= 1) {
        size_t smem = ((size_t)T + (size_t)wpb * 3 * T) * sizeof(float);
        if (smem <= 200 * 1024) { *smem_out = smem; return wpb; }
    }
    return 0;
}

extern "C" int nlosb200_compat_albedo(const float* means, const float* features, const float* camera_pos, int N, int sh_dim,
                                      int sh_degree, float* records, void* stream)
{
    if (N <= 0) return 0;
    NLOS_REQUIRE(sh_degree >= 0 && sh_degree <= 3, "compat_albedo: active_sh_degree=%d not in [0,3]", sh_degree);
    NLOS_REQUIRE(sh_dim >= (sh_degree + 1) * (sh_degree + 1),
                 "compat_albedo: features hold %d SH coefficients but degree %d needs %d (the reference reads out of bounds here, "
                 "volume_renderer.cu:110)", sh_dim, sh_degree, (sh_degree + 1) * (sh_degree + 1));
    NLOS_REQUIRE(means && features && camera_pos && records, "compat_albedo: null pointer");
    compat_albedo_kernel<<<(N + 255) / 256, 256, 0, (cudaStream_t)stream>>>(means, features, camera_pos, N, sh_dim, sh_degree, records);
    NLOS_CHECK_LAUNCH("compat_albedo_kernel");
    return 0;
}

extern "C" int nlosb200_compat_albedo_backward(const float* means, const float* features, const float* camera_pos, int N, int sh_dim,
                                               int sh_degree, const float* dacc, float* d_means, float* d_features, void* stream)
{
    if (N <= 0) return 0;
    NLOS_REQUIRE(sh_degree >= 0 && sh_degree <= 3, "compat_albedo_backward: active_sh_degree=%d not in [0,3]", sh_degree);
    NLOS_REQUIRE(sh_dim >= (sh_degree + 1) * (sh_degree + 1), "compat_albedo_backward: %d SH coefficients, degree %d needs %d", sh_dim, sh_degree,
                 (sh_degree + 1) * (sh_degree + 1));
    NLOS_REQUIRE(means && features && camera_pos && dacc && d_means && d_features, "compat_albedo_backward: null pointer");
    compat_albedo_backward_kernel<<<(N + 255) / 256, 256, 0, (cudaStream_t)stream>>>(means, features, camera_pos, N, sh_dim, sh_degree, dacc, d_means,
                                                                                   d_features);
    NLOS_CHECK_LAUNCH("compat_albedo_backward_kernel");
    return 0;
}

extern "C" int nlosb200_render_rays(const float* ray_origins, const float* ray_directions, const float* t_samples, const float* records,
                                    const int* filter, int cap, int n_rays, int T, int N, float c, float deltaT, int use_occlusion,
                                    float* rho_density, float* density, float* transmittance, void* stream)
{
    if (n_rays <= 0 || T <= 0) return 0;
    NLOS_REQUIRE(ray_origins && ray_directions && t_samples && filter && rho_density && density && transmittance && (records || N == 0),
                 "render_rays: null pointer");
    NLOS_REQUIRE(((uintptr_t)records % 16) == 0, "render_rays: records must be 16-byte aligned");
    size_t smem = 0;
    int wpb = compat_wpb(T, &smem);
    NLOS_REQUIRE(wpb > 0, "render_rays: %d samples per ray need more than 200 KB of shared memory per ray", T);
    cudaStream_t s = (cudaStream_t)stream;
    dim3 grid((n_rays + wpb - 1) / wpb), block(wpb * 32);
    cudaError_t e;
    if (use_occlusion) {
        e = cudaFuncSetAttribute(compat_render_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        NLOS_REQUIRE(e == cudaSuccess, "render_rays: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
        compat_render_kernel<true><<<grid, block, smem, s>>>(ray_origins, ray_directions, t_samples, records, filter, cap, n_rays, T, N, c,
                                                             deltaT, rho_density, density, transmittance);
    } else {
        e = cudaFuncSetAttribute(compat_render_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        NLOS_REQUIRE(e == cudaSuccess, "render_rays: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
        compat_render_kernel<false><<<grid, block, smem, s>>>(ray_origins, ray_directions, t_samples, records, filter, cap, n_rays, T, N, c,
                                                              deltaT, rho_density, density, transmittance);
    }
    NLOS_CHECK_LAUNCH("compat_render_kernel");
    return 0;
}

extern "C" int nlosb200_render_rays_backward(const float* ray_origins, const float* ray_directions, const float* t_samples,
                                             const float* records, const int* filter, int cap, int n_rays, int T, int N, float c,
                                             float deltaT, int use_occlusion, const float* rho_density, const float* transmittance,
                                             const float* grad_rho, const float* grad_density, const float* grad_transmittance, float* dacc,
                                             void* stream)
{
    NLOS_REQUIRE(dacc || N == 0, "render_rays_backward: null dacc");
    cudaStream_t s = (cudaStream_t)stream;
    if (N > 0) {
        cudaError_t e0 = cudaMemsetAsync(dacc, 0, (size_t)N * NLOS_GRAD_FLOATS * sizeof(float), s);
        NLOS_REQUIRE(e0 == cudaSuccess, "render_rays_backward: memset: %s", cudaGetErrorString(e0));
    }
    if (n_rays <= 0 || T <= 0 || N <= 0) return 0;
    NLOS_REQUIRE(ray_origins && ray_directions && t_samples && filter && records, "render_rays_backward: null pointer");
    NLOS_REQUIRE(!use_occlusion || (rho_density && transmittance), "render_rays_backward: occlusion needs the saved rho/transmittance");
    size_t smem = 0;
    int wpb = compat_wpb(T, &smem);
    NLOS_REQUIRE(wpb > 0, "render_rays_backward: %d samples per ray need more than 200 KB of shared memory per ray", T);
    dim3 grid((n_rays + wpb - 1) / wpb), block(wpb * 32);
    cudaError_t e;
    if (use_occlusion) {
        e = cudaFuncSetAttribute(compat_render_backward_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        NLOS_REQUIRE(e == cudaSuccess, "render_rays_backward: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
        compat_render_backward_kernel<true><<<grid, block, smem, s>>>(ray_origins, ray_directions, t_samples, records, filter, cap, n_rays, T,
                                                                      N, c, deltaT, rho_density, transmittance, grad_rho, grad_density,
                                                                      grad_transmittance, dacc);
    } else {
        e = cudaFuncSetAttribute(compat_render_backward_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        NLOS_REQUIRE(e == cudaSuccess, "render_rays_backward: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
        compat_render_backward_kernel<false><<<grid, block, smem, s>>>(ray_origins, ray_directions, t_samples, records, filter, cap, n_rays, T,
                                                                       N, c, deltaT, rho_density, transmittance, grad_rho, grad_density,
                                                                       grad_transmittance, dacc);
    }
    NLOS_CHECK_LAUNCH("compat_render_backward_kernel");
    return 0;
}

extern "C" int nlosb200_render_rays_analytic(const float* ray_origins, const float* ray_directions, float t_min, float t_max,
                                             const int* filter, int cap, const float* means, const float* log_scales,
                                             const float* rotations, const float* opacities, const float* features, const float* camera_pos,
                                             int n_rays, int N, int sh_dim, int sh_degree, float scaling_modifier, float sigma_threshold,
                                             int max_sections, float* out, void* stream)
{
    if (n_rays <= 0) return 0;
    NLOS_REQUIRE(sh_degree >= 0 && sh_degree <= 3, "render_rays_analytic: active_sh_degree=%d not in [0,3]", sh_degree);
    NLOS_REQUIRE(sh_dim >= (sh_degree + 1) * (sh_degree + 1), "render_rays_analytic: features hold %d SH coefficients, degree %d needs %d",
                 sh_dim, sh_degree, (sh_degree + 1) * (sh_degree + 1));
    NLOS_REQUIRE(max_sections >= 0 && max_sections <= CP_MAX_SECTIONS, "render_rays_analytic: max_sections=%d not in [0,%d]", max_sections,
                 CP_MAX_SECTIONS);
    NLOS_REQUIRE(ray_origins && ray_directions && filter && out && camera_pos, "render_rays_analytic: null pointer");
    NLOS_REQUIRE(N == 0 || (means && log_scales && rotations && opacities && features), "render_rays_analytic: null parameter pointer");
    compat_analytic_kernel<<<(n_rays + CP_WPB - 1) / CP_WPB, CP_WPB * 32, 0, (cudaStream_t)stream>>>(
        ray_origins, ray_directions, t_min, t_max, filter, cap, means, log_scales, rotations, opacities, features, camera_pos, n_rays, N, sh_dim,
        sh_degree, scaling_modifier, sigma_threshold, max_sections, out);
    NLOS_CHECK_LAUNCH("compat_analytic_kernel");
    return 0;
}
