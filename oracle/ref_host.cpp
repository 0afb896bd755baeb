/* ref_host.cpp -- host driver around the REFERENCE's own kernel code.  TEST INFRASTRUCTURE ONLY.
 *
 * oracle/build_ref.py compiles this file with g++ against
 *   - the reference's device headers, read where they lie:
 *       /root/reference/submodules/cuda_renderer/include/{cuda_utils,bbox_compute,
 *       spherical_harmonics,analytic_integration}.cuh
 *   - oracle/_ref/ref_kernels_gen.inc, generated at build time (git-ignored) by
 *     extracting the three __global__ kernel bodies verbatim from
 *       src/ray_aabb.cu (filter_gaussians_kernel), src/volume_renderer.cu
 *       (volume_render_kernel) and src/volume_renderer_analytic.cu
 *       (volume_render_analytic_kernel)
 *   - oracle/ref_shim/ (cuda_runtime.h stand-in: __device__, float3, threadIdx...).
 * No reference source is stored in this repository.  Only the launch sequences
 * of the reference's host wrappers are restated here (cited below); every
 * arithmetic operation executed is the reference's own text.
 */
#include "cuda_runtime.h"
#include "cuda_utils.cuh"
#include "bbox_compute.cuh"
#include "spherical_harmonics.cuh"
#include "analytic_integration.cuh"
#include "ref_kernels_gen.inc"

#include <cstring>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

#define REF_TPB 256

template <class F>
static void launch_1d(int n, F body)
{
    /* grid ceil(n/256) x block 256, one thread per element, as every reference launch does */
#pragma omp parallel for schedule(dynamic, 16)
    for (int idx = 0; idx < ((n + REF_TPB - 1) / REF_TPB) * REF_TPB; ++idx) {
        blockDim.x = REF_TPB;
        blockIdx.x = idx / REF_TPB;
        threadIdx.x = idx % REF_TPB;
        body();
    }
}

extern "C" {

int ref_max_gaussians_per_ray(void) { return MAX_GAUSSIANS_PER_RAY; }
int ref_max_sections_per_ray(void) { return MAX_SECTIONS_PER_RAY; }

/* volume_renderer.cu:224-236 */
void ref_bboxes(const float *means, const float *scales, const float *rots, int N, float mod, float sigma, float *out)
{
    launch_1d(N, [&] { compute_gaussian_bboxes_kernel(means, scales, rots, N, mod, sigma, out); });
}

/* ray_aabb.cu:63-102 : indices pre-filled with -1, counts zero, result = cat(count, indices) */
void ref_filter(const float *ro, const float *rd, const float *bboxes, int n_rays, int N, int *out /*[n_rays,257]*/)
{
    std::vector<int> idx((size_t)n_rays * MAX_GAUSSIANS_PER_RAY, -1), cnt((size_t)n_rays, 0);
    launch_1d(n_rays, [&] { filter_gaussians_kernel(ro, rd, bboxes, n_rays, N, idx.data(), cnt.data()); });
    for (int r = 0; r < n_rays; ++r) {
        out[(size_t)r * (MAX_GAUSSIANS_PER_RAY + 1)] = cnt[r];
        std::memcpy(out + (size_t)r * (MAX_GAUSSIANS_PER_RAY + 1) + 1, idx.data() + (size_t)r * MAX_GAUSSIANS_PER_RAY,
                    sizeof(int) * MAX_GAUSSIANS_PER_RAY);
    }
}

/* volume_renderer.cu:189-305 : bboxes (sigma 3.0) -> filter -> volume_render_kernel, outputs zero-initialised */
void ref_render_rays(const float *ro, const float *rd, const float *t_samples,
                     const float *means, const float *scales, const float *rots, const float *opac,
                     const float *feats, const float *cam, int n_rays, int T, int N, int sh_degree, int sh_dim,
                     float c, float deltaT, float mod, int use_occlusion, int rendering_type,
                     float *rho, float *density, float *trans, int *filter_out /* nullable [n_rays,257] */)
{
    std::vector<float> bb((size_t)N * 6);
    ref_bboxes(means, scales, rots, N, mod, 3.0f, bb.data());
    std::vector<int> filt((size_t)n_rays * (MAX_GAUSSIANS_PER_RAY + 1));
    ref_filter(ro, rd, bb.data(), n_rays, N, filt.data());
    if (filter_out) std::memcpy(filter_out, filt.data(), sizeof(int) * filt.size());
    std::memset(rho, 0, sizeof(float) * (size_t)n_rays * T);
    std::memset(density, 0, sizeof(float) * (size_t)n_rays * T);
    std::memset(trans, 0, sizeof(float) * (size_t)n_rays * T);
    launch_1d(n_rays, [&] {
        volume_render_kernel(ro, rd, t_samples, filt.data(), means, scales, rots, opac, feats, cam, n_rays, T, N,
                             sh_degree, sh_dim, c, deltaT, mod, use_occlusion != 0, rendering_type, rho, density, trans);
    });
}

/* the same kernel with the per-ray lists GIVEN (frozen ray membership: finite differences of the reference forward must not move rays in
 * or out of a 3-sigma AABB -- the hard cull is a discontinuity of the reference) */
void ref_render_rays_with_filter(const float *ro, const float *rd, const float *t_samples,
                                 const float *means, const float *scales, const float *rots, const float *opac,
                                 const float *feats, const float *cam, int n_rays, int T, int N, int sh_degree, int sh_dim,
                                 float c, float deltaT, float mod, int use_occlusion, int rendering_type, const int *filter /*[n_rays,257]*/,
                                 float *rho, float *density, float *trans)
{
    std::memset(rho, 0, sizeof(float) * (size_t)n_rays * T);
    std::memset(density, 0, sizeof(float) * (size_t)n_rays * T);
    std::memset(trans, 0, sizeof(float) * (size_t)n_rays * T);
    launch_1d(n_rays, [&] {
        volume_render_kernel(ro, rd, t_samples, filter, means, scales, rots, opac, feats, cam, n_rays, T, N,
                             sh_degree, sh_dim, c, deltaT, mod, use_occlusion != 0, rendering_type, rho, density, trans);
    });
}

/* volume_renderer_analytic.cu:178-241 */
void ref_render_rays_analytic(const float *ro, const float *rd, float t_min, float t_max, const int *filter,
                              const float *means, const float *scales, const float *rots, const float *opac,
                              const float *feats, const float *cam, int n_rays, int N, int sh_degree, int sh_dim,
                              float c, float deltaT, float mod, float sigma_threshold, int rendering_type, float *out)
{
    std::memset(out, 0, sizeof(float) * (size_t)n_rays);
    launch_1d(n_rays, [&] {
        volume_render_analytic_kernel(ro, rd, t_min, t_max, filter, means, scales, rots, opac, feats, cam, n_rays, N,
                                      sh_degree, sh_dim, c, deltaT, mod, sigma_threshold, rendering_type, out);
    });
}

int ref_num_threads(void)
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
void ref_set_num_threads(int n)
{
#ifdef _OPENMP
    omp_set_num_threads(n);
#else
    (void)n;
#endif
}
}
