// preprocess.cu -- per-Gaussian activation / packing kernels and the scan-tile table builder.
//
// Replaces (reference, /root/reference/submodules/cuda_renderer):
//   include/bbox_compute.cuh:76-120  compute_gaussian_bboxes_kernel
//   src/volume_renderer.cu:86-111    the per-sample re-activation of scale / opacity / rotation / albedo
// and adds the backward chain M,mu,op,alb -> raw parameters that the reference
// never implemented (gaussian_model/cuda_autograd.py:139-172 returns zeros).
#include "common.cuh"
#include "../../include/nlosb200.h"

#include <stdarg.h>

namespace nlos {
static thread_local char g_err[512] = "";
void set_error(const char* fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
}  // namespace nlos

extern "C" const char* nlosb200_last_error(void) { return nlos::g_err; }
extern "C" int nlosb200_version(void) { return 100; }
extern "C" int nlosb200_record_floats(void) { return NLOS_REC_FLOATS; }
extern "C" int nlosb200_grad_floats(void) { return NLOS_GRAD_FLOATS; }
extern "C" long long nlosb200_tile_floats(int A) { return (long long)(NLOS_NTAB * A + NLOS_PINFO_ROWS) * NLOS_TILE; }

using namespace nlos;

// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) preprocess_kernel(const float* __restrict__ means, const float* __restrict__ log_scales,
                                                         const float* __restrict__ rots, const float* __restrict__ opac,
                                                         const float* __restrict__ feats, int N, int sh_dim, float mod,
                                                         float sigma, float mcut, float tw_min, float* __restrict__ records,
                                                         float4* __restrict__ shell, float* __restrict__ bboxes)
{
    int g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= N) return;
    float R[9];
    quat_to_rotmat_x(rots[4 * g + 0], rots[4 * g + 1], rots[4 * g + 2], rots[4 * g + 3], R);
    // volume_renderer.cu:92-96 / bbox_compute.cuh:93-97 : expf(s) * scaling_modifier
    float sx = __fmul_rn(det_expf(log_scales[3 * g + 0]), mod);
    float sy = __fmul_rn(det_expf(log_scales[3 * g + 1]), mod);
    float sz = __fmul_rn(det_expf(log_scales[3 * g + 2]), mod);
    float mx = means[3 * g + 0], my = means[3 * g + 1], mz = means[3 * g + 2];
    float ex = extent_x(R[0], R[1], R[2], sx, sy, sz, sigma);
    float ey = extent_x(R[3], R[4], R[5], sx, sy, sz, sigma);
    float ez = extent_x(R[6], R[7], R[8], sx, sy, sz, sigma);
    float bmin[3] = {__fsub_rn(mx, ex), __fsub_rn(my, ey), __fsub_rn(mz, ez)};
    float bmax[3] = {__fadd_rn(mx, ex), __fadd_rn(my, ey), __fadd_rn(mz, ez)};
    if (bboxes) {
        float* b = bboxes + 6 * (size_t)g;
        b[0] = bmin[0]; b[1] = bmin[1]; b[2] = bmin[2];
        b[3] = bmax[0]; b[4] = bmax[1]; b[5] = bmax[2];
    }
    float rad = __fmul_rn(mcut, fmaxf(fmaxf(sx, sy), sz));
    // the sign bit of the shell radius flags a time-wide Gaussian: smallest scale >= tw_min (= 1.3 bins; 0 switches the flag off)
    if (shell) shell[g] = make_float4(mx, my, mz, (tw_min > 0.f && fminf(fminf(sx, sy), sz) >= tw_min) ? -rad : rad);
    if (!records) return;
    // cuda_utils.cuh:124-151 : t = R^T (x - mu), m2 = sum (t_j / (s_j + 1e-8))^2
    float is0 = 1.0f / (sx + 1e-8f), is1 = 1.0f / (sy + 1e-8f), is2 = 1.0f / (sz + 1e-8f);
    float op = 1.0f / (1.0f + expf(-opac[g]));                                      // volume_renderer.cu:105
    float alb = fmaxf(fmaf(NLOS_SH_C0, feats[(size_t)g * sh_dim], 0.5f), 0.0f);     // volume_renderer.cu:109-111 (degree 0)
    float4* rec = reinterpret_cast<float4*>(records + (size_t)g * NLOS_REC_FLOATS);
    rec[0] = make_float4(mx, my, mz, op);
    rec[1] = make_float4(R[0] * is0, R[3] * is0, R[6] * is0, R[1] * is1);
    rec[2] = make_float4(R[4] * is1, R[7] * is1, R[2] * is2, R[5] * is2);
    rec[3] = make_float4(R[8] * is2, bmin[0], bmin[1], bmin[2]);
    rec[4] = make_float4(bmax[0], bmax[1], bmax[2], alb);
    rec[5] = make_float4(rad, sx, sy, sz);
    // half extents of the AABB about the mean as the render kernels see them (max distance of a box face from mu, rounded up) and their norm
    float hx = fmaxf(mx - bmin[0], bmax[0] - mx) * 1.00001f + 1e-9f, hy = fmaxf(my - bmin[1], bmax[1] - my) * 1.00001f + 1e-9f,
          hz = fmaxf(mz - bmin[2], bmax[2] - mz) * 1.00001f + 1e-9f;
    rec[6] = make_float4(hx, hy, hz, sqrtf(hx * hx + hy * hy + hz * hz) * 1.00001f);
    rec[7] = make_float4(0.f, 0.f, 0.f, 0.f);
}

extern "C" int nlosb200_preprocess(const float* means, const float* log_scales, const float* rotations, const float* opacities,
                                   const float* features, int N, int sh_dim, float scaling_modifier, float sigma_bbox, float mcut,
                                   float time_wide_min, float* records, float* shell, float* bboxes, void* stream)
{
    if (N <= 0) return 0;
    NLOS_REQUIRE(means && log_scales && rotations, "preprocess: null parameter pointer");
    NLOS_REQUIRE(!records || (opacities && features && sh_dim >= 1), "preprocess: records need opacities and features");
    NLOS_REQUIRE(!records || ((uintptr_t)records % 128) == 0, "preprocess: records must be 128-byte aligned");
    preprocess_kernel<<<(N + 255) / 256, 256, 0, (cudaStream_t)stream>>>(means, log_scales, rotations, opacities, features, N, sh_dim,
                                                                         scaling_modifier, sigma_bbox, mcut, time_wide_min, records,
                                                                         reinterpret_cast<float4*>(shell), bboxes);
    NLOS_CHECK_LAUNCH("preprocess_kernel");
    return 0;
}

// ------------------------------------------------------------------------------------------------
// backward chain: dacc[g] = { dL/dM (9, row major), dL/dmu (3), dL/dop, dL/dalb0, -, - }  ->  raw parameter gradients
__global__ void __launch_bounds__(256) preprocess_backward_kernel(const float* __restrict__ log_scales, const float* __restrict__ rots,
                                                                  const float* __restrict__ opac, const float* __restrict__ feats, int N,
                                                                  int sh_dim, float mod, const float* __restrict__ dacc,
                                                                  float* __restrict__ d_means, float* __restrict__ d_scales,
                                                                  float* __restrict__ d_rots, float* __restrict__ d_opac,
                                                                  float* __restrict__ d_feats, int accumulate, const float* __restrict__ dsh = nullptr,
                                                                  int sh_degree = 0)
{
    int g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= N) return;
    const float4* da4 = reinterpret_cast<const float4*>(dacc + (size_t)g * NLOS_GRAD_FLOATS);
    float4 a0 = da4[0], a1 = da4[1], a2 = da4[2], a3 = da4[3];
    float dM[9] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w, a2.x};
    float dmu[3] = {a2.y, a2.z, a2.w};
    float dop = a3.x, dalb = a3.y;

    float qw = rots[4 * g + 0], qx = rots[4 * g + 1], qy = rots[4 * g + 2], qz = rots[4 * g + 3];
    float R[9], inv_norm;
    bool has_rot = quat_to_rotmat_x(qw, qx, qy, qz, R, &inv_norm);
    float s[3], is[3];
#pragma unroll
    for (int j = 0; j < 3; ++j) {
        s[j] = __fmul_rn(det_expf(log_scales[3 * g + j]), mod);
        is[j] = 1.0f / (s[j] + 1e-8f);
    }
    // M[j][i] = R[i][j] * is[j]
    float dR[9];
    float dls[3];
#pragma unroll
    for (int j = 0; j < 3; ++j) {
        float acc = 0.f;
#pragma unroll
        for (int i = 0; i < 3; ++i) {
            dR[3 * i + j] = dM[3 * j + i] * is[j];
            acc = fmaf(dM[3 * j + i], R[3 * i + j], acc);
        }
        // ds_j = -sum_i dM[j][i] R[i][j] / (s_j+eps)^2 ;  dls_j = ds_j * s_j
        dls[j] = -acc * is[j] * is[j] * s[j];
    }
    float dq[4] = {0.f, 0.f, 0.f, 0.f};
    if (has_rot) {
        float w = qw * inv_norm, x = qx * inv_norm, y = qy * inv_norm, z = qz * inv_norm;
        const float g00 = dR[0], g01 = dR[1], g02 = dR[2], g10 = dR[3], g11 = dR[4], g12 = dR[5], g20 = dR[6], g21 = dR[7], g22 = dR[8];
        float dw = 2.f * (-z * g01 + y * g02 + z * g10 - x * g12 - y * g20 + x * g21);
        float dx = 2.f * (y * g01 + z * g02 + y * g10 - 2.f * x * g11 - w * g12 + z * g20 + w * g21 - 2.f * x * g22);
        float dy = 2.f * (-2.f * y * g00 + x * g01 + w * g02 + x * g10 + z * g12 - w * g20 + z * g21 - 2.f * y * g22);
        float dz = 2.f * (-2.f * z * g00 - w * g01 + x * g02 + w * g10 - 2.f * z * g11 + y * g12 + x * g20 + y * g21);
        float dot = w * dw + x * dx + y * dy + z * dz;
        dq[0] = (dw - w * dot) * inv_norm;
        dq[1] = (dx - x * dot) * inv_norm;
        dq[2] = (dy - y * dot) * inv_norm;
        dq[3] = (dz - z * dot) * inv_norm;
    }
    float op = 1.0f / (1.0f + expf(-opac[g]));
    float dopac = dop * op * (1.0f - op);
    float f0 = feats[(size_t)g * sh_dim];
    float df0 = (fmaf(NLOS_SH_C0, f0, 0.5f) > 0.f) ? dalb * NLOS_SH_C0 : 0.f;
    if (sh_degree > 0) {            // view-dependent albedo: the render backward already reduced dL/d(SH coefficients) over the scan points
        const int K = (sh_degree + 1) * (sh_degree + 1);
        for (int i = 0; i < 3; ++i) { d_means[3 * g + i] = (accumulate ? d_means[3 * g + i] : 0.f) + dmu[i]; d_scales[3 * g + i] = (accumulate ? d_scales[3 * g + i] : 0.f) + dls[i]; }
        for (int i = 0; i < 4; ++i) d_rots[4 * g + i] = (accumulate ? d_rots[4 * g + i] : 0.f) + dq[i];
        d_opac[g] = (accumulate ? d_opac[g] : 0.f) + dopac;
        for (int k = 0; k < sh_dim; ++k) {
            float v = k < K ? dsh[(size_t)g * 16 + k] : 0.f;
            d_feats[(size_t)g * sh_dim + k] = (accumulate ? d_feats[(size_t)g * sh_dim + k] : 0.f) + v;
        }
        return;
    }
    if (accumulate) {
        for (int i = 0; i < 3; ++i) { d_means[3 * g + i] += dmu[i]; d_scales[3 * g + i] += dls[i]; }
        for (int i = 0; i < 4; ++i) d_rots[4 * g + i] += dq[i];
        d_opac[g] += dopac;
        d_feats[(size_t)g * sh_dim] += df0;
    } else {
        for (int i = 0; i < 3; ++i) { d_means[3 * g + i] = dmu[i]; d_scales[3 * g + i] = dls[i]; }
        for (int i = 0; i < 4; ++i) d_rots[4 * g + i] = dq[i];
        d_opac[g] = dopac;
        d_feats[(size_t)g * sh_dim] = df0;
        for (int k = 1; k < sh_dim; ++k) d_feats[(size_t)g * sh_dim + k] = 0.f;
    }
}

extern "C" int nlosb200_preprocess_backward(const float* log_scales, const float* rotations, const float* opacities, const float* features,
                                            int N, int sh_dim, float scaling_modifier, const float* dacc, float* d_means,
                                            float* d_log_scales, float* d_rotations, float* d_opacities, float* d_features, int accumulate,
                                            void* stream)
{
    if (N <= 0) return 0;
    NLOS_REQUIRE(dacc && d_means && d_log_scales && d_rotations && d_opacities && d_features, "preprocess_backward: null pointer");
    NLOS_REQUIRE(((uintptr_t)dacc % 16) == 0, "preprocess_backward: dacc must be 16-byte aligned");
    preprocess_backward_kernel<<<(N + 255) / 256, 256, 0, (cudaStream_t)stream>>>(log_scales, rotations, opacities, features, N, sh_dim,
                                                                                  scaling_modifier, dacc, d_means, d_log_scales, d_rotations,
                                                                                  d_opacities, d_features, accumulate);
    NLOS_CHECK_LAUNCH("preprocess_backward_kernel");
    return 0;
}

extern "C" int nlosb200_preprocess_backward_sh(const float* log_scales, const float* rotations, const float* opacities, const float* features,
                                               int N, int sh_dim, int sh_degree, float scaling_modifier, const float* dacc, const float* dsh,
                                               float* d_means, float* d_log_scales, float* d_rotations, float* d_opacities, float* d_features,
                                               int accumulate, void* stream)
{
    if (N <= 0) return 0;
    NLOS_REQUIRE(dacc && d_means && d_log_scales && d_rotations && d_opacities && d_features, "preprocess_backward_sh: null pointer");
    NLOS_REQUIRE(sh_degree >= 0 && sh_degree <= 3 && (sh_degree == 0 || (dsh && sh_dim >= (sh_degree + 1) * (sh_degree + 1))),
                 "preprocess_backward_sh: degree %d with %d coefficients / null dsh", sh_degree, sh_dim);
    NLOS_REQUIRE(((uintptr_t)dacc % 16) == 0, "preprocess_backward_sh: dacc must be 16-byte aligned");
    preprocess_backward_kernel<<<(N + 255) / 256, 256, 0, (cudaStream_t)stream>>>(log_scales, rotations, opacities, features, N, sh_dim,
                                                                                  scaling_modifier, dacc, d_means, d_log_scales, d_rotations,
                                                                                  d_opacities, d_features, accumulate, dsh, sh_degree);
    NLOS_CHECK_LAUNCH("preprocess_backward_kernel(sh)");
    return 0;
}

// compat entry: bboxes only (include/bbox_compute.cuh:76-120; gaussian_model.py:140-178 get_bboxes)
extern "C" int nlosb200_compute_bboxes(const float* means, const float* log_scales, const float* rotations, int N, float scaling_modifier,
                                       float sigma, float* bboxes, void* stream)
{
    if (N <= 0) return 0;
    NLOS_REQUIRE(means && log_scales && rotations && bboxes, "compute_bboxes: null pointer");
    preprocess_kernel<<<(N + 255) / 256, 256, 0, (cudaStream_t)stream>>>(means, log_scales, rotations, nullptr, nullptr, N, 1, scaling_modifier,
                                                                         sigma, 0.f, 0.f, nullptr, nullptr, bboxes);
    NLOS_CHECK_LAUNCH("preprocess_kernel(bboxes)");
    return 0;
}

// ------------------------------------------------------------------------------------------------
// Scan-tile tables.  A tile = 32 consecutive scan points (caller orders them tile-major); per tile one
// contiguous block theta [A][3][32] + phi [A][4][32] (common.cuh) + [NLOS_PINFO_ROWS][32] floats that the render kernels pull into
// shared memory with ONE TMA bulk copy.  The lattice follows gaussian_model/cuda_autograd.py:254-268:
// theta_i / phi_j = linspace(min, max, A); the caller passes the sin/cos tables it built with torch so
// that this path and the oracle consume identical ray directions.
__global__ void __launch_bounds__(128) build_scan_tiles_kernel(const float* __restrict__ scan_o, const float* __restrict__ sin_th,
                                                               const float* __restrict__ cos_th, const float* __restrict__ sin_ph,
                                                               const float* __restrict__ cos_ph, const float* __restrict__ lattice,
                                                               const float* __restrict__ hscale, const int* __restrict__ valid_mask, int S, int A,
                                                               float* __restrict__ tiles)
{
    int p = blockIdx.x * blockDim.x + threadIdx.x;
    int n_tiles = (S + NLOS_TILE - 1) / NLOS_TILE;
    if (p >= n_tiles * NLOS_TILE) return;
    int tile = p / NLOS_TILE, lane = p % NLOS_TILE;
    size_t tile_floats = (size_t)(NLOS_NTAB * A + NLOS_PINFO_ROWS) * NLOS_TILE;
    float* tb = tiles + tile * tile_floats;
    float* pi = tb + (size_t)NLOS_NTAB * A * NLOS_TILE;
    bool valid = p < S && (!valid_mask || valid_mask[p] != 0);
    float istmax = 1.f;
    for (int i = 0; i < A; ++i) {
        float st = valid ? sin_th[(size_t)p * A + i] : 1.f, ct = valid ? cos_th[(size_t)p * A + i] : 1.f;
        istmax = fmaxf(istmax, 1.0f / fmaxf(fabsf(st), 1e-12f));
        float sp = valid ? sin_ph[(size_t)p * A + i] : 1.f, cp = valid ? cos_ph[(size_t)p * A + i] : 1.f;
        tb[TAB_TH(A, i) + TH_SIN + lane] = st;
        tb[TAB_TH(A, i) + TH_COS + lane] = ct;
        tb[TAB_TH(A, i) + TH_TAN + lane] = __fdiv_rn(st, ct);
        tb[TAB_PH(A, i) + PH_SIN + lane] = sp;
        tb[TAB_PH(A, i) + PH_COS + lane] = cp;
        tb[TAB_PH(A, i) + PH_ISIN + lane] = __frcp_rn(sp);
        tb[TAB_PH(A, i) + PH_ICOS + lane] = __frcp_rn(cp);
    }
    float th0 = 0.f, dth = 0.f, ph0 = 0.f, dph = 0.f;
    if (valid) { th0 = lattice[4 * p + 0]; dth = lattice[4 * p + 1]; ph0 = lattice[4 * p + 2]; dph = lattice[4 * p + 3]; }
    const float PI_F = 3.14159265358979f;
    bool fast = valid && A > 1 && dth > 1e-9f && dph > 1e-9f && th0 >= 0.f && th0 + dth * (A - 1) <= PI_F && ph0 >= -PI_F &&
                ph0 + dph * (A - 1) <= PI_F;
    pi[PI_OX * NLOS_TILE + lane] = valid ? scan_o[3 * p + 0] : 0.f;
    pi[PI_OY * NLOS_TILE + lane] = valid ? scan_o[3 * p + 1] : 0.f;
    pi[PI_OZ * NLOS_TILE + lane] = valid ? scan_o[3 * p + 2] : 0.f;
    pi[PI_TH0 * NLOS_TILE + lane] = th0;
    pi[PI_IDTH * NLOS_TILE + lane] = fast ? 1.0f / dth : 0.f;
    pi[PI_PH0 * NLOS_TILE + lane] = ph0;
    pi[PI_IDPH * NLOS_TILE + lane] = fast ? 1.0f / dph : 0.f;
    pi[PI_HSCALE * NLOS_TILE + lane] = valid ? hscale[p] : 0.f;
    pi[PI_FLAGS * NLOS_TILE + lane] = __int_as_float((valid ? 1 : 0) | (fast ? 2 : 0));
    pi[PI_ISTMAX * NLOS_TILE + lane] = istmax;
    pi[PI_DTH * NLOS_TILE + lane] = dth;
    pi[PI_DPH * NLOS_TILE + lane] = dph;
}

extern "C" int nlosb200_build_scan_tiles(const float* scan_o, const float* sin_th, const float* cos_th, const float* sin_ph,
                                         const float* cos_ph, const float* lattice, const float* hscale, const int* valid, int S, int A,
                                         float* tiles, void* stream)
{
    if (S <= 0) return 0;
    NLOS_REQUIRE(A >= 1 && A <= 128, "build_scan_tiles: A=%d out of range [1,128]", A);
    NLOS_REQUIRE(scan_o && sin_th && cos_th && sin_ph && cos_ph && lattice && hscale && tiles, "build_scan_tiles: null pointer");
    NLOS_REQUIRE(((uintptr_t)tiles % 128) == 0, "build_scan_tiles: tiles must be 128-byte aligned");
    int n = ((S + NLOS_TILE - 1) / NLOS_TILE) * NLOS_TILE;
    build_scan_tiles_kernel<<<(n + 127) / 128, 128, 0, (cudaStream_t)stream>>>(scan_o, sin_th, cos_th, sin_ph, cos_ph, lattice, hscale, valid, S, A,
                                                                               tiles);
    NLOS_CHECK_LAUNCH("build_scan_tiles_kernel");
    return 0;
}
