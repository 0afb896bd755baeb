// common.cuh -- shared device helpers for libnlosb200 (sm_100a only).
//
// Conventions
//   * "record": the 32-float (128 B, 128 B aligned) per-Gaussian block written by
//     the preprocess kernel and gathered into shared memory with TMA bulk copies.
//   * functions suffixed _x ("exact") use only correctly-rounded single IEEE
//     operations (__fadd_rn/__fmul_rn/__fmaf_rn/__fsqrt_rn/__frcp_rn) in a fixed
//     order, so that culling decisions are bit-identical to the CPU oracle
//     (oracle/nlos_oracle.c, compiled with -ffp-contract=off).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define NLOS_REC_FLOATS 32
#define NLOS_GRAD_FLOATS 16
#define NLOS_TILE 32          // scan points per tile == lanes per warp
#define NLOS_NTAB 7           // per-tile lattice tables
#define NLOS_PINFO_ROWS 12
#define NLOS_ENTRY_SHIFT 26   // list entry = gaussian id | (sub-interval index << 26)

// record layout (float index)
#define REC_MU 0      // 0..2 mean
#define REC_OP 3      // sigmoid(opacity logit)
#define REC_M 4       // 4..12  M = diag(1/(s+1e-8)) R^T, row major  (m2 = |M (x-mu)|^2)
#define REC_BMIN 13   // 13..15 3-sigma AABB min  (bbox_compute.cuh:23-120)
#define REC_BMAX 16   // 16..18 3-sigma AABB max
#define REC_ALB 19    // degree-0 albedo  max(C0*f0 + 0.5, 0)
#define REC_RAD 20    // mcut * max(scale)
#define REC_S 21      // 21..23 activated scales
#define REC_EXT 24    // 24..26 half extents of the 3-sigma AABB about the mean (rounded up), 27 their norm Rb

// lattice tables per tile (NLOS_NTAB * A rows of 32 lanes):
//   theta block  [A][3][32] : row i -> sin(theta_i), cos(theta_i), tan(theta_i)
//   phi block    [A][4][32] : row j -> sin(phi_j), cos(phi_j), 1/sin(phi_j), 1/cos(phi_j)
// so that ONE address per lattice index reaches all of its values with immediate offsets.
#define TAB_TH(A, i) ((i) * 96)                      // float offset of theta row i
#define TAB_PH(A, j) (3 * (A) * 32 + (j) * 128)      // float offset of phi row j
#define TH_SIN 0
#define TH_COS 32
#define TH_TAN 64
#define PH_SIN 0
#define PH_COS 32
#define PH_ISIN 64
#define PH_ICOS 96
// pinfo rows: [NLOS_PINFO_ROWS][32]
#define PI_OX 0
#define PI_OY 1
#define PI_OZ 2
#define PI_TH0 3
#define PI_IDTH 4
#define PI_PH0 5
#define PI_IDPH 6
#define PI_HSCALE 7
#define PI_FLAGS 8    // bit0 valid, bit1 fast footprint ok
#define PI_DTH 10     // lattice steps (the linspace's, theta / phi)
#define PI_DPH 11
#define PI_ISTMAX 9   // max_i 1 / sin(theta_i) of the scan point's lattice (bound used by the borderline band of the slab test)

#define NLOS_SH_C0 0.28209479177387814f

namespace nlos {

// ---------------------------------------------------------------- exact math
// det_expf: identical operation sequence to oracle/nlos_oracle.c:det_expf
__device__ __forceinline__ float det_expf(float x)
{
    if (x > 88.0f) x = 88.0f;
    if (x < -87.0f) x = -87.0f;
    float t = __fmul_rn(x, 1.4426950408889634f);
    float n = rintf(t);
    float r = __fmaf_rn(n, -0.693145751953125f, x);
    r = __fmaf_rn(n, -1.42860682030941723212e-6f, r);
    float p = 1.9875691500e-4f;
    p = __fmaf_rn(p, r, 1.3981999507e-3f);
    p = __fmaf_rn(p, r, 8.3334519073e-3f);
    p = __fmaf_rn(p, r, 4.1665795894e-2f);
    p = __fmaf_rn(p, r, 1.6666665459e-1f);
    p = __fmaf_rn(p, r, 5.0000001201e-1f);
    p = __fmaf_rn(p, __fmul_rn(r, r), r);
    p = __fadd_rn(p, 1.0f);
    int e = (int)n;
    float s = __int_as_float((e + 127) << 23);
    return __fmul_rn(p, s);
}

// cuda_utils.cuh:54-85 quat_to_rotmat (w,x,y,z; normalised inside; identity if |q|<1e-8),
// evaluated with the oracle's operation order.  Returns false for the identity fallback.
__device__ __forceinline__ bool quat_to_rotmat_x(float w, float x, float y, float z, float* R, float* inv_norm_out = nullptr)
{
    float n2 = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(w, w), __fmul_rn(x, x)), __fmul_rn(y, y)), __fmul_rn(z, z));
    float norm = __fsqrt_rn(n2);
    if (norm < 1e-8f) {
        R[0] = 1; R[1] = 0; R[2] = 0; R[3] = 0; R[4] = 1; R[5] = 0; R[6] = 0; R[7] = 0; R[8] = 1;
        if (inv_norm_out) *inv_norm_out = 0.f;
        return false;
    }
    w = __fdiv_rn(w, norm); x = __fdiv_rn(x, norm); y = __fdiv_rn(y, norm); z = __fdiv_rn(z, norm);
    if (inv_norm_out) *inv_norm_out = __frcp_rn(norm);
    R[0] = __fsub_rn(1.0f, __fmul_rn(2.0f, __fadd_rn(__fmul_rn(y, y), __fmul_rn(z, z))));
    R[1] = __fmul_rn(2.0f, __fsub_rn(__fmul_rn(x, y), __fmul_rn(w, z)));
    R[2] = __fmul_rn(2.0f, __fadd_rn(__fmul_rn(x, z), __fmul_rn(w, y)));
    R[3] = __fmul_rn(2.0f, __fadd_rn(__fmul_rn(x, y), __fmul_rn(w, z)));
    R[4] = __fsub_rn(1.0f, __fmul_rn(2.0f, __fadd_rn(__fmul_rn(x, x), __fmul_rn(z, z))));
    R[5] = __fmul_rn(2.0f, __fsub_rn(__fmul_rn(y, z), __fmul_rn(w, x)));
    R[6] = __fmul_rn(2.0f, __fsub_rn(__fmul_rn(x, z), __fmul_rn(w, y)));
    R[7] = __fmul_rn(2.0f, __fadd_rn(__fmul_rn(y, z), __fmul_rn(w, x)));
    R[8] = __fsub_rn(1.0f, __fmul_rn(2.0f, __fadd_rn(__fmul_rn(x, x), __fmul_rn(y, y))));
    return true;
}

// 3-sigma extent of one axis: sigma * sqrt((Ra sx)^2 + (Rb sy)^2 + (Rc sz)^2)   (bbox_compute.cuh:23-74)
__device__ __forceinline__ float extent_x(float ra, float rb, float rc, float sx, float sy, float sz, float sigma)
{
    float a = __fmul_rn(ra, sx), b = __fmul_rn(rb, sy), c = __fmul_rn(rc, sz);
    float s = __fadd_rn(__fadd_rn(__fmul_rn(a, a), __fmul_rn(b, b)), __fmul_rn(c, c));
    return __fmul_rn(sigma, __fsqrt_rn(s));
}

// cuda_utils.cuh:97-121 ray_aabb_intersect, bit-exact: inv = 1/(d + 1e-8) per axis
__device__ __forceinline__ bool ray_aabb_x(float ox, float oy, float oz, float dx, float dy, float dz,
                                           const float* __restrict__ b, float* tmin_o = nullptr, float* tmax_o = nullptr)
{
    float ix = __frcp_rn(__fadd_rn(dx, 1e-8f)), iy = __frcp_rn(__fadd_rn(dy, 1e-8f)), iz = __frcp_rn(__fadd_rn(dz, 1e-8f));
    float t0x = __fmul_rn(__fsub_rn(b[0], ox), ix), t0y = __fmul_rn(__fsub_rn(b[1], oy), iy), t0z = __fmul_rn(__fsub_rn(b[2], oz), iz);
    float t1x = __fmul_rn(__fsub_rn(b[3], ox), ix), t1y = __fmul_rn(__fsub_rn(b[4], oy), iy), t1z = __fmul_rn(__fsub_rn(b[5], oz), iz);
    float tmin = fmaxf(fmaxf(fminf(t0x, t1x), fminf(t0y, t1y)), fminf(t0z, t1z));
    float tmax = fminf(fminf(fmaxf(t0x, t1x), fmaxf(t0y, t1y)), fmaxf(t0z, t1z));
    if (tmin_o) *tmin_o = tmin;
    if (tmax_o) *tmax_o = tmax;
    return tmax >= tmin && tmax >= 0.0f;
}

// ---------------------------------------------------------------- spherical harmonics (degree 0..3)
// include/spherical_harmonics.cuh:20-80 of the reference: real SH as polynomials of the (normalised) view direction, ALL-POSITIVE sign
// convention (not the PlenOctree signs of gaussian_model/sh_utils.py), coefficient order l = 0..degree, m = -l..l.
__device__ __forceinline__ void sh_basis_all(int degree, float x, float y, float z, float* Y)
{
    Y[0] = 0.28209479177387814f;
    if (degree >= 1) {
        Y[1] = 0.4886025119029199f * y; Y[2] = 0.4886025119029199f * z; Y[3] = 0.4886025119029199f * x;
    }
    if (degree >= 2) {
        Y[4] = 1.0925484305920792f * x * y; Y[5] = 1.0925484305920792f * y * z;
        Y[6] = 0.31539156525252005f * (3.0f * z * z - 1.0f);
        Y[7] = 1.0925484305920792f * x * z; Y[8] = 0.5462742152960396f * (x * x - y * y);
    }
    if (degree >= 3) {
        Y[9] = 0.5900435899266435f * y * (3.0f * x * x - y * y); Y[10] = 2.890611442640554f * x * y * z;
        Y[11] = 0.4570457994644658f * y * (5.0f * z * z - 1.0f); Y[12] = 0.3731763325901154f * z * (5.0f * z * z - 3.0f);
        Y[13] = 0.4570457994644658f * x * (5.0f * z * z - 1.0f); Y[14] = 1.445305721320277f * z * (x * x - y * y);
        Y[15] = 0.5900435899266435f * x * (x * x - 3.0f * y * y);
    }
}
// gradient of  sum_k f[k] Y_k(x, y, z)  with respect to (x, y, z), the polynomials taken as functions of three free variables
__device__ __forceinline__ void sh_grad_all(int degree, const float* f, float x, float y, float z, float& gx, float& gy, float& gz)
{
    gx = gy = gz = 0.f;
    if (degree >= 1) {
        gy += 0.4886025119029199f * f[1]; gz += 0.4886025119029199f * f[2]; gx += 0.4886025119029199f * f[3];
    }
    if (degree >= 2) {
        const float a = 1.0925484305920792f, b = 0.31539156525252005f, c = 0.5462742152960396f;
        gx += a * (f[4] * y + f[7] * z) + 2.0f * c * f[8] * x;
        gy += a * (f[4] * x + f[5] * z) - 2.0f * c * f[8] * y;
        gz += a * (f[5] * y + f[7] * x) + 6.0f * b * f[6] * z;
    }
    if (degree >= 3) {
        const float a = 0.5900435899266435f, b = 2.890611442640554f, c = 0.4570457994644658f, d = 0.3731763325901154f, e = 1.445305721320277f;
        const float zz = 5.0f * z * z - 1.0f;
        gx += a * f[9] * 6.0f * x * y + b * f[10] * y * z + c * f[13] * zz + e * f[14] * 2.0f * x * z + a * f[15] * 3.0f * (x * x - y * y);
        gy += a * f[9] * 3.0f * (x * x - y * y) + b * f[10] * x * z + c * f[11] * zz - e * f[14] * 2.0f * y * z - a * f[15] * 6.0f * x * y;
        gz += b * f[10] * x * y + c * f[11] * 10.0f * y * z + d * f[12] * (15.0f * z * z - 3.0f) + c * f[13] * 10.0f * x * z + e * f[14] * (x * x - y * y);
    }
}
// albedo of one (Gaussian, scan point) pair:  max(SH(dir(mu - o)) + 0.5, 0)  with dir = c / (|c| + 1e-8)  (volume_renderer.cu:107-111,
// cuda_utils.cuh:48-51).  c = mu - o, D = |c|.  Kept out of line: only scenes with active_sh_degree > 0 pay for it.
static __device__ __noinline__ float sh_pair_albedo(const float* __restrict__ f, int degree, float cx, float cy, float cz, float D)
{
    const float s = 1.0f / (D + 1e-8f);
    float Y[16];
    sh_basis_all(degree, cx * s, cy * s, cz * s, Y);
    const int K = (degree + 1) * (degree + 1);
    float r = 0.f;
    for (int i = 0; i < K; ++i) r = fmaf(__ldg(f + i), Y[i], r);
    return fmaxf(r + 0.5f, 0.0f);
}
// backward of the above for one pair: dpre = dL/d(SH value) (0 where the clamp is active).  Writes dsh[k] = dpre Y_k and ADDS dL/dmu to (gmx, gmy, gmz):
// d dir = s dc - dir (dir . dc) / D  with s = 1 / (D + 1e-8).
static __device__ __noinline__ void sh_pair_backward(const float* __restrict__ f, int degree, float cx, float cy, float cz, float D, float dpre,
                                              float* dsh, float& gmx, float& gmy, float& gmz)
{
    const float s = 1.0f / (D + 1e-8f);
    const float x = cx * s, y = cy * s, z = cz * s;
    float Y[16], fl[16];
    sh_basis_all(degree, x, y, z, Y);
    const int K = (degree + 1) * (degree + 1);
    for (int i = 0; i < 16; ++i) { fl[i] = i < K ? __ldg(f + i) : 0.f; dsh[i] = i < K ? dpre * Y[i] : 0.f; }
    float gx, gy, gz;
    sh_grad_all(degree, fl, x, y, z, gx, gy, gz);
    const float vd = (x * gx + y * gy + z * gz) / fmaxf(D, 1e-30f);
    gmx = fmaf(dpre, s * gx - x * vd, gmx);
    gmy = fmaf(dpre, s * gy - y * vd, gmy);
    gmz = fmaf(dpre, s * gz - z * vd, gmz);
}

// ---------------------------------------------------------------- fast math (SFU aware)
__device__ __forceinline__ float ex2_approx(float x)
{
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ float rcp_approx(float x)
{
    float y;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ float sqrt_approx(float x)
{
    float y;
    asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

// atan2 with |error| < 2e-6 rad (one MUFU.RCP + 6 FMA); used only for conservative
// footprint rectangles, which are widened by a far larger margin.
__device__ __forceinline__ float atan2_fast(float y, float x)
{
    float ax = fabsf(x), ay = fabsf(y);
    float mx = fmaxf(ax, ay), mn = fminf(ax, ay);
    float a = mn * rcp_approx(fmaxf(mx, 1e-30f));
    float s = a * a;
    float p = -0.01172120f;
    p = fmaf(p, s, 0.05265332f);
    p = fmaf(p, s, -0.11643287f);
    p = fmaf(p, s, 0.19354346f);
    p = fmaf(p, s, -0.33262347f);
    p = fmaf(p, s, 0.99997726f);
    float r = p * a;
    if (ay > ax) r = 1.57079632679f - r;
    if (x < 0.f) r = 3.14159265359f - r;
    return y < 0.f ? -r : r;
}

// atan2(y, x) for y > 0 with inv_y ~ 1 / y known: the same polynomial; the reciprocal that is still needed (of |x|, when |x| > y) does not depend
// on y, so it is off the dependency chain  sum of squares -> rsqrt -> y  of the caller
__device__ __forceinline__ float atan2_fast_inv(float y, float x, float inv_y)
{
    const float ax = fabsf(x);
    const bool steep = y >= ax;
    const float a = steep ? ax * inv_y : y * rcp_approx(fmaxf(ax, 1e-30f));
    const float s = a * a;
    float p = -0.01172120f;
    p = fmaf(p, s, 0.05265332f);
    p = fmaf(p, s, -0.11643287f);
    p = fmaf(p, s, 0.19354346f);
    p = fmaf(p, s, -0.33262347f);
    p = fmaf(p, s, 0.99997726f);
    float r = p * a;
    if (steep) r = 1.57079632679f - r;
    return x < 0.f ? 3.14159265359f - r : r;
}

// erf with |abs error| <= ~1.2e-7 : odd polynomial for |x|<1 , 1-exp(poly) beyond (CUDA-libm style split)
__device__ __forceinline__ float erf_fast(float x)
{
    return erff(x);
}

// packed fp32x2 arithmetic (sm_100: FMUL2 / FADD2 / FFMA2 -- two fp32 operations per issued instruction)
typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 pack2(float lo, float hi)
{
    f32x2 r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void unpack2(f32x2 v, float& lo, float& hi)
{
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ f32x2 mul2(f32x2 a, f32x2 b)
{
    f32x2 r;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
__device__ __forceinline__ f32x2 add2(f32x2 a, f32x2 b)
{
    f32x2 r;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c)
{
    f32x2 r;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
    return r;
}

// in-place forms (the accumulator is both source and destination: keeps unrolled register arrays in place)
__device__ __forceinline__ void acc_add2(f32x2& a, f32x2 b)
{
    asm("add.rn.f32x2 %0, %0, %1;" : "+l"(a) : "l"(b));
}
__device__ __forceinline__ void acc_mul2(f32x2& a, f32x2 b)
{
    asm("mul.rn.f32x2 %0, %0, %1;" : "+l"(a) : "l"(b));
}
__device__ __forceinline__ void acc_fma2(f32x2& c, f32x2 a, f32x2 b)
{
    asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(c) : "l"(a), "l"(b));
}

__device__ __forceinline__ float warp_sum(float v)
{
    v += __shfl_xor_sync(0xffffffffu, v, 16);
    v += __shfl_xor_sync(0xffffffffu, v, 8);
    v += __shfl_xor_sync(0xffffffffu, v, 4);
    v += __shfl_xor_sync(0xffffffffu, v, 2);
    v += __shfl_xor_sync(0xffffffffu, v, 1);
    return v;
}

// ---------------------------------------------------------------- TMA bulk copy + mbarrier (PTX)
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init()
{
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void fence_proxy_async()
{
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity)
{
    uint32_t addr = smem_u32(bar);
    uint32_t ok = 0;
    while (!ok) {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(ok)
            : "r"(addr), "r"(parity)
            : "memory");
    }
}
// 1-D TMA bulk copy global -> shared, completion signalled on an mbarrier (SASS: UBLKCP)
__device__ __forceinline__ void bulk_g2s(void* smem_dst, const void* gmem_src, uint32_t bytes, uint64_t* bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(smem_dst)),
                 "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}

}  // namespace nlos

// ---------------------------------------------------------------- host-side error plumbing
#include <stdio.h>
#include <string.h>
namespace nlos {
void set_error(const char* fmt, ...);
}
#define NLOS_CHECK_LAUNCH(name)                                                   \
    do {                                                                          \
        cudaError_t e__ = cudaGetLastError();                                     \
        if (e__ != cudaSuccess) {                                                 \
            nlos::set_error("%s: %s", name, cudaGetErrorString(e__));             \
            return -2;                                                            \
        }                                                                         \
    } while (0)
#define NLOS_REQUIRE(cond, ...)                                                   \
    do {                                                                          \
        if (!(cond)) {                                                            \
            nlos::set_error(__VA_ARGS__);                                         \
            return -1;                                                            \
        }                                                                         \
    } while (0)
