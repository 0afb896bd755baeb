// transient_shared.cuh -- parameters and device helpers shared by the two scan-tile kernels of the batched transient renderer
// (transient.cu: register-accumulator kernel for wide footprints; transient_narrow.cu: direct-to-window kernel for footprints of a few rays).
#pragma once
#include "common.cuh"
#include "../../include/nlosb200.h"

using namespace nlos;

#ifndef TR_WPB
#define TR_WPB 12          // warps per CTA
#endif
#define TR_MINB 1          // CTAs per SM the register allocation is sized for
#define TR_RING 4          // records per ring stage
#define TR_STAGES 2
#define TR_NH 32           // forward: register accumulators cover bins [kc - TR_NH, kc + TR_NH)
#define TR_GUARD 4         // zero bins on either side of a warp's window (backward reads up to 3 bins past an entry)
#define LOG2E_F 1.4426950408889634f
#define SQRT_PI_2 1.2533141373155001f
#define NO_BASE (-0x40000000)
// analytic mode: gpm = 0.72 / (ray width in bins)^2 up to which the 1-node / 2-node moment-matched mixtures are used
#ifndef TR_GPM_K0
#define TR_GPM_K0 0.115f   // >= 2.5 bins
#endif
#ifndef TR_GPM_K1
#define TR_GPM_K1 1.125f   // >= 0.8 bins
#endif

enum { MODE_NUM = 0, MODE_ANA = 1, MODE_COUNT = 2 };

struct TransientParams {
    const float* tiles;
    int n_tiles, A;
    const float* records;
    int N;
    const unsigned* list;            // entries g | e << NLOS_ENTRY_SHIFT, tile-major, ordered by bucket
    const unsigned* list_bins;       // per entry: tile-level shell klo | khi << 16 (binning.cu:shell_fill_kernel)
    const long long* cell_offsets;   // [n_tiles*cpt + 1], cpt = n_classes * n_buckets cells per tile (class-major); this launch walks cells [cell0, cell0 + n_buckets)
    int cpt;
    // a launch walks one or two SEGMENTS = entry classes rendered by the same kernel: cells [seg_cell0[s], seg_cell0[s] + n_buckets) of every tile
    int n_seg, seg_cell0[2];
    long long list_capacity;         // entries the list buffers hold
    int* work;                       // [n_seg*n_tiles] next slice of every (segment, tile) entry list (zeroed before the launch)
    float t0, dt, inv_dt;
    int T, window, bucket, n_buckets, slice_len;
    float flush_scale;   // c*deltaT for the numerical mode (volume_renderer.cu:182), 1 for the analytic-exact mode
    float mcut2;
    int seg_one_pass[2]; // analytic mode, time-wide entries: every ray is a 1-node ray, H / gH are the UNWEIGHTED bin rows (no 1/(r^2+1e-8)):
                         // the fixed 7-tap filter of analytic_filter_kernel and the radial weight are applied to the whole row outside
    float* seg_H[2];        // forward out per segment  [n_tiles*32, T]  (accumulated with REDs; caller zeroes)
    const float* seg_gH[2]; // backward in per segment  [n_tiles*32, T]
    float* dacc;         // backward out  [N, 16]          (accumulated with REDs; caller zeroes)
    const float* sh;     // raw SH coefficients [N, sh_dim] (used when sh_degree > 0: per (Gaussian, scan point) albedo, volume_renderer.cu:107-111)
    int sh_dim, sh_degree;
    size_t part_stride;  // 0, or N*16: deterministic mode -- dacc / dsh are per-TILE partial buffers [n_tiles][N][16], summed in tile order afterwards
    float* dsh;          // backward out [N, 16]: dL/d(SH coefficients) when sh_degree > 0 (REDs; caller zeroes)
    unsigned short* hit_stash;      // narrow kernel: per (list position, scan point) culling words, written by the forward and read by the backward of
                                    // the same lists (null: every pass decides for itself); needs A <= 64
    unsigned char* hit_dump;        // tests only (nlosb200_debug_hit_dump): dense slab-test membership [n_tiles*32][A*A][N], or null
    unsigned long long* counters;   // MODE_COUNT: [0] hit (g,ray) pairs, [1] W_s, [2] evaluated sample-evals, [3] W_b
};

__device__ __forceinline__ int clamp_ceil(float x, int lo, int hi)
{
    x = ceilf(x);
    return x <= (float)lo ? lo : (x >= (float)hi ? hi : (int)x);
}
__device__ __forceinline__ int clamp_floor(float x, int lo, int hi)
{
    x = floorf(x);
    return x <= (float)lo ? lo : (x >= (float)hi ? hi : (int)x);
}
// Conservative lattice-index bounds of the footprint rectangle without FRND / F2I (both quarter-rate XU instructions; the four clamps of a pair
// were 4.7 % of the narrow forward's stall samples): rint by the 1.5 * 2^23 trick.  lattice_lo(x) <= ceil(x) and lattice_hi(y) >= floor(y)
// (equal unless x is within 0.01 above / y within 0.01 below an integer: one more candidate row or column, decided by the slab test as ever).
__device__ __forceinline__ int rint_small(float x)                    // (int)rintf(x) for |x| < 2^22, NaN -> garbage: callers clamp first
{
    return __float_as_int(__fadd_rn(x, 12582912.0f)) - 0x4B400000;
}
__device__ __forceinline__ int lattice_lo(float x, int hi)            // in [0, hi]
{
    x = fminf(fmaxf(x, 0.f), (float)hi) + 0.49f;
    return __float_as_int(__fadd_rn(x, 12582912.0f)) - 0x4B400000;
}
__device__ __forceinline__ int lattice_hi(float y, int hi)            // in [-1, hi]
{
    y = fminf(fmaxf(y, -1.f), (float)hi) - 0.49f;
    return __float_as_int(__fadd_rn(y, 12582912.0f)) - 0x4B400000;
}
// 1/x to ~1 ulp: MUFU.RCP + one Newton step instead of the ~10-instruction IEEE division
__device__ __forceinline__ float rcp_nr(float x)
{
    float r = rcp_approx(x);
    return r * fmaf(-x, r, 2.0f);
}
__device__ __forceinline__ float rsqrt_approx(float x)
{
    float y;
    asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ float lo_of(f32x2 v) { float a, b; unpack2(v, a, b); return a; }
__device__ __forceinline__ float hi_of(f32x2 v) { float a, b; unpack2(v, a, b); return b; }

// moments of one packed {up, down} stream after n executed steps:  x_s = w e at step s,
//   A = sum x,  B = sum_s A_s = sum x (n - s),  C = sum_s B_s = sum x (n-s)(n-s+1)/2
// up bin kc+s has tau = tc + s dt, down bin kc-1-s has tau = tc - dt - s dt.
__device__ __forceinline__ void stream_moments(f32x2 A, f32x2 B, f32x2 C, float nf, float tc, float dt, float& S0, float& S1, float& S2)
{
    // evaluated on the packed {up, down} halves, then the two halves are added
    const f32x2 n2 = pack2(nf, nf), neg1 = pack2(-1.0f, -1.0f), two = pack2(2.0f, 2.0f);
    const f32x2 nB = mul2(B, neg1);
    const f32x2 M1 = fma2(n2, A, nB);                                                       // sum x s
    const f32x2 M2 = fma2(n2, fma2(n2, A, add2(nB, nB)), fma2(two, C, nB));                 // sum x s^2
    const float td = tc - dt, dt2 = dt * dt;
    const f32x2 T1 = pack2(tc, td), DT = pack2(dt, -dt);
    const f32x2 T2 = mul2(T1, T1), TT = mul2(mul2(T1, DT), two);
    const f32x2 s1 = fma2(T1, A, mul2(DT, M1));
    const f32x2 s2 = fma2(T2, A, fma2(TT, M1, mul2(pack2(dt2, dt2), M2)));
    float a0, a1, b0, b1, c0, c1;
    unpack2(A, a0, a1); unpack2(s1, b0, b1); unpack2(s2, c0, c1);
    S0 += a0 + a1;
    S1 += b0 + b1;
    S2 += c0 + c1;
}

// transposing butterfly: 16 + 1 shuffles reduce 16 per-lane values over the warp (instead of 16 x 5); after the step with lane-xor 16/8/4/2
// each lane keeps half of its slots, so slot index = lane bits 4..1 and lanes 2i, 2i+1 both return the total of slot i
__device__ __forceinline__ float warp_reduce16(float (&v)[16], int lane)
{
#pragma unroll
    for (int half = 8, bit = 16; half >= 1; half >>= 1, bit >>= 1) {
        const bool up = lane & bit;
#pragma unroll
        for (int i = 0; i < half; ++i) {
            const float send = up ? v[i] : v[i + half], keep = up ? v[i + half] : v[i];
            v[i] = keep + __shfl_xor_sync(0xffffffffu, send, bit);
        }
    }
    return v[0] + __shfl_xor_sync(0xffffffffu, v[0], 1);
}

// ---- host side, shared by the two kernels
static inline size_t transient_smem_bytes(int A, int window)
{
    size_t f = (size_t)(NLOS_NTAB * A + NLOS_PINFO_ROWS) * NLOS_TILE + (size_t)TR_WPB * (window + 2 * TR_GUARD) * NLOS_TILE +
               (size_t)TR_WPB * TR_STAGES * TR_RING * NLOS_REC_FLOATS;
    return f * 4 + (1 + TR_WPB * TR_STAGES) * 8 + 32 + 16 + (size_t)TR_WPB * TR_STAGES * TR_RING * NLOS_TILE * 2;   // (+ the culling-word staging)
}
static inline int transient_num_sms()
{
    static int n_sm = 0;
    if (n_sm == 0) {
        int dev = 0;
        cudaGetDevice(&dev);
        if (cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n_sm <= 0) n_sm = 148;
    }
    return n_sm;
}
// transient_narrow.cu
int launch_transient_narrow(const TransientParams& P, int mode, bool bwd, cudaStream_t s);
