// transient.cu -- batched confocal transient renderer  H[scan point, time bin]  and its backward pass.
//
// One launch renders a whole slab of the relay-wall grid.  Per scan point the result equals what the
// reference produces with one CUDARenderModule.forward call per optimiser step
// (/root/reference/gaussian_model/cuda_autograd.py:254-314 -> submodules/cuda_renderer/src/volume_renderer.cu:16-185,
// no-occlusion branch :139-182) times volume_position[1]^2 (nlos_helpers.py:275-276), without the
// 256-Gaussians-per-ray cap (ray_aabb.cu:6,39).  The backward pass is what cuda_autograd.py:139-172 leaves as zeros.
//
// Work decomposition (B200: 148 SMs, one CTA of TR_WPB warps per SM):
//   CTA  = one scan tile (32 scan points) ; its lattice tables arrive with ONE TMA bulk copy
//   warp = one slice of the tile's entry list (binning.cu: entries ordered by the first bin of the Gaussian's shell)
//   lane = scan point.  Each warp owns a private LINEAR WINDOW [window bins][32] in shared memory (bank == lane ->
//   conflict-free without atomics) that slides with the list: bins that fall behind are flushed with coalesced float
//   REDs (forward) or refilled with the incoming dL/dH weights (backward).  Records are gathered with per-record TMA
//   bulk copies through a 2-stage mbarrier ring.
//
// Per (Gaussian, scan point)  -- "pair":
//   1. conservative (theta,phi) footprint rectangle from the 3-sigma AABB; every candidate ray gets the exact slab
//      test (same predicate as cuda_utils.cuh:97-121); hits are collected in a 32-bit mask per 4x8 block;
//   2. every hit ray is a Gaussian in r on the uniform bin grid:  e_k = exp2(al + gam (t_k - r*)^2).  It is evaluated
//      CENTRE-OUT from the pair's centre bin kc with the second-order geometric recurrence  e <- e g ; g <- g q
//      (q = exp2(2 gam dt^2)), the UP direction (bins kc, kc+1, ..) and the DOWN direction (kc-1, kc-2, ..) packed in
//      the two halves of one fp32x2 register (FMUL2/FADD2 of sm_100): 3 packed instructions per 2 sample-evals, 4 MUFU
//      per ray, and -- because the seed sits at the peak -- errors are relative to values that only shrink;
//   3. forward: the pair's bins live in REGISTERS (ACC[s] = {bin kc+s, bin kc-1-s}, static indices by full unrolling),
//      summed over all rays of the pair and merged into the shared window ONCE per pair;
//      backward: the three moments sum_k w_k e_k tau^n (n = 0,1,2) per ray from running prefix sums (A += w e; B += A;
//      C += B), chain rule to dL/dM, dL/dmu, dL/dop, dL/dalb per lane, ONE warp-shuffle reduction per entry, 14 REDs;
//   4. the analytic-exact mode integrates every bin with Gauss-Legendre nodes (2 or 4 by the ray's width in bins;
//      <= 1e-5 of the peak, far below the 1e-4 parity tolerance): each node is one more recurrence stream on the same
//      registers.  Rays narrower than 0.6 bins, or whose seeds are not representable, take the closed-form erf path
//      (analytic) / the direct 1-MUFU-per-bin path (numerical).
#include "common.cuh"
#include "../../include/nlosb200.h"

using namespace nlos;

#define TR_WPB 12          // warps per CTA
#define TR_MINB 1          // CTAs per SM the register allocation is sized for
#define TR_RING 4          // records per ring stage
#define TR_STAGES 2
#define TR_NH 32           // forward: register accumulators cover bins [kc - TR_NH, kc + TR_NH)
#define TR_GUARD 4         // zero bins on either side of a warp's window (backward reads up to 3 bins past an entry)
#define LOG2E_F 1.4426950408889634f
#define SQRT_PI_2 1.2533141373155001f
#define NO_BASE (-0x40000000)

enum { MODE_NUM = 0, MODE_ANA = 1, MODE_COUNT = 2 };

struct TransientParams {
    const float* tiles;
    int n_tiles, A;
    const float* records;
    int N;
    const unsigned* list;            // entries g | e << NLOS_ENTRY_SHIFT, tile-major, ordered by bucket
    const long long* cell_offsets;   // [n_tiles*n_buckets + 1]
    const int* tile_items;           // [n_tiles + 1] prefix of CTA work items
    float t0, dt, inv_dt;
    int T, window, bucket, n_buckets, slice_len;
    float flush_scale;   // c*deltaT for the numerical mode (volume_renderer.cu:182), 1 for the analytic-exact mode
    float mcut2;
    float* H;            // forward out   [n_tiles*32, T]  (accumulated with REDs; caller zeroes)
    const float* gH;     // backward in   [n_tiles*32, T]
    float* dacc;         // backward out  [N, 16]          (accumulated with REDs; caller zeroes)
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

// one hit ray of a pair: the Gaussian in r it restricts to (log2 domain) and its bin interval inside the entry
struct RayQ {
    float rs;      // r* : closest approach (Mahalanobis) along the ray
    float gam;     // -0.5 log2(e) C          (C = |M d|^2)
    float al;      // -0.5 log2(e) m2min
    float sth;     // sin(theta) quadrature weight of the ray
    float kl, kh;  // bins of the mcut cut clipped to the entry's sub-interval (kl > kh: nothing to do)
};

// moments of one packed {up, down} stream after n executed steps:  x_s = w e at step s,
//   A = sum x,  B = sum_s A_s = sum x (n - s),  C = sum_s B_s = sum x (n-s)(n-s+1)/2
// up bin kc+s has tau = tc + s dt, down bin kc-1-s has tau = tc - dt - s dt.
__device__ __forceinline__ void stream_moments(f32x2 A, f32x2 B, f32x2 C, float nf, float tc, float dt, float& S0, float& S1, float& S2)
{
    float Au, Ad, Bu, Bd, Cu, Cd;
    unpack2(A, Au, Ad); unpack2(B, Bu, Bd); unpack2(C, Cu, Cd);
    const float M1u = fmaf(nf, Au, -Bu), M1d = fmaf(nf, Ad, -Bd);                                  // sum x s
    const float M2u = fmaf(nf, fmaf(nf, Au, -2.0f * Bu), fmaf(2.0f, Cu, -Bu));                     // sum x s^2
    const float M2d = fmaf(nf, fmaf(nf, Ad, -2.0f * Bd), fmaf(2.0f, Cd, -Bd));
    const float td = tc - dt, dt2 = dt * dt;
    S0 += Au + Ad;
    S1 += fmaf(tc, Au, dt * M1u) + fmaf(td, Ad, -dt * M1d);
    S2 += fmaf(tc * tc, Au, fmaf(2.0f * tc * dt, M1u, dt2 * M2u)) + fmaf(td * td, Ad, fmaf(-2.0f * td * dt, M1d, dt2 * M2d));
}

template <int MODE, bool BWD>
__global__ void __launch_bounds__(TR_WPB * 32, TR_MINB) transient_kernel(const TransientParams P)
{
    constexpr bool FWD = !BWD && MODE != MODE_COUNT;
    constexpr bool GRAD = BWD && MODE != MODE_COUNT;
    constexpr int RC = (MODE == MODE_ANA) ? 1 : 2;          // hit rays consumed per iteration (analytic: one ray = 2 or 4 node streams)
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int A = P.A;
    const int W = P.window, BS = P.bucket, L = P.window - P.bucket;
    const int tile_floats = (NLOS_NTAB * A + NLOS_PINFO_ROWS) * NLOS_TILE;
    const int wstride = (W + 2 * TR_GUARD) * NLOS_TILE;
    float* s_tab = reinterpret_cast<float*>(smem_raw);
    float* s_pi = s_tab + NLOS_NTAB * A * NLOS_TILE;
    float* s_H = s_tab + tile_floats;                                   // [TR_WPB][guard | window | guard][32]
    float* s_ring = s_H + TR_WPB * wstride;                             // [TR_WPB][TR_STAGES][TR_RING][32]
    uint64_t* s_bar = reinterpret_cast<uint64_t*>(s_ring + TR_WPB * TR_STAGES * TR_RING * NLOS_REC_FLOATS);   // [1 + TR_WPB*TR_STAGES]

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int item = blockIdx.x;
    const int n_items = P.tile_items[P.n_tiles];
    if (item >= n_items) return;

    // tile of this CTA item: largest t with tile_items[t] <= item
    int tile;
    {
        int lo = 0, hi = P.n_tiles - 1;
        while (lo < hi) {
            int mid = (lo + hi + 1) >> 1;
            if (P.tile_items[mid] <= item) lo = mid; else hi = mid - 1;
        }
        tile = lo;
    }
    if (threadIdx.x == 0) {
        mbar_init(&s_bar[0], 1);
        for (int i = 0; i < TR_WPB * TR_STAGES; ++i) mbar_init(&s_bar[1 + i], 1);
        mbar_fence_init();
        mbar_arrive_expect_tx(&s_bar[0], (uint32_t)tile_floats * 4u);
        bulk_g2s(s_tab, P.tiles + (size_t)tile * tile_floats, (uint32_t)tile_floats * 4u, &s_bar[0]);
    }
    __syncthreads();

    // ---- this warp's slice of the tile's entry list
    const long long tile_beg = P.cell_offsets[(size_t)tile * P.n_buckets];
    const long long tile_tot = P.cell_offsets[(size_t)(tile + 1) * P.n_buckets] - tile_beg;
    const long long first = (long long)((item - P.tile_items[tile]) * TR_WPB + warp) * P.slice_len;
    const int n = first < tile_tot ? (int)min((long long)P.slice_len, tile_tot - first) : 0;
    const long long beg = tile_beg + first;
    mbar_wait(&s_bar[0], 0);   // lattice tables + scan-point info have landed
    if (n == 0) return;

    float* Hw = s_H + warp * wstride + TR_GUARD * NLOS_TILE;            // bin `base` of this warp's window
    float* ring = s_ring + warp * TR_STAGES * TR_RING * NLOS_REC_FLOATS;
    uint64_t* bar = s_bar + 1 + warp * TR_STAGES;

    // ---- per-lane scan point
    const float ox = s_pi[PI_OX * 32 + lane], oy = s_pi[PI_OY * 32 + lane], oz = s_pi[PI_OZ * 32 + lane];
    const int flags = __float_as_int(s_pi[PI_FLAGS * 32 + lane]);
    const bool valid = flags & 1, fast = flags & 2;
    const float t0 = P.t0, dt = P.dt, inv_dt = P.inv_dt, dt2 = P.dt * P.dt;
    const int T = P.T;

    if (MODE != MODE_COUNT)
        for (int i = lane; i < wstride; i += 32) Hw[i - TR_GUARD * NLOS_TILE] = 0.f;
    __syncwarp();

    // ---- linear window [base, base + W): when it advances, the bins that fall behind are flushed (forward) and the
    //      rest is shifted down; the backward pass refills the new bins with the weights dL/dH * hscale / (r^2+1e-8)
    const float* const s_hs = s_pi + PI_HSCALE * 32;
    int base = NO_BASE;
    auto flush = [&](int cnt) {                       // forward: H[p,k] += acc * hscale_p / (r_k^2 + 1e-8) * flush_scale  (cuda_autograd.py:308-314)
        for (int idx = lane; idx < cnt * NLOS_TILE; idx += 32) {
            const int pp = idx / cnt, kk = idx - pp * cnt, k = base + kk;
            const float v = Hw[kk * NLOS_TILE + pp];
            if (v != 0.f && k < T) {
                const float rk = fmaf((float)k, dt, t0);
                atomicAdd(P.H + ((size_t)tile * NLOS_TILE + pp) * T + k, v * (s_hs[pp] * P.flush_scale) / (rk * rk + 1e-8f));
            }
        }
    };
    auto refill = [&](int nbase, int kk_from, int cnt) {       // backward: W[k][p] = dL/dH[p,k] * hscale_p / (r_k^2 + 1e-8) * flush_scale
        for (int idx = lane; idx < cnt * NLOS_TILE; idx += 32) {
            const int pp = idx / cnt, kk = kk_from + (idx - pp * cnt), k = nbase + kk;
            float v = 0.f;
            if (k >= 0 && k < T) {
                const float rk = fmaf((float)k, dt, t0);
                v = P.gH[((size_t)tile * NLOS_TILE + pp) * T + k] * (s_hs[pp] * P.flush_scale) / (rk * rk + 1e-8f);
            }
            Hw[kk * NLOS_TILE + pp] = v;
        }
    };
    auto advance = [&](int nbase) {
        if (base == NO_BASE) {
            if (BWD) refill(nbase, 0, W);
        } else {
            const int d = min(nbase - base, W);
            if (!BWD) flush(d);
            __syncwarp();
            for (int k = 0; k < W - d; ++k) Hw[k * NLOS_TILE + lane] = Hw[(k + d) * NLOS_TILE + lane];
            if (!BWD)
                for (int k = W - d; k < W; ++k) Hw[k * NLOS_TILE + lane] = 0.f;
            __syncwarp();
            if (BWD) refill(nbase, W - d, d);
        }
        base = nbase;
        __syncwarp();
    };
    unsigned long long c_hits = 0, c_ws = 0, c_ev = 0, c_wb = 0;

    // ---- TMA ring over the list slice
    const int n_batches = (n + TR_RING - 1) / TR_RING;
    unsigned ent_next = 0, ent_cur = 0;
    auto issue = [&](int b) {
        int stage = b & 1;
        int cnt = min(TR_RING, n - b * TR_RING);
        if (lane == 0) {
            fence_proxy_async();
            mbar_arrive_expect_tx(&bar[stage], (uint32_t)cnt * NLOS_REC_FLOATS * 4u);
        }
        __syncwarp();
        ent_next = 0;
        if (lane < cnt) {
            ent_next = P.list[beg + b * TR_RING + lane];
            bulk_g2s(ring + (stage * TR_RING + lane) * NLOS_REC_FLOATS,
                     P.records + (size_t)(ent_next & ((1u << NLOS_ENTRY_SHIFT) - 1u)) * NLOS_REC_FLOATS, NLOS_REC_FLOATS * 4u, &bar[stage]);
        }
    };
    issue(0);
    for (int b = 0; b < n_batches; ++b) {
        ent_cur = ent_next;
        if (b + 1 < n_batches) issue(b + 1);   // (overwrites ent_next; ent_cur keeps this batch's entries)
        mbar_wait(&bar[b & 1], (b >> 1) & 1);
        const int cnt = min(TR_RING, n - b * TR_RING);
        for (int q = 0; q < cnt; ++q) {
            const unsigned ent = __shfl_sync(0xffffffffu, ent_cur, q);
            const int esub = (int)(ent >> NLOS_ENTRY_SHIFT);
            const float4* rec = reinterpret_cast<const float4*>(ring + ((b & 1) * TR_RING + q) * NLOS_REC_FLOATS);
            const float4 v0 = rec[0], v1 = rec[1], v2 = rec[2], v3 = rec[3], v4 = rec[4], v5 = rec[5];
            const float M00 = v1.x, M01 = v1.y, M02 = v1.z, M10 = v1.w, M11 = v2.x, M12 = v2.y, M20 = v2.z, M21 = v2.w, M22 = v3.x;
            const float opalb = v0.w * v4.w;
            const float cx = v0.x - ox, cy = v0.y - oy, cz = v0.z - oz;                             // mu - o
            // ---- the tile-level shell of this Gaussian, bit-identical to binning.cu:shell_bins (warp min/max over the scan points)
            int s_lo, s_hi;
            bool split;
            float d2;
            {
                d2 = __fadd_rn(__fadd_rn(__fmul_rn(cx, cx), __fmul_rn(cy, cy)), __fmul_rn(cz, cz));
                const unsigned umin = __reduce_min_sync(0xffffffffu, valid ? __float_as_uint(d2) : 0x7f800000u);
                const unsigned umax = __reduce_max_sync(0xffffffffu, valid ? __float_as_uint(d2) : 0u);
                const float rlo = __fsub_rn(__fsqrt_rn(__uint_as_float(umin)), v5.x), rhi = __fadd_rn(__fsqrt_rn(__uint_as_float(umax)), v5.x);
                const float klo_f = floorf(__fadd_rn(__fmul_rn(__fsub_rn(rlo, t0), inv_dt), 0.5f));
                const float khi_f = floorf(__fadd_rn(__fmul_rn(__fsub_rn(rhi, t0), inv_dt), 0.5f));
                const int klo_t = klo_f < 0.0f ? 0 : (klo_f > (float)(T - 1) ? T : (int)klo_f);
                const int khi_t = khi_f < 0.0f ? -1 : (khi_f > (float)(T - 1) ? T - 1 : (int)khi_f);
                s_lo = klo_t + esub * L;                                                         // this entry's sub-interval
                s_hi = min(khi_t, s_lo + L - 1);
                split = khi_t - klo_t + 1 > L;                                                   // the Gaussian has more than one entry
            }
            if (s_lo > s_hi) continue;                                                           // (cannot happen for a listed entry)
            if (MODE != MODE_COUNT) {
                const int nbase = (s_lo / BS) * BS;
                if (nbase != base) advance(nbase);
            }
            const float fRlo = (float)s_lo, fRhi = (float)s_hi;
            const float r_in = fmaf(fRlo - 0.5f, dt, t0), r_out = fmaf(fRhi + 0.5f, dt, t0);    // bin k covers [t0+(k-.5)dt, t0+(k+.5)dt]
            // accumulators of the backward pass for this pair
            float gax = 0.f, gay = 0.f, gaz = 0.f;                                                  // dL/da
            float g00 = 0.f, g01 = 0.f, g02 = 0.f, g10 = 0.f, g11 = 0.f, g12 = 0.f, g20 = 0.f, g21 = 0.f, g22 = 0.f;   // sum dL/db (x) d
            float gwa = 0.f;                                                                        // dL/d(op*alb)
            bool act = valid && (MODE == MODE_COUNT || opalb > 0.f);
            float D = 0.f;
            if (act) {
                // radial reject of this scan point against the entry's bins
                D = sqrtf(d2);
                act = (D + v5.x >= r_in) && (D - v5.x <= r_out);
            }
            if (act) {
                const float lox = v3.y - ox, loy = v3.z - oy, loz = v3.w - oz;                      // AABB relative to the scan point
                const float hix = v4.x - ox, hiy = v4.y - oy, hiz = v4.z - oz;
                const float ax = -(M00 * cx + M01 * cy + M02 * cz), ay = -(M10 * cx + M11 * cy + M12 * cz),
                            az = -(M20 * cx + M21 * cy + M22 * cz);                                 // a = M (o - mu)
                // ---- conservative footprint rectangle in lattice index space
                int i0 = 0, i1 = A - 1, j0 = 0, j1 = A - 1;
                if (fast) {
                    const float th0 = s_pi[PI_TH0 * 32 + lane], idth = s_pi[PI_IDTH * 32 + lane];
                    const float ph0 = s_pi[PI_PH0 * 32 + lane], idph = s_pi[PI_IDPH * 32 + lane];
                    float qx = fmaxf(fmaxf(lox, -hix), 0.f), qy = fmaxf(fmaxf(loy, -hiy), 0.f);
                    float rmin = sqrtf(qx * qx + qy * qy);
                    float bx = fmaxf(fabsf(lox), fabsf(hix)), by = fmaxf(fabsf(loy), fabsf(hiy));
                    float rmax = sqrtf(bx * bx + by * by);
                    float thlo = atan2_fast(hiz > 0.f ? rmin : rmax, hiz);
                    float thhi = atan2_fast(loz > 0.f ? rmax : rmin, loz);
                    i0 = clamp_ceil((thlo - th0) * idth - 0.02f, 0, A);
                    i1 = clamp_floor((thhi - th0) * idth + 0.02f, -1, A - 1);
                    if (loy > 0.f) {
                        float phhi = atan2_fast(lox > 0.f ? hiy : loy, lox);
                        float phlo = atan2_fast(hix > 0.f ? loy : hiy, hix);
                        j0 = clamp_ceil((phlo - ph0) * idph - 0.02f, 0, A);
                        j1 = clamp_floor((phhi - ph0) * idph + 0.02f, -1, A - 1);
                    }
                }
                // ---- the pair's centre bin: up bins kc + s (s < n_up), down bins kc - 1 - s (s < n_dn)
                const int kc = (int)rintf(fminf(fmaxf((D - t0) * inv_dt, fRlo), fRhi + 1.0f));
                const int n_up = s_hi - kc + 1, n_dn = kc - s_lo;
                const float fkc = (float)kc;
                const float tkc = fmaf(fkc, dt, t0);
                float* const hpc = Hw + (kc - base) * NLOS_TILE + lane;      // bin kc ; bin kc+s at hpc[32 s], bin kc-1-s at hpc[-32 (s+1)]
                f32x2 ACC[FWD ? TR_NH : 1];
                int nflush = 0;
                if (FWD) {
#pragma unroll
                    for (int s = 0; s < TR_NH; ++s) ACC[s] = 0ull;
                }

                // ---- ray -> Gaussian in r (shared by all modes)
                auto ray_setup = [&](int i, int j, bool has, RayQ& R, float& bx, float& by, float& bz) {
                    const float sth = s_tab[(TAB_STH * A + i) * 32 + lane], cth = s_tab[(TAB_CTH * A + i) * 32 + lane];
                    const float sph = s_tab[(TAB_SPH * A + j) * 32 + lane], cph = s_tab[(TAB_CPH * A + j) * 32 + lane];
                    const float dx = sth * cph, dy = sth * sph, dz = cth;
                    bx = M00 * dx + M01 * dy + M02 * dz; by = M10 * dx + M11 * dy + M12 * dz; bz = M20 * dx + M21 * dy + M22 * dz;   // b = M d
                    const float C = bx * bx + by * by + bz * bz;
                    const float ab = ax * bx + ay * by + az * bz;
                    const float iC = rcp_nr(C);
                    const float rsu = -ab * iC;                                             // closest approach r*
                    const float px = fmaf(rsu, bx, ax), py = fmaf(rsu, by, ay), pz = fmaf(rsu, bz, az);
                    const float m2min = px * px + py * py + pz * pz;
                    const float h2 = (P.mcut2 - m2min) * iC;
                    const float hh = h2 * rsqrt_approx(fmaxf(h2, 1e-30f)) * inv_dt;          // half width of the cut in bins
                    const float xc = (rsu - t0) * inv_dt;
                    const float off = (MODE == MODE_ANA) ? 0.5f : 0.0f;                        // bins are [k-.5, k+.5] for the integral
                    const float klu = fmaxf(ceilf(xc - hh - off), fRlo), khu = fminf(floorf(xc + hh + off), fRhi);
                    const bool ok = has && (h2 > 0.f) && (klu <= khu);
                    R.rs = rsu;
                    R.gam = -0.5f * LOG2E_F * C;
                    R.al = -0.5f * LOG2E_F * m2min;
                    R.sth = sth;
                    R.kl = ok ? klu : 3.0e9f;
                    R.kh = ok ? khu : -3.0e9f;
                };
                // ---- chain rule of one ray from its moments S_n = sum_k w_k e_k tau_k^n  (tau = r_k - r*)
                auto ray_chain = [&](int i, int j, float rsu, float S0, float S1, float S2) {
                    const float sth = s_tab[(TAB_STH * A + i) * 32 + lane], cth = s_tab[(TAB_CTH * A + i) * 32 + lane];
                    const float sph = s_tab[(TAB_SPH * A + j) * 32 + lane], cph = s_tab[(TAB_CPH * A + j) * 32 + lane];
                    const float dx = sth * cph, dy = sth * sph, dz = cth;
                    const float bx = M00 * dx + M01 * dy + M02 * dz, by = M10 * dx + M11 * dy + M12 * dz, bz = M20 * dx + M21 * dy + M22 * dz;
                    const float px = fmaf(rsu, bx, ax), py = fmaf(rsu, by, ay), pz = fmaf(rsu, bz, az);
                    gwa = fmaf(sth, S0, gwa);
                    // dL/da = -wt (S0 perp + S1 b) ; dL/db = -wt (r* (S0 perp + S1 b) + S1 perp + S2 b)
                    const float ux = S0 * px + S1 * bx, uy = S0 * py + S1 * by, uz = S0 * pz + S1 * bz;
                    const float vx = fmaf(rsu, ux, S1 * px + S2 * bx), vy = fmaf(rsu, uy, S1 * py + S2 * by), vz = fmaf(rsu, uz, S1 * pz + S2 * bz);
                    const float w_ = opalb * sth;
                    gax = fmaf(-w_, ux, gax); gay = fmaf(-w_, uy, gay); gaz = fmaf(-w_, uz, gaz);
                    const float bxw = -w_ * vx, byw = -w_ * vy, bzw = -w_ * vz;
                    g00 = fmaf(bxw, dx, g00); g01 = fmaf(bxw, dy, g01); g02 = fmaf(bxw, dz, g02);
                    g10 = fmaf(byw, dx, g10); g11 = fmaf(byw, dy, g11); g12 = fmaf(byw, dz, g12);
                    g20 = fmaf(bzw, dx, g20); g21 = fmaf(bzw, dy, g21); g22 = fmaf(bzw, dz, g22);
                };
                // ---- seeds of one stream at the centre bin.  tc = t(kc) + node offset - r*
                //      E = {e(kc), e(kc-1)} * weight ; G = {e(kc+1)/e(kc), e(kc-2)/e(kc-1)} ; returns q = e-ratio of ratios
                auto seed = [&](float gam, float al, float tc, float weight, f32x2& E, f32x2& G, f32x2& Q) {
                    const float gd = gam * dt;
                    const float e0 = ex2_approx(fmaf(gam * tc, tc, al)) * weight;
                    const float gu = ex2_approx(gd * fmaf(2.0f, tc, dt));
                    const float gm = ex2_approx(gd * fmaf(-2.0f, tc, dt));
                    const float qv = gu * gm;
                    E = pack2(e0, e0 * gm);
                    G = pack2(gu, gm * qv);
                    Q = pack2(qv, qv);
                };
                // ---- forward: n steps (rounded up to 4) of two streams into the register accumulators
                auto run_fwd = [&](f32x2 E0, f32x2 G0, f32x2 Q0, f32x2 E1, f32x2 G1, f32x2 Q1, int nst) {
#pragma unroll
                    for (int s = 0; s < (FWD ? TR_NH : 0); ++s) {
                        if ((s & 3) == 0 && s >= nst) break;
                        ACC[s] = add2(ACC[s], add2(E0, E1));
                        E0 = mul2(E0, G0); E1 = mul2(E1, G1);
                        G0 = mul2(G0, Q0); G1 = mul2(G1, Q1);
                    }
                    nflush = max(nflush, (nst + 3) & ~3);
                };
                // ---- backward: prefix-sum moments of two streams against the window's weights
                //      returns the number of executed steps
                auto run_bwd = [&](f32x2& E0, f32x2 G0, f32x2 Q0, f32x2& E1, f32x2 G1, f32x2 Q1, int nst, f32x2& A0, f32x2& B0, f32x2& C0, f32x2& A1,
                                   f32x2& B1, f32x2& C1) -> int {
                    A0 = 0ull; B0 = 0ull; C0 = 0ull; A1 = 0ull; B1 = 0ull; C1 = 0ull;
                    const float* pu = hpc;
                    const float* pd = hpc - NLOS_TILE;
                    if (!split) {
                        // reads run at most 3 bins past the entry (guard bins / bins beyond the Gaussian's shell); a direction that would
                        // leave the window + guard is switched off at a 4-step boundary
                        const int zu = (n_up + TR_GUARD) & ~3, zd = (n_dn + TR_GUARD) & ~3;
                        const int zmin = min(zu, zd);
                        int s = 0;
                        for (; s < nst; s += 4) {
                            if (s >= zmin) {
                                float a, b;
                                unpack2(E0, a, b); E0 = pack2(s >= zu ? 0.f : a, s >= zd ? 0.f : b);
                                unpack2(E1, a, b); E1 = pack2(s >= zu ? 0.f : a, s >= zd ? 0.f : b);
                                if (s >= zu) pu = hpc;                      // keep the (ignored) reads inside the window
                                if (s >= zd) pd = hpc;
                            }
#pragma unroll
                            for (int jx = 0; jx < 4; ++jx) {
                                const f32x2 w2 = pack2(pu[jx * NLOS_TILE], pd[-jx * NLOS_TILE]);
                                const f32x2 x0 = mul2(w2, E0), x1 = mul2(w2, E1);
                                A0 = add2(A0, x0); A1 = add2(A1, x1);
                                B0 = add2(B0, A0); B1 = add2(B1, A1);
                                C0 = add2(C0, B0); C1 = add2(C1, B1);
                                E0 = mul2(E0, G0); E1 = mul2(E1, G1);
                                G0 = mul2(G0, Q0); G1 = mul2(G1, Q1);
                            }
                            pu += 4 * NLOS_TILE; pd -= 4 * NLOS_TILE;
                        }
                        return s;
                    } else {
                        // split entries: exact clipping to the entry's sub-interval, step by step
                        const int ns = min(nst, max(n_up, n_dn));
                        for (int s = 0; s < ns; ++s) {
                            const float wu = s < n_up ? pu[0] : 0.f, wd = s < n_dn ? pd[0] : 0.f;
                            const f32x2 w2 = pack2(wu, wd);
                            const f32x2 x0 = mul2(w2, E0), x1 = mul2(w2, E1);
                            A0 = add2(A0, x0); A1 = add2(A1, x1);
                            B0 = add2(B0, A0); B1 = add2(B1, A1);
                            C0 = add2(C0, B0); C1 = add2(C1, B1);
                            E0 = mul2(E0, G0); E1 = mul2(E1, G1);
                            G0 = mul2(G0, Q0); G1 = mul2(G1, Q1);
                            if (s + 1 < n_up) pu += NLOS_TILE;
                            if (s + 1 < n_dn) pd -= NLOS_TILE;
                        }
                        return ns;
                    }
                };
                // ---- direct evaluation of bins [ka, kb] of one ray (numerical: 1 MUFU per bin; analytic: closed-form erf per bin edge)
                //      forward: into the window; backward: exact moments
                auto direct = [&](const RayQ& R, int ka, int kb, float& S0, float& S1, float& S2) {
                    if (ka > kb) return;
                    const float wt = opalb * R.sth;
                    if (MODE == MODE_NUM) {
                        float t = fmaf((float)ka, dt, t0) - R.rs;
                        float* hp = Hw + (ka - base) * NLOS_TILE + lane;
                        for (int k = ka; k <= kb; ++k) {
                            const float ev = ex2_approx(fmaf(R.gam, t * t, R.al));
                            if (!BWD) {
                                *hp = fmaf(wt, ev, *hp);
                            } else {
                                const float we = *hp * ev, wet = we * t;
                                S0 += we; S1 += wet; S2 = fmaf(wet, t, S2);
                            }
                            t += dt;
                            hp += NLOS_TILE;
                        }
                    } else if (MODE == MODE_ANA) {
                        // integral of exp(-C tau^2 / 2) e^{-m2min/2} over bin k = pref [erf(sq tau1) - erf(sq tau0)]  
                        const float C = R.gam * (-2.0f / LOG2E_F);
                        const float sq = sqrtf(0.5f * C), iC = 1.0f / C;
                        const float pk = ex2_approx(R.al);                                       // e^{-m2min/2}
                        const float pref = SQRT_PI_2 * rsqrt_approx(C) * pk;
                        float tau = fmaf((float)ka - 0.5f, dt, t0) - R.rs;                       // lower edge of bin ka
                        float Ep = erf_fast(sq * tau);
                        float fp = BWD ? pk * ex2_approx(R.gam * tau * tau) : 0.f;
                        float* hp = Hw + (ka - base) * NLOS_TILE + lane;
                        for (int k = ka; k <= kb; ++k) {
                            const float tn = tau + dt;
                            const float En = erf_fast(sq * tn);
                            const float Ik = pref * (En - Ep);
                            if (!BWD) {
                                *hp = fmaf(wt, Ik, *hp);
                            } else {
                                // int tau f = -[f]/C ; int tau^2 f = (int f - [tau f]) / C
                                const float fn = pk * ex2_approx(R.gam * tn * tn);
                                const float w = *hp;
                                S0 = fmaf(w, Ik, S0);
                                S1 = fmaf(w, -(fn - fp) * iC, S1);
                                S2 = fmaf(w, (Ik - (tn * fn - tau * fp)) * iC, S2);
                                fp = fn;
                            }
                            Ep = En; tau = tn;
                            hp += NLOS_TILE;
                        }
                    }
                };

                int wb_lo = 0x7fffffff, wb_hi = -1;
                for (int ib = i0; ib <= i1; ib += 4)
                    for (int jb = j0; jb <= j1; jb += 8) {
                        // ---- 1. exact slab test of every candidate ray of this 4x8 block -> hit mask (bit = 8*row + column)
                        unsigned mask = 0;
                        const int ni = min(4, i1 - ib + 1), nj = min(8, j1 - jb + 1);
                        for (int ii = 0; ii < ni; ++ii) {
                            const float tth = s_tab[(TAB_TTH * A + ib + ii) * 32 + lane];
                            const float tz0 = loz * tth, tz1 = hiz * tth;
                            const float zmn = fminf(tz0, tz1), zmx = fmaxf(tz0, tz1);
                            for (int jj = 0; jj < nj; ++jj) {
                                // cuda_utils.cuh:97-121 with every t multiplied by sin(theta) > 0
                                const float icp = s_tab[(TAB_ICPH * A + jb + jj) * 32 + lane], isp = s_tab[(TAB_ISPH * A + jb + jj) * 32 + lane];
                                const float tx0 = lox * icp, tx1 = hix * icp, ty0 = loy * isp, ty1 = hiy * isp;
                                const float tmin = fmaxf(fmaxf(fminf(tx0, tx1), fminf(ty0, ty1)), zmn);
                                const float tmax = fminf(fminf(fmaxf(tx0, tx1), fmaxf(ty0, ty1)), zmx);
                                const bool hit = (tmax >= tmin) && (tmax >= 0.f);
                                mask |= (hit ? 1u : 0u) << (ii * 8 + jj);
                                if (MODE == MODE_COUNT && hit) {
                                    // slab interval stretched about its midpoint by 5.3/3 (t = t'/sin(theta))
                                    const float ist = 1.0f / s_tab[(TAB_STH * A + ib + ii) * 32 + lane];
                                    const float mid = 0.5f * (tmin + tmax) * ist, half = 0.5f * (tmax - tmin) * ist * (5.3f / 3.0f);
                                    const int ka = clamp_ceil((mid - half - t0) * inv_dt, s_lo, s_hi + 1),
                                              kb = clamp_floor((mid + half - t0) * inv_dt, s_lo - 1, s_hi);
                                    c_hits += (r_in <= mid && mid < r_out) ? 1 : 0;
                                    if (kb >= ka) { c_ws += kb - ka + 1; wb_lo = min(wb_lo, ka); wb_hi = max(wb_hi, kb); }
                                }
                            }
                        }
                        // ---- 2. consume the hits
                        while (mask) {
                            RayQ R[RC];
                            int bits[RC];
                            bool any = false;
                            int bit = 0;
#pragma unroll
                            for (int u = 0; u < RC; ++u) {
                                const bool has = mask != 0;
                                if (has) { bit = __ffs(mask) - 1; mask &= mask - 1; }
                                bits[u] = bit;
                                float bx, by, bz;
                                ray_setup(ib + (bit >> 3), jb + (bit & 7), has, R[u], bx, by, bz);
                                any = any || (R[u].kl <= R[u].kh);
                            }
                            if (!any) continue;
                            if (MODE == MODE_COUNT) {
#pragma unroll
                                for (int u = 0; u < RC; ++u)
                                    if (R[u].kl <= R[u].kh) {
                                        const int hu = max((int)R[u].kh - kc + 1, kc - (int)R[u].kl);
                                        c_ev += (unsigned long long)(2 * ((max(hu, 0) + 3) & ~3));
                                    }
                            } else if (MODE == MODE_NUM) {
                                // ---- numerical mode: two rays = two streams
                                f32x2 E[2], G[2], Q[2];
                                bool fr[2];
                                int nst = 0;
                                const int cap = FWD ? TR_NH : 0x7fffffff;
#pragma unroll
                                for (int u = 0; u < 2; ++u) {
                                    const bool vu = R[u].kl <= R[u].kh;
                                    const float tc = tkc - R[u].rs;
                                    const float Lc = fmaf(R[u].gam * tc, tc, R[u].al);
                                    fr[u] = vu && (Lc > -100.f) && (R[u].gam * dt2 > -6.f);       // seed representable, recurrence well-conditioned
                                    seed(R[u].gam, R[u].al, tc, BWD ? 1.0f : opalb * R[u].sth, E[u], G[u], Q[u]);
                                    if (!fr[u]) { E[u] = 0ull; G[u] = 0ull; }
                                    const int hu = fr[u] ? max((int)R[u].kh - kc + 1, kc - (int)R[u].kl) : 0;
                                    nst = max(nst, min(hu, cap));
                                }
                                float S0[2] = {0.f, 0.f}, S1[2] = {0.f, 0.f}, S2[2] = {0.f, 0.f};
                                if (nst > 0) {
                                    if (!BWD) {
                                        run_fwd(E[0], G[0], Q[0], E[1], G[1], Q[1], nst);
                                    } else {
                                        f32x2 A0, B0, C0, A1, B1, C1;
                                        const int ne = run_bwd(E[0], G[0], Q[0], E[1], G[1], Q[1], nst, A0, B0, C0, A1, B1, C1);
                                        stream_moments(A0, B0, C0, (float)ne, tkc - R[0].rs, dt, S0[0], S1[0], S2[0]);
                                        stream_moments(A1, B1, C1, (float)ne, tkc - R[1].rs, dt, S0[1], S1[1], S2[1]);
                                    }
                                }
#pragma unroll
                                for (int u = 0; u < 2; ++u) {
                                    if (!(R[u].kl <= R[u].kh)) continue;
                                    const int kl = (int)R[u].kl, kh = (int)R[u].kh;
                                    if (!fr[u]) {
                                        direct(R[u], kl, kh, S0[u], S1[u], S2[u]);
                                    } else if (FWD) {                                      // bins beyond the register accumulators
                                        direct(R[u], kl, min(kh, kc - TR_NH - 1), S0[u], S1[u], S2[u]);
                                        direct(R[u], max(kl, kc + TR_NH), kh, S0[u], S1[u], S2[u]);
                                    }
                                    if (BWD) ray_chain(ib + (bits[u] >> 3), jb + (bits[u] & 7), R[u].rs, S0[u], S1[u], S2[u]);
                                }
                            } else {
                                // ---- analytic-exact mode: one ray, Gauss-Legendre nodes +-x (pairs) per bin
                                const RayQ& r = R[0];
                                const int kl = (int)r.kl, kh = (int)r.kh;
                                const float gp = -r.gam * dt2;                               // 0.72 / (sigma in bins)^2
                                const float tc = tkc - r.rs;
                                const float tfar = fabsf(tc) + 0.5f * dt;
                                const bool fr = (gp <= 2.0f) && (fmaf(r.gam * tfar, tfar, r.al) > -100.f);
                                float S0 = 0.f, S1 = 0.f, S2 = 0.f;
                                if (!fr) {
                                    direct(r, kl, kh, S0, S1, S2);
                                } else {
                                    const int npass = gp <= 0.08f ? 1 : 2;                   // GL2 (sigma >= 3 bins) or GL4
                                    const int cap = FWD ? TR_NH : 0x7fffffff;
                                    const int nst = min(max(kh - kc + 1, kc - kl), cap);
                                    const float wbase = (BWD ? 1.0f : opalb * r.sth) * dt;
                                    for (int ps = 0; ps < npass; ++ps) {
                                        const float xn = (npass == 1 ? 0.28867513459481287f : (ps == 0 ? 0.16999052179242816f : 0.43056815579702626f)) * dt;
                                        const float wn = (npass == 1 ? 0.5f : (ps == 0 ? 0.32607257743127305f : 0.17392742256872692f)) * wbase;
                                        f32x2 E0, G0, Q0, E1, G1, Q1;
                                        seed(r.gam, r.al, tc + xn, wn, E0, G0, Q0);
                                        seed(r.gam, r.al, tc - xn, wn, E1, G1, Q1);
                                        if (!BWD) {
                                            run_fwd(E0, G0, Q0, E1, G1, Q1, nst);
                                        } else {
                                            f32x2 A0, B0, C0, A1, B1, C1;
                                            const int ne = run_bwd(E0, G0, Q0, E1, G1, Q1, nst, A0, B0, C0, A1, B1, C1);
                                            stream_moments(A0, B0, C0, (float)ne, tc + xn, dt, S0, S1, S2);
                                            stream_moments(A1, B1, C1, (float)ne, tc - xn, dt, S0, S1, S2);
                                        }
                                    }
                                    if (FWD) {
                                        direct(r, kl, min(kh, kc - TR_NH - 1), S0, S1, S2);
                                        direct(r, max(kl, kc + TR_NH), kh, S0, S1, S2);
                                    }
                                }
                                if (BWD) ray_chain(ib + (bits[0] >> 3), jb + (bits[0] & 7), r.rs, S0, S1, S2);
                            }
                        }
                    }
                if (MODE == MODE_COUNT && wb_hi >= wb_lo) c_wb += wb_hi - wb_lo + 1;
                // ---- forward: merge the pair's register accumulators into the window (clipped to the entry's sub-interval)
                if (FWD && nflush > 0) {
                    const int nu = min(nflush, n_up), nd = min(nflush, n_dn);
                    const int nm = max(nu, nd);
#pragma unroll
                    for (int s = 0; s < (FWD ? TR_NH : 0); ++s) {
                        if ((s & 3) == 0 && s >= nm) break;
                        float lo, hi;
                        unpack2(ACC[s], lo, hi);
                        if (s < nu) hpc[s * NLOS_TILE] += lo;
                        if (s < nd) hpc[-(s + 1) * NLOS_TILE] += hi;
                    }
                }
            }
            if (GRAD) {
                // a = M (o - mu) = -M c :  dL/dM += dL/da (x) (o - mu) ;  dL/dmu = -M^T dL/da
                float o[13];
                o[0] = fmaf(-gax, cx, g00); o[1] = fmaf(-gax, cy, g01); o[2] = fmaf(-gax, cz, g02);
                o[3] = fmaf(-gay, cx, g10); o[4] = fmaf(-gay, cy, g11); o[5] = fmaf(-gay, cz, g12);
                o[6] = fmaf(-gaz, cx, g20); o[7] = fmaf(-gaz, cy, g21); o[8] = fmaf(-gaz, cz, g22);
                o[9] = -(M00 * gax + M10 * gay + M20 * gaz);
                o[10] = -(M01 * gax + M11 * gay + M21 * gaz);
                o[11] = -(M02 * gax + M12 * gay + M22 * gaz);
                o[12] = gwa;
#pragma unroll
                for (int i = 0; i < 13; ++i) o[i] = warp_sum(o[i]);
                const int g = (int)(ent & ((1u << NLOS_ENTRY_SHIFT) - 1u));
                if (lane < 14) {
                    // slots: 0..8 dM, 9..11 dmu, 12 dop = alb * d(op*alb), 13 dalb = op * d(op*alb)
                    float val = 0.f;
#pragma unroll
                    for (int i = 0; i < 12; ++i) val = (lane == i) ? o[i] : val;
                    if (lane == 12) val = o[12] * v4.w;
                    if (lane == 13) val = o[12] * v0.w;
                    if (val != 0.f) atomicAdd(P.dacc + (size_t)g * NLOS_GRAD_FLOATS + lane, val);
                }
            }
        }
        __syncwarp();
    }

    if (MODE == MODE_COUNT) {
        for (int o = 16; o > 0; o >>= 1) {
            c_hits += __shfl_xor_sync(0xffffffffu, c_hits, o);
            c_ws += __shfl_xor_sync(0xffffffffu, c_ws, o);
            c_ev += __shfl_xor_sync(0xffffffffu, c_ev, o);
            c_wb += __shfl_xor_sync(0xffffffffu, c_wb, o);
        }
        if (lane == 0) {
            atomicAdd(P.counters + 0, c_hits);
            atomicAdd(P.counters + 1, c_ws);
            atomicAdd(P.counters + 2, c_ev);
            atomicAdd(P.counters + 3, c_wb);
        }
        return;
    }
    if (!BWD && base != NO_BASE) {
        __syncwarp();
        flush(W);
    }
}

static size_t transient_smem_bytes(int A, int window)
{
    size_t f = (size_t)(NLOS_NTAB * A + NLOS_PINFO_ROWS) * NLOS_TILE + (size_t)TR_WPB * (window + 2 * TR_GUARD) * NLOS_TILE +
               (size_t)TR_WPB * TR_STAGES * TR_RING * NLOS_REC_FLOATS;
    return f * 4 + (1 + TR_WPB * TR_STAGES) * 8 + 16;
}

template <int MODE, bool BWD>
static int launch_transient(const TransientParams& P, int n_cta_items, cudaStream_t s, const char* name)
{
    size_t smem = transient_smem_bytes(P.A, P.window);
    NLOS_REQUIRE(smem <= 227 * 1024, "%s: %zu B of shared memory needed (A=%d, window=%d) > 227 KB", name, smem, P.A, P.window);
    cudaError_t e = cudaFuncSetAttribute(transient_kernel<MODE, BWD>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    NLOS_REQUIRE(e == cudaSuccess, "%s: cudaFuncSetAttribute: %s", name, cudaGetErrorString(e));
    transient_kernel<MODE, BWD><<<n_cta_items, TR_WPB * 32, smem, s>>>(P);
    NLOS_CHECK_LAUNCH(name);
    return 0;
}

static int fill_params(TransientParams& P, const float* tiles, int n_tiles, int A, const float* records, int N, const int* list,
                       const long long* cell_offsets, const int* tile_items, float t0, float dt, int T, int window, int bucket, int slice_len,
                       float c_deltaT, float mcut, int mode)
{
    NLOS_REQUIRE(n_tiles > 0 && N > 0 && T > 0 && A >= 1 && A <= 128, "transient: bad sizes");
    NLOS_REQUIRE(window >= 32 && window % 8 == 0 && bucket >= 1 && 2 * bucket <= window && slice_len > 0,
                 "transient: bad window/bucket/slice_len");
    NLOS_REQUIRE(mode == MODE_NUM || mode == MODE_ANA, "transient: mode must be 0 (numerical) or 1 (analytic-exact)");
    NLOS_REQUIRE(tiles && records && list && cell_offsets && tile_items, "transient: null pointer");
    NLOS_REQUIRE(((uintptr_t)tiles % 16) == 0 && ((uintptr_t)records % 16) == 0, "transient: tiles/records must be 16-byte aligned");
    P.tiles = tiles; P.n_tiles = n_tiles; P.A = A; P.records = records; P.N = N; P.list = reinterpret_cast<const unsigned*>(list);
    P.cell_offsets = cell_offsets; P.tile_items = tile_items;
    P.t0 = t0; P.dt = dt; P.inv_dt = 1.0f / dt; P.T = T; P.window = window; P.bucket = bucket; P.n_buckets = (T + bucket - 1) / bucket;
    P.slice_len = slice_len; P.flush_scale = mode == MODE_NUM ? c_deltaT : 1.0f; P.mcut2 = mcut * mcut;
    P.H = nullptr; P.gH = nullptr; P.dacc = nullptr; P.counters = nullptr;
    return 0;
}

extern "C" int nlosb200_transient_warps_per_block(void) { return TR_WPB; }

extern "C" int nlosb200_transient_forward(const float* tiles, int n_tiles, int A, const float* records, int N, const int* list,
                                          const long long* cell_offsets, const int* tile_items, int n_cta_items, float t0, float dt, int T,
                                          int window, int bucket, int slice_len, float c_deltaT, float mcut, int mode, float* H, void* stream)
{
    TransientParams P;
    int rc = fill_params(P, tiles, n_tiles, A, records, N, list, cell_offsets, tile_items, t0, dt, T, window, bucket, slice_len, c_deltaT, mcut,
                         mode);
    if (rc) return rc;
    NLOS_REQUIRE(H, "transient_forward: null H");
    cudaStream_t s = (cudaStream_t)stream;
    cudaError_t e = cudaMemsetAsync(H, 0, (size_t)n_tiles * NLOS_TILE * T * sizeof(float), s);
    NLOS_REQUIRE(e == cudaSuccess, "transient_forward: memset: %s", cudaGetErrorString(e));
    if (n_cta_items <= 0) return 0;
    P.H = H;
    return mode == MODE_NUM ? launch_transient<MODE_NUM, false>(P, n_cta_items, s, "transient_kernel<num,fwd>")
                            : launch_transient<MODE_ANA, false>(P, n_cta_items, s, "transient_kernel<ana,fwd>");
}

extern "C" int nlosb200_transient_backward(const float* tiles, int n_tiles, int A, const float* records, int N, const int* list,
                                           const long long* cell_offsets, const int* tile_items, int n_cta_items, float t0, float dt, int T,
                                           int window, int bucket, int slice_len, float c_deltaT, float mcut, int mode, const float* grad_H,
                                           float* dacc, void* stream)
{
    TransientParams P;
    int rc = fill_params(P, tiles, n_tiles, A, records, N, list, cell_offsets, tile_items, t0, dt, T, window, bucket, slice_len, c_deltaT, mcut,
                         mode);
    if (rc) return rc;
    NLOS_REQUIRE(grad_H && dacc, "transient_backward: null pointer");
    cudaStream_t s = (cudaStream_t)stream;
    cudaError_t e = cudaMemsetAsync(dacc, 0, (size_t)N * NLOS_GRAD_FLOATS * sizeof(float), s);
    NLOS_REQUIRE(e == cudaSuccess, "transient_backward: memset: %s", cudaGetErrorString(e));
    if (n_cta_items <= 0) return 0;
    P.gH = grad_H;
    P.dacc = dacc;
    return mode == MODE_NUM ? launch_transient<MODE_NUM, true>(P, n_cta_items, s, "transient_kernel<num,bwd>")
                            : launch_transient<MODE_ANA, true>(P, n_cta_items, s, "transient_kernel<ana,bwd>");
}

// work counters of SURVEY.md 8(d): counters[0] hit (g,ray) pairs, [1] W_s, [2] sample-evals this library evaluates, [3] W_b
extern "C" int nlosb200_transient_count(const float* tiles, int n_tiles, int A, const float* records, int N, const int* list,
                                        const long long* cell_offsets, const int* tile_items, int n_cta_items, float t0, float dt, int T,
                                        int window, int bucket, int slice_len, float mcut, unsigned long long* counters, void* stream)
{
    TransientParams P;
    int rc = fill_params(P, tiles, n_tiles, A, records, N, list, cell_offsets, tile_items, t0, dt, T, window, bucket, slice_len, 1.0f, mcut,
                         MODE_NUM);
    if (rc) return rc;
    NLOS_REQUIRE(counters, "transient_count: null counters");
    cudaStream_t s = (cudaStream_t)stream;
    cudaError_t e = cudaMemsetAsync(counters, 0, 4 * sizeof(unsigned long long), s);
    NLOS_REQUIRE(e == cudaSuccess, "transient_count: memset: %s", cudaGetErrorString(e));
    if (n_cta_items <= 0) return 0;
    P.counters = counters;
    return launch_transient<MODE_COUNT, false>(P, n_cta_items, s, "transient_kernel<count>");
}
