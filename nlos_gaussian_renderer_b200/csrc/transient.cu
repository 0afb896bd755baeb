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
//   lane = scan point.  Each warp owns a private SLIDING WINDOW accumulator [window bins][32] in shared memory
//   (ring indexed by bin mod window, bank == lane -> conflict-free read-modify-write without atomics) that
//   advances with the list: bins that fall behind are flushed with coalesced float REDs (forward) or refilled with
//   the incoming dL/dH weights (backward).  A Gaussian is therefore set up ONCE per (tile, scan point) however many
//   bins it covers.  Records are gathered with per-record TMA bulk copies through a 2-stage mbarrier ring.
//
// Per (Gaussian, scan point):
//   1. conservative (theta,phi) footprint rectangle from the 3-sigma AABB; every candidate ray gets the exact slab
//      test (same predicate as cuda_utils.cuh:97-121); hits are collected in a 32-bit mask per 4x8 block;
//   2. hits are consumed RC at a time ("chunk"): centred quadratic m2(r) = m2min + C (r - r*)^2, bin interval of
//      the 5.3-sigma cut; ONE loop over the union of the chunk's bin intervals in which every ray advances
//      e_{k+1} = e_k g_k,  g_{k+1} = g_k q  (a Gaussian on a uniform grid is a second-order geometric recurrence,
//      q = exp(-C dt^2)), re-seeded with MUFU.EX2 every TR_SEG bins, evaluated two rays at a time with the packed
//      fp32x2 instructions of sm_100 (FMUL2/FADD2/FFMA2): 1.5 issue slots per sample-eval and ONE
//      shared-memory read-modify-write per (chunk, bin).  Rays the recurrence cannot carry (seed underflow,
//      Gaussians much narrower than a bin) take a direct 1-MUFU-per-bin path;
//   3. backward: per ray three moments sum_k w_k e_k k^n (n = 0,1,2) in registers, chain rule to dL/dM, dL/dmu,
//      dL/dop, dL/dalb per lane, ONE warp-shuffle reduction per entry, 14 float REDs.
#include "common.cuh"
#include "../../include/nlosb200.h"

using namespace nlos;

#define TR_WPB 12          // warps per CTA
#define TR_MINB 1          // CTAs per SM the register allocation is sized for
#define TR_RING 4          // records per ring stage
#define TR_STAGES 2
#define TR_SEG 16          // bins between MUFU re-seeds of a recurrence
#define LOG2E_F 1.4426950408889634f
#define TWO_OVER_SQRT_PI 1.1283791670955126f

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

template <int MODE, bool BWD>
struct ChunkWidth {
    static constexpr int value = (MODE == MODE_ANA && BWD) ? 2 : 4;     // rays per row chunk (register budget)
};


template <int MODE, bool BWD>
__global__ void __launch_bounds__(TR_WPB * 32, TR_MINB) transient_kernel(const TransientParams P)
{
    constexpr int RC = ChunkWidth<MODE, BWD>::value;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int A = P.A;
    const int W = P.window, BS = P.bucket, L = P.window - P.bucket;
    const unsigned wmagic = 0xFFFFFFFFu / (unsigned)P.window + 1u;
    auto wrap = [&](int k) { return k - (int)__umulhi((unsigned)k, wmagic) * W; };      // k mod window for 0 <= k < 2^20
    const int tile_floats = (NLOS_NTAB * A + NLOS_PINFO_ROWS) * NLOS_TILE;
    float* s_tab = reinterpret_cast<float*>(smem_raw);
    float* s_pi = s_tab + NLOS_NTAB * A * NLOS_TILE;
    float* s_H = s_tab + tile_floats;                                   // [TR_WPB][window][32]
    float* s_ring = s_H + TR_WPB * W * NLOS_TILE;                       // [TR_WPB][TR_STAGES][TR_RING][32]
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

    float* Hw = s_H + warp * W * NLOS_TILE;
    float* ring = s_ring + warp * TR_STAGES * TR_RING * NLOS_REC_FLOATS;
    uint64_t* bar = s_bar + 1 + warp * TR_STAGES;

    // ---- per-lane scan point
    const float ox = s_pi[PI_OX * 32 + lane], oy = s_pi[PI_OY * 32 + lane], oz = s_pi[PI_OZ * 32 + lane];
    const float th0 = s_pi[PI_TH0 * 32 + lane], idth = s_pi[PI_IDTH * 32 + lane];
    const float ph0 = s_pi[PI_PH0 * 32 + lane], idph = s_pi[PI_IDPH * 32 + lane];
    const int flags = __float_as_int(s_pi[PI_FLAGS * 32 + lane]);
    const bool valid = flags & 1, fast = flags & 2;
    const float t0 = P.t0, dt = P.dt, inv_dt = P.inv_dt, dt2 = P.dt * P.dt;
    const int T = P.T;

    if (MODE != MODE_COUNT && !BWD)
        for (int i = lane; i < W * NLOS_TILE; i += 32) Hw[i] = 0.f;
    __syncwarp();

    // ---- sliding window [base, base + W): flush (forward) / refill (backward) the bins that change when it advances
    const float* const s_hs = s_pi + PI_HSCALE * 32;
    auto flush = [&](int k_from, int cnt) {          // forward: H[p,k] += acc * hscale_p / (r_k^2 + 1e-8) * flush_scale  (cuda_autograd.py:308-314)
        for (int idx = lane; idx < cnt * NLOS_TILE; idx += 32) {
            const int pp = idx / cnt, k = k_from + (idx - pp * cnt);
            float* slot = Hw + wrap(k) * NLOS_TILE + pp;
            const float v = *slot;
            *slot = 0.f;
            if (v != 0.f && k < T) {
                const float rk = fmaf((float)k, dt, t0);
                atomicAdd(P.H + ((size_t)tile * NLOS_TILE + pp) * T + k, v * (s_hs[pp] * P.flush_scale) / (rk * rk + 1e-8f));
            }
        }
    };
    auto refill = [&](int k_from, int cnt) {         // backward: W[k][p] = dL/dH[p,k] * hscale_p / (r_k^2 + 1e-8) * flush_scale
        for (int idx = lane; idx < cnt * NLOS_TILE; idx += 32) {
            const int pp = idx / cnt, k = k_from + (idx - pp * cnt);
            float v = 0.f;
            if (k >= 0 && k < T) {
                const float rk = fmaf((float)k, dt, t0);
                v = P.gH[((size_t)tile * NLOS_TILE + pp) * T + k] * (s_hs[pp] * P.flush_scale) / (rk * rk + 1e-8f);
            }
            Hw[wrap(k) * NLOS_TILE + pp] = v;
        }
    };
    int base = -0x40000000;                           // "no window yet"
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
            }
            if (s_lo > s_hi) continue;                                                           // (cannot happen for a listed entry)
            if (MODE != MODE_COUNT) {
                const int nbase = (s_lo / BS) * BS;
                if (nbase != base) {
                    if (!BWD) {
                        if (base > -0x40000000) flush(base, min(nbase - base, W));
                    } else {
                        const int from = max(base + W, nbase);
                        refill(from, nbase + W - from);
                    }
                    base = nbase;
                    __syncwarp();
                }
            }
            const float fRlo = (float)s_lo, fRhi = (float)s_hi;
            const float r_in = fmaf(fRlo - 0.5f, dt, t0), r_out = fmaf(fRhi + 0.5f, dt, t0);    // bin k covers [t0+(k-.5)dt, t0+(k+.5)dt]
            // accumulators of the backward pass for this (Gaussian, scan point)
            float gax = 0.f, gay = 0.f, gaz = 0.f;                                                  // dL/da
            float g00 = 0.f, g01 = 0.f, g02 = 0.f, g10 = 0.f, g11 = 0.f, g12 = 0.f, g20 = 0.f, g21 = 0.f, g22 = 0.f;   // sum dL/db (x) d
            float gwa = 0.f;                                                                        // dL/d(op*alb)
            bool act = valid && (MODE == MODE_COUNT || opalb > 0.f);
            if (act) {
                // radial reject of this scan point against the entry's bins
                const float D = sqrtf(d2);
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
                        // ---- 2. consume the hits RC at a time
                        while (mask) {
                            float rs[RC], gam[RC], al[RC], kl[RC], kh[RC], wt[RC];
                            int bits[RC];
                            float kminf = 3.0e9f, kmaxf = -3.0e9f;
                            int bit = 0;
#pragma unroll
                            for (int u = 0; u < RC; ++u) {
                                const bool has = mask != 0;
                                if (has) { bit = __ffs(mask) - 1; mask &= mask - 1; }
                                bits[u] = bit;
                                const int i = ib + (bit >> 3), j = jb + (bit & 7);
                                const float sth = s_tab[(TAB_STH * A + i) * 32 + lane], cth = s_tab[(TAB_CTH * A + i) * 32 + lane];
                                const float sph = s_tab[(TAB_SPH * A + j) * 32 + lane], cph = s_tab[(TAB_CPH * A + j) * 32 + lane];
                                const float dx = sth * cph, dy = sth * sph, dz = cth;
                                const float bx = M00 * dx + M01 * dy + M02 * dz, by = M10 * dx + M11 * dy + M12 * dz,
                                            bz = M20 * dx + M21 * dy + M22 * dz;                            // b = M d
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
                                rs[u] = rsu;
                                gam[u] = -0.5f * LOG2E_F * C;
                                al[u] = -0.5f * LOG2E_F * m2min;
                                wt[u] = opalb * sth;
                                kl[u] = ok ? klu : 3.0e9f;
                                kh[u] = ok ? khu : -3.0e9f;
                                kminf = fminf(kminf, kl[u]);
                                kmaxf = fmaxf(kmaxf, kh[u]);
                            }
                            if (kminf > kmaxf) continue;
                            const int kmin = (int)kminf, kmax = (int)kmaxf;
                            const float tk0 = fmaf(kminf, dt, t0);
                            if (MODE == MODE_COUNT) {
#pragma unroll
                                for (int u = 0; u < RC; ++u)
                                    if (kl[u] <= kh[u]) c_ev += (unsigned long long)(kmax - kmin + 1);
                            } else if (MODE == MODE_NUM) {
                                // ---- numerical mode: recurrence  e_{k+1} = e_k g_k ; g_{k+1} = g_k q   per ray, common bin loop
                                float e[RC], g[RC], qq[RC];
                                float S0[RC], S1[RC], S2[RC];
                                bool fr[RC];
                                bool anyfast = false, anyslow = false;
#pragma unroll
                                for (int u = 0; u < RC; ++u) {
                                    const float tu = tk0 - rs[u];
                                    const float Lg = fmaf(gam[u] * tu, tu, al[u]);
                                    const bool vu = kl[u] <= kh[u];
                                    fr[u] = vu && (Lg > -100.f) && (gam[u] * dt2 > -6.f);     // seed representable, recurrence well-conditioned
                                    anyfast = anyfast || fr[u];
                                    anyslow = anyslow || (vu && !fr[u]);
                                    qq[u] = ex2_approx(2.0f * gam[u] * dt2);
                                    S0[u] = 0.f; S1[u] = 0.f; S2[u] = 0.f;
                                }
                                const float kfc = 0.5f * (kmaxf - kminf);                    // moments are taken about the middle of the chunk's bins
                                if (anyfast) {
                                    float kf = -kfc;
                                    const f32x2 q01 = pack2(qq[0], qq[1]), q23 = pack2(qq[2], qq[3]);
                                    f32x2 A01 = 0ull, A23 = 0ull, B01 = 0ull, B23 = 0ull, C01 = 0ull, C23 = 0ull;     // packed moments
                                    for (int ks = kmin; ks <= kmax;) {
                                        const int nseg = min(min(TR_SEG, kmax - ks + 1), W - wrap(ks));      // never across the ring's wrap
                                        const float tks = fmaf((float)ks, dt, t0);
                                        float* hp = Hw + wrap(ks) * NLOS_TILE + lane;
#pragma unroll
                                        for (int u = 0; u < RC; ++u) {
                                            const float tu = tks - rs[u];
                                            const float gt = gam[u] * tu;
                                            const float ev = ex2_approx(fmaf(gt, tu, al[u]));
                                            const float gv = ex2_approx(fmaf(2.0f * gt, dt, gam[u] * dt2));
                                            e[u] = fr[u] ? (BWD ? ev : wt[u] * ev) : 0.f;
                                            g[u] = fr[u] ? gv : 0.f;
                                        }
                                        // packed fp32x2: rays (0,1) and (2,3) share one FMUL2 / FADD2 / FFMA2 each
                                        f32x2 e01 = pack2(e[0], e[1]), e23 = pack2(e[2], e[3]);
                                        f32x2 g01 = pack2(g[0], g[1]), g23 = pack2(g[2], g[3]);
                                        if (!BWD) {
#pragma unroll 4
                                            for (int kk = 0; kk < nseg; ++kk) {
                                                float lo, hi;
                                                unpack2(add2(e01, e23), lo, hi);
                                                e01 = mul2(e01, g01); e23 = mul2(e23, g23);
                                                g01 = mul2(g01, q01); g23 = mul2(g23, q23);
                                                *hp += lo + hi;
                                                hp += NLOS_TILE;
                                            }
                                        } else {
#pragma unroll 2
                                            for (int kk = 0; kk < nseg; ++kk) {
                                                const float w = *hp;
                                                const f32x2 w2 = pack2(w, w), k2 = pack2(kf, kf);
                                                const f32x2 we01 = mul2(w2, e01), we23 = mul2(w2, e23);
                                                const f32x2 wk01 = mul2(we01, k2), wk23 = mul2(we23, k2);
                                                A01 = add2(A01, we01); A23 = add2(A23, we23);
                                                B01 = add2(B01, wk01); B23 = add2(B23, wk23);
                                                C01 = fma2(wk01, k2, C01); C23 = fma2(wk23, k2, C23);
                                                e01 = mul2(e01, g01); e23 = mul2(e23, g23);
                                                g01 = mul2(g01, q01); g23 = mul2(g23, q23);
                                                kf += 1.0f;
                                                hp += NLOS_TILE;
                                            }
                                        }
                                        ks += nseg;
                                    }
                                    if (BWD) {
                                        unpack2(A01, S0[0], S0[1]); unpack2(A23, S0[2], S0[3]);
                                        unpack2(B01, S1[0], S1[1]); unpack2(B23, S1[2], S1[3]);
                                        unpack2(C01, S2[0], S2[1]); unpack2(C23, S2[2], S2[3]);
                                    }
                                }
                                if (BWD) {
                                    // moments about the chunk centre -> moments of tau = r_k - r*
                                    const float tcen = fmaf(kfc, dt, tk0);
#pragma unroll
                                    for (int u = 0; u < RC; ++u) {
                                        const float tc = tcen - rs[u];
                                        const float m0 = S0[u], m1 = S1[u] * dt, m2 = S2[u] * dt2;
                                        S1[u] = fmaf(tc, m0, m1);
                                        S2[u] = fmaf(tc, fmaf(tc, m0, 2.0f * m1), m2);
                                    }
                                }
                                if (anyslow) {
                                    // direct evaluation, 1 MUFU per bin, for rays the recurrence cannot carry
#pragma unroll
                                    for (int u = 0; u < RC; ++u) {
                                        if (!(kl[u] <= kh[u]) || fr[u]) continue;
                                        const int klo = (int)kl[u], khi = (int)kh[u];
                                        float t = fmaf((float)klo, dt, t0) - rs[u];
                                        float a0 = 0.f, a1 = 0.f, a2 = 0.f;
                                        for (int k = klo; k <= khi; ++k) {
                                            float* hp = Hw + wrap(k) * NLOS_TILE + lane;
                                            const float ev = ex2_approx(fmaf(gam[u], t * t, al[u]));
                                            if (!BWD) {
                                                *hp = fmaf(wt[u], ev, *hp);
                                            } else {
                                                const float we = *hp * ev, wet = we * t;
                                                a0 += we; a1 += wet; a2 = fmaf(wet, t, a2);
                                            }
                                            t += dt;
                                        }
                                        if (BWD) { S0[u] = a0; S1[u] = a1; S2[u] = a2; }
                                    }
                                }
                                if (BWD) {
#pragma unroll
                                    for (int u = 0; u < RC; ++u) {
                                        if (!(kl[u] <= kh[u])) continue;
                                        const int i = ib + (bits[u] >> 3), j = jb + (bits[u] & 7);
                                        const float sth = s_tab[(TAB_STH * A + i) * 32 + lane], cth = s_tab[(TAB_CTH * A + i) * 32 + lane];
                                        const float sph = s_tab[(TAB_SPH * A + j) * 32 + lane], cph = s_tab[(TAB_CPH * A + j) * 32 + lane];
                                        const float dx = sth * cph, dy = sth * sph, dz = cth;
                                        const float bx = M00 * dx + M01 * dy + M02 * dz, by = M10 * dx + M11 * dy + M12 * dz,
                                                    bz = M20 * dx + M21 * dy + M22 * dz;
                                        const float px = fmaf(rs[u], bx, ax), py = fmaf(rs[u], by, ay), pz = fmaf(rs[u], bz, az);
                                        gwa = fmaf(sth, S0[u], gwa);
                                        // dL/da = -wt (S0 perp + S1 b) ; dL/db = -wt (r* (S0 perp + S1 b) + S1 perp + S2 b)
                                        const float ux = S0[u] * px + S1[u] * bx, uy = S0[u] * py + S1[u] * by, uz = S0[u] * pz + S1[u] * bz;
                                        const float vx = fmaf(rs[u], ux, S1[u] * px + S2[u] * bx), vy = fmaf(rs[u], uy, S1[u] * py + S2[u] * by),
                                                    vz = fmaf(rs[u], uz, S1[u] * pz + S2[u] * bz);
                                        const float w_ = wt[u];
                                        gax = fmaf(-w_, ux, gax); gay = fmaf(-w_, uy, gay); gaz = fmaf(-w_, uz, gaz);
                                        const float bxw = -w_ * vx, byw = -w_ * vy, bzw = -w_ * vz;
                                        g00 = fmaf(bxw, dx, g00); g01 = fmaf(bxw, dy, g01); g02 = fmaf(bxw, dz, g02);
                                        g10 = fmaf(byw, dx, g10); g11 = fmaf(byw, dy, g11); g12 = fmaf(byw, dz, g12);
                                        g20 = fmaf(bzw, dx, g20); g21 = fmaf(bzw, dy, g21); g22 = fmaf(bzw, dz, g22);
                                    }
                                }
                            } else {
                                // ---- MODE_ANA : exact bin integrals, erf shared between adjacent bin edges
                                float sq[RC], pref[RC], u0[RC], Ep[RC];
                                float Dp[RC], DUp[RC], F[RC], P1[RC], P2[RC];
                                const float ue = fmaf(-0.5f, dt, tk0);                           // lower edge of bin kmin
#pragma unroll
                                for (int u = 0; u < RC; ++u) {
                                    const bool vu = kl[u] <= kh[u];
                                    const float C = gam[u] * (-2.0f / LOG2E_F);
                                    sq[u] = vu ? sqrtf(0.5f * C) : 0.f;
                                    pref[u] = vu ? 1.2533141373155001f * rsqrt_approx(C) * ex2_approx(al[u]) : 0.f;     // sqrt(pi/(2C)) e^{-m2min/2}
                                    u0[u] = ue - rs[u];
                                    const float z = sq[u] * u0[u];
                                    Ep[u] = erf_fast(z);
                                    if (BWD) {
                                        Dp[u] = TWO_OVER_SQRT_PI * ex2_approx(-LOG2E_F * z * z);
                                        DUp[u] = Dp[u] * u0[u];
                                        F[u] = 0.f; P1[u] = 0.f; P2[u] = 0.f;
                                    }
                                }
                                {
                                    float kf = 1.0f;
                                    if (!BWD) {
#pragma unroll 2
                                        for (int k = kmin; k <= kmax; ++k) {
                                            float acc = 0.f;
#pragma unroll
                                            for (int u = 0; u < RC; ++u) {
                                                const float E = erf_fast(sq[u] * fmaf(kf, dt, u0[u]));
                                                acc = fmaf(wt[u] * pref[u], E - Ep[u], acc);
                                                Ep[u] = E;
                                            }
                                            Hw[wrap(k) * NLOS_TILE + lane] += acc;
                                            kf += 1.0f;
                                        }
                                    } else {
                                        for (int k = kmin; k <= kmax; ++k) {
                                            const float w = Hw[wrap(k) * NLOS_TILE + lane];
#pragma unroll
                                            for (int u = 0; u < RC; ++u) {
                                                const float uu = fmaf(kf, dt, u0[u]);
                                                const float z = sq[u] * uu;
                                                const float E = erf_fast(z);
                                                const float Dn = TWO_OVER_SQRT_PI * ex2_approx(-LOG2E_F * z * z);
                                                const float DUn = Dn * uu;
                                                F[u] = fmaf(w, E - Ep[u], F[u]);
                                                P1[u] = fmaf(w, Dn - Dp[u], P1[u]);
                                                P2[u] = fmaf(w, DUn - DUp[u], P2[u]);
                                                Ep[u] = E; Dp[u] = Dn; DUp[u] = DUn;
                                            }
                                            kf += 1.0f;
                                        }
                                    }
                                }
                                if (BWD) {
#pragma unroll
                                    for (int u = 0; u < RC; ++u) {
                                        if (!(kl[u] <= kh[u])) continue;
                                        const int i = ib + (bits[u] >> 3), j = jb + (bits[u] & 7);
                                        const float sth = s_tab[(TAB_STH * A + i) * 32 + lane], cth = s_tab[(TAB_CTH * A + i) * 32 + lane];
                                        const float sph = s_tab[(TAB_SPH * A + j) * 32 + lane], cph = s_tab[(TAB_CPH * A + j) * 32 + lane];
                                        const float dx = sth * cph, dy = sth * sph, dz = cth;
                                        const float bx = M00 * dx + M01 * dy + M02 * dz, by = M10 * dx + M11 * dy + M12 * dz,
                                                    bz = M20 * dx + M21 * dy + M22 * dz;
                                        const float rsu = rs[u], w_ = wt[u];
                                        const float px = fmaf(rsu, bx, ax), py = fmaf(rsu, by, ay), pz = fmaf(rsu, bz, az);
                                        const float C = gam[u] * (-2.0f / LOG2E_F);
                                        gwa = fmaf(sth * pref[u], F[u], gwa);
                                        const float iC = 1.0f / C;
                                        const float gpref = w_ * F[u], gsq = w_ * pref[u] * P2[u], grs = -w_ * pref[u] * sq[u] * P1[u];
                                        const float gm = -0.5f * gpref * pref[u];
                                        const float k1 = grs * iC;                                  // multiplies (perp + r* b) and b
                                        const float k2 = gsq / (2.0f * sq[u]) - gpref * pref[u] * iC;     // multiplies b
                                        // dL/da = 2 gm perp - k1 b ;  dL/db = 2 gm r* perp - k1 (perp + r* b) + k2 b
                                        const float dax = 2.f * gm * px - k1 * bx, day = 2.f * gm * py - k1 * by, daz = 2.f * gm * pz - k1 * bz;
                                        const float dbx = 2.f * gm * rsu * px - k1 * fmaf(rsu, bx, px) + k2 * bx;
                                        const float dby = 2.f * gm * rsu * py - k1 * fmaf(rsu, by, py) + k2 * by;
                                        const float dbz = 2.f * gm * rsu * pz - k1 * fmaf(rsu, bz, pz) + k2 * bz;
                                        gax += dax; gay += day; gaz += daz;
                                        g00 = fmaf(dbx, dx, g00); g01 = fmaf(dbx, dy, g01); g02 = fmaf(dbx, dz, g02);
                                        g10 = fmaf(dby, dx, g10); g11 = fmaf(dby, dy, g11); g12 = fmaf(dby, dz, g12);
                                        g20 = fmaf(dbz, dx, g20); g21 = fmaf(dbz, dy, g21); g22 = fmaf(dbz, dz, g22);
                                    }
                                }
                            }
                        }
                    }
                if (MODE == MODE_COUNT && wb_hi >= wb_lo) c_wb += wb_hi - wb_lo + 1;
            }
            if (BWD && MODE != MODE_COUNT) {
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
    if (!BWD && base > -0x40000000) {
        __syncwarp();
        flush(base, W);
    }
}

static size_t transient_smem_bytes(int A, int window)
{
    size_t f = (size_t)(NLOS_NTAB * A + NLOS_PINFO_ROWS) * NLOS_TILE + (size_t)TR_WPB * window * NLOS_TILE +
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
