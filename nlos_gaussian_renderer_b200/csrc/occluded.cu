// occluded.cu -- batched [scan point, time bin] transient WITH occlusion, and its backward pass.
//
// Reference semantics (per RAY, /root/reference/submodules/cuda_renderer):
//   src/ray_aabb.cu:10-61          per ray the first `cap` (256) Gaussians, in index order, whose 3-sigma AABB passes the slab test
//   src/volume_renderer.cu:80-137  per sample s, in order:  D = sum_g pdf_g sig(o_g) ;  W = sum_g (1 - exp(-pdf_g sig(o_g) c dT)) albedo_g ;
//                                  rho_s = T W ;  T *= exp(-D c dT) ;  stop (zero-fill) once T < 1e-4
//   gaussian_model/cuda_autograd.py:302-314, nlos_helpers.py:275-276   H[p, s] = vp_y^2 dtheta dphi sum_rays rho_s sin(theta) / (r_s^2 + 1e-8)
// The reference loops ONE scan point per optimiser step (main.py:198-203) and materialises three [N_rays, T] tensors per call; its backward
// is zeros.  Here a whole slab of scan points is rendered by three kernels and nothing of size [rays, T] ever exists in memory:
//
//   occ_lists_kernel     CTA = scan point.  Every thread takes one Gaussian of a 256-wide index chunk, finds the few lattice rays its AABB can
//                        meet (footprint rectangle from the direction of the AABB centre, as transient_narrow.cu), runs the slab test
//                        (cuda_utils.cuh:97-121; separable form + the reference's own expression for borderline candidates = bit-exact
//                        decisions) and sets bit `thread` in the ray's 256-bit chunk bitmap in shared memory; the bitmaps are then walked in
//                        bit order, so every ray's list comes out in ASCENDING Gaussian index and is cut at `cap` exactly like the reference's
//                        (O(N) per scan point instead of the reference's O(N_rays N)); the scan stops as soon as every ray is full.
//   occ_forward_kernel   warp = ray.  Listed Gaussians are set up 32 at a time (one per lane), each is evaluated on the samples inside its
//                        5.3-sigma cut only, density and weighted sum accumulate in the warp's shared-memory rows; transmittance by a warp
//                        prefix scan; sin(theta) T W is added to the warp's histogram row and the CTA flushes hscale / (r^2+1e-8) x its rows
//                        to H[p, :] once.
//   occ_backward_kernel  recomputes the ray (never stored), folds dL/dH into the per-sample weights dL/dW and dL/dD (suffix scan for the effect
//                        of a sample's density on every later transmittance), re-evaluates every listed Gaussian against them (moments of
//                        (t - r*) reduced with warp shuffles) and flushes the chain rule of one (ray, Gaussian) pair with 14 REDs.
#include "common.cuh"
#include "rayquad.cuh"
#include "../../include/nlosb200.h"

using namespace nlos;

#define OC_THREADS 256
#define OC_WARPS (OC_THREADS / 32)
#define OC_WORDS (OC_THREADS / 32)      // bitmap words per ray and chunk

struct ScanPoint {
    float ox, oy, oz, th0, idth, ph0, idph, hscale, kband;
    bool valid, fast;
};

// this scan point's lattice columns -> shared memory rows [7][A]: sin th, cos th, tan th | sin ph, cos ph, 1/sin ph, 1/cos ph
__device__ __forceinline__ ScanPoint load_scan_point(const float* __restrict__ tiles, int A, int P, float* __restrict__ tab)
{
    const int tile = P / NLOS_TILE, lane = P % NLOS_TILE;
    const size_t tile_floats = (size_t)(NLOS_NTAB * A + NLOS_PINFO_ROWS) * NLOS_TILE;
    const float* tb = tiles + tile * tile_floats;
    const float* pi = tb + (size_t)NLOS_NTAB * A * NLOS_TILE;
    for (int i = threadIdx.x; i < A; i += blockDim.x) {
        tab[0 * A + i] = tb[TAB_TH(A, i) + TH_SIN + lane];
        tab[1 * A + i] = tb[TAB_TH(A, i) + TH_COS + lane];
        tab[2 * A + i] = tb[TAB_TH(A, i) + TH_TAN + lane];
        tab[3 * A + i] = tb[TAB_PH(A, i) + PH_SIN + lane];
        tab[4 * A + i] = tb[TAB_PH(A, i) + PH_COS + lane];
        tab[5 * A + i] = tb[TAB_PH(A, i) + PH_ISIN + lane];
        tab[6 * A + i] = tb[TAB_PH(A, i) + PH_ICOS + lane];
    }
    ScanPoint S;
    S.ox = pi[PI_OX * NLOS_TILE + lane]; S.oy = pi[PI_OY * NLOS_TILE + lane]; S.oz = pi[PI_OZ * NLOS_TILE + lane];
    S.th0 = pi[PI_TH0 * NLOS_TILE + lane]; S.idth = pi[PI_IDTH * NLOS_TILE + lane];
    S.ph0 = pi[PI_PH0 * NLOS_TILE + lane]; S.idph = pi[PI_IDPH * NLOS_TILE + lane];
    S.hscale = pi[PI_HSCALE * NLOS_TILE + lane];
    S.kband = 1.5e-8f * pi[PI_ISTMAX * NLOS_TILE + lane];
    const int flags = __float_as_int(pi[PI_FLAGS * NLOS_TILE + lane]);
    S.valid = flags & 1; S.fast = flags & 2;
    return S;
}

__device__ __forceinline__ int oc_clamp_ceil(float x, int lo, int hi) { x = ceilf(x); return x <= (float)lo ? lo : (x >= (float)hi ? hi : (int)x); }
__device__ __forceinline__ int oc_clamp_floor(float x, int lo, int hi) { x = floorf(x); return x <= (float)lo ? lo : (x >= (float)hi ? hi : (int)x); }

// ------------------------------------------------------------------------------------------------------------------------------------
// lists [n_sp][A*A][cap+1]: column 0 = count, then ascending Gaussian ids (unused slots are NOT written)
__global__ void __launch_bounds__(OC_THREADS) occ_lists_kernel(const float* __restrict__ tiles, int A, const float* __restrict__ records, int N, int cap,
                                                               int sp0, int n_sp, int* __restrict__ lists)
{
    extern __shared__ __align__(16) unsigned char oc_smem[];
    float* tab = reinterpret_cast<float*>(oc_smem);                                  // [7][A]
    int* cnt = reinterpret_cast<int*>(tab + 7 * A);                                  // [A*A]
    unsigned* bm = reinterpret_cast<unsigned*>(cnt + A * A);                         // [A*A][OC_WORDS]
    const int tid = threadIdx.x, R = A * A;
    for (int spi = blockIdx.x; spi < n_sp; spi += gridDim.x) {
        __syncthreads();
        const ScanPoint S = load_scan_point(tiles, A, sp0 + spi, tab);
        for (int i = tid; i < R; i += OC_THREADS) cnt[i] = 0;
        for (int i = tid; i < R * OC_WORDS; i += OC_THREADS) bm[i] = 0u;
        __syncthreads();
        int* L = lists + (size_t)spi * R * (cap + 1);
        if (S.valid) {
            for (int base = 0; base < N; base += OC_THREADS) {
                const int g = base + tid;
                if (g < N) {
                    const float4* rec = reinterpret_cast<const float4*>(records + (size_t)g * NLOS_REC_FLOATS);
                    const float4 v0 = rec[0], v3 = rec[3], v4 = rec[4], v6 = rec[6];
                    const float cx = v0.x - S.ox, cy = v0.y - S.oy, cz = v0.z - S.oz;
                    const float lox = v3.y - S.ox, loy = v3.z - S.oy, loz = v3.w - S.oz, hix = v4.x - S.ox, hiy = v4.y - S.oy, hiz = v4.z - S.oz;
                    // footprint rectangle (derivation: transient_narrow.cu)
                    int i0 = 0, i1 = A - 1, j0 = 0, j1 = A - 1;
                    const float rho2 = fmaf(cx, cx, cy * cy), d2 = fmaf(cz, cz, rho2), Rb = v6.w;
                    const float rho = sqrtf(rho2), rmin = rho - Rb;
                    if (S.fast && rmin > 0.5f * rho) {
                        const float mg = Rb * Rb / (rmin * rmin), acx = fabsf(cx), acy = fabsf(cy);
                        const float thc = atan2f(rho, cz);
                        const float eth = (fabsf(cz) / rho * (acx * v6.x + acy * v6.y) + rho * v6.z) / d2 * 1.0001f + mg;
                        i0 = oc_clamp_ceil((thc - eth - S.th0) * S.idth - 0.02f, 0, A);
                        i1 = oc_clamp_floor((thc + eth - S.th0) * S.idth + 0.02f, -1, A - 1);
                        if (loy > 0.f) {
                            const float phc = atan2f(cy, cx);
                            const float eph = (acy * v6.x + acx * v6.y) / rho2 * 1.0001f + mg;
                            j0 = oc_clamp_ceil((phc - eph - S.ph0) * S.idph - 0.02f, 0, A);
                            j1 = oc_clamp_floor((phc + eph - S.ph0) * S.idph + 0.02f, -1, A - 1);
                        }
                    }
                    const float box[6] = {v3.y, v3.z, v3.w, v4.x, v4.y, v4.z};
                    for (int j = j0; j <= j1; ++j) {
                        const float icp = tab[6 * A + j], isp = tab[5 * A + j];
                        const float tx0 = lox * icp, tx1 = hix * icp, ty0 = loy * isp, ty1 = hiy * isp;
                        const float xymn = fmaxf(fminf(tx0, tx1), fminf(ty0, ty1)), xymx = fminf(fmaxf(tx0, tx1), fmaxf(ty0, ty1));
                        const float kcol = fmaxf(fabsf(icp), fabsf(isp));
                        for (int i = i0; i <= i1; ++i) {
                            const float tth = tab[2 * A + i];
                            const float tz0 = loz * tth, tz1 = hiz * tth;
                            const float tmin = fmaxf(xymn, fminf(tz0, tz1)), tmax = fminf(xymx, fmaxf(tz0, tz1));
                            const float diff = tmax - tmin;
                            bool hit = diff >= 0.f && tmax >= 0.f;
                            const float kblk = fmaf(fmaxf(kcol, fabsf(tth)), S.kband, 1e-6f);
                            if (kblk > 0.05f || fabsf(diff) <= 2.5f * kblk * fabsf(tmax))       // borderline: the reference's own expression decides
                                hit = ray_aabb_x(S.ox, S.oy, S.oz, __fmul_rn(tab[i], tab[4 * A + j]), __fmul_rn(tab[i], tab[3 * A + j]), tab[A + i], box);
                            if (hit) atomicOr(&bm[(i * A + j) * OC_WORDS + (tid >> 5)], 1u << (tid & 31));
                        }
                    }
                }
                __syncthreads();
                bool allfull = true;
                for (int ray = tid; ray < R; ray += OC_THREADS) {
                    int c = cnt[ray];
#pragma unroll
                    for (int w = 0; w < OC_WORDS; ++w) {
                        unsigned m = bm[ray * OC_WORDS + w];
                        if (m) {
                            bm[ray * OC_WORDS + w] = 0u;
                            while (m && c < cap) {
                                const int b = __ffs(m) - 1;
                                m &= m - 1;
                                L[(size_t)ray * (cap + 1) + 1 + c] = base + w * 32 + b;
                                ++c;
                            }
                        }
                    }
                    cnt[ray] = c;
                    allfull = allfull && c >= cap;
                }
                if (__syncthreads_and(allfull)) break;      // every ray holds `cap` Gaussians: later indices cannot enter (ray_aabb.cu:39)
            }
        }
        __syncthreads();
        for (int ray = tid; ray < R; ray += OC_THREADS) L[(size_t)ray * (cap + 1)] = cnt[ray];
    }
}

// ------------------------------------------------------------------------------------------------------------------------------------
struct OccParams {
    const float* tiles;
    int A;
    const float* records;
    int N, cap;
    const int* lists;
    int sp0, n_sp, ray_blocks;
    const float* t_samples;
    int T;
    float t0, inv_dt, cdt, mcut2;
    int occ;             // 1: shared transmittance (volume_renderer.cu:80-137); 0: the no-occlusion branch (:139-182) on the same capped lists
    const float* sh;
    int sh_dim, sh_degree;
    float* H;            // [n_tiles*32, T] forward out (REDs)
    const float* gH;     // backward in
    float* dacc;         // [N,16] backward out (REDs)
    float* dsh;          // [N,16]
};

// one lane's Gaussian of the current batch: ray quadratic + albedo of this scan point + sample range inside the cut
struct LaneG {
    RayQuad Q;
    int ka, kb, g;
    float a3[3], b3[3], p3[3];
};

template <bool BWD>
__device__ __forceinline__ void lane_setup(const OccParams& P, const ScanPoint& S, const int* __restrict__ row, int idx, int n, float dx, float dy,
                                           float dz, LaneG& G)
{
    G.Q = RayQuad{0.f, 0.f, 0.f, 0.f, 0.f};
    G.ka = 1; G.kb = 0; G.g = -1;
    if (idx >= n) return;
    const int g = row[1 + idx];
    if (g < 0 || g >= P.N) return;
    const float* rec = P.records + (size_t)g * NLOS_REC_FLOATS;
    if (BWD) ray_quad(rec, S.ox, S.oy, S.oz, dx, dy, dz, G.Q, G.a3, G.b3, G.p3);
    else ray_quad(rec, S.ox, S.oy, S.oz, dx, dy, dz, G.Q);
    G.g = g;
    if (P.sh_degree > 0) {
        const float cx = rec[0] - S.ox, cy = rec[1] - S.oy, cz = rec[2] - S.oz;
        G.Q.alb = sh_pair_albedo(P.sh + (size_t)g * P.sh_dim, P.sh_degree, cx, cy, cz, sqrtf(cx * cx + cy * cy + cz * cz));
    }
    // samples with m2 <= mcut^2: |t - rs| <= sqrt((mcut^2 - m2min) / C)
    const float C = G.Q.gam * (-2.0f / LOG2E_F), m2 = G.Q.al * (-2.0f / LOG2E_F);
    const float h2 = (P.mcut2 - m2) / C;
    if (!(h2 > 0.f) || G.Q.op == 0.f) { G.g = BWD ? g : -1; return; }
    const float hw = sqrtf(h2);
    G.ka = oc_clamp_ceil((G.Q.rs - hw - P.t0) * P.inv_dt, 0, P.T);
    G.kb = oc_clamp_floor((G.Q.rs + hw - P.t0) * P.inv_dt, -1, P.T - 1);
}

// density / weighted-sum rows of one ray (rows zero on entry); returns the touched sample range
__device__ __forceinline__ void ray_accumulate(const OccParams& P, const ScanPoint& S, const int* __restrict__ row, int n, float dx, float dy, float dz,
                                               const float* __restrict__ s_t, float* __restrict__ s_den, float* __restrict__ s_ws, int lane, int& klo,
                                               int& khi)
{
    klo = P.T; khi = -1;
    for (int base = 0; base < n; base += 32) {
        LaneG G;
        lane_setup<false>(P, S, row, base + lane, n, dx, dy, dz, G);
        const int m = min(32, n - base);
        for (int q = 0; q < m; ++q) {
            const int ka = __shfl_sync(0xffffffffu, G.ka, q), kb = __shfl_sync(0xffffffffu, G.kb, q);
            if (ka > kb) continue;
            const float rs = __shfl_sync(0xffffffffu, G.Q.rs, q), gam = __shfl_sync(0xffffffffu, G.Q.gam, q);
            const float al = __shfl_sync(0xffffffffu, G.Q.al, q), op = __shfl_sync(0xffffffffu, G.Q.op, q);
            const float alb = __shfl_sync(0xffffffffu, G.Q.alb, q);
            klo = min(klo, ka); khi = max(khi, kb);
            for (int s = ka + lane; s <= kb; s += 32) {
                const float t = s_t[s] - rs;
                const float v = op * ex2_approx(fmaf(gam, t * t, al));                 // pdf * sigmoid(opacity)
                s_den[s] += v;
                s_ws[s] += (P.occ ? one_minus_exp_neg(v * P.cdt) : v * P.cdt) * alb;    // volume_renderer.cu:114-117 / :174-182
            }
            __syncwarp();
        }
    }
}

// transmittance at the samples [klo, khi] of one ray: T_s = prod_{s' < s} exp(-D_s' c dT) (T = 1 before the first touched sample), written over
// s_den's partner row s_tr; returns `alive` = number of samples the reference outputs (the sample whose update drops T below 1e-4 is the last)
__device__ __forceinline__ int ray_transmittance(const OccParams& P, const float* __restrict__ s_den, float* __restrict__ s_tr, int klo, int khi, int lane)
{
    float carry = 0.f;                      // sum of D c dT over the samples before this chunk
    int alive = P.T;
    if (!P.occ) {
        for (int s = klo + lane; s <= khi; s += 32) s_tr[s] = 1.0f;
        __syncwarp();
        return alive;
    }
    for (int c0 = klo; c0 <= khi; c0 += 32) {
        const int s = c0 + lane;
        const float x = s <= khi ? s_den[s] * P.cdt : 0.f;
        float inc = x;
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
            const float y = __shfl_up_sync(0xffffffffu, inc, off);
            if (lane >= off) inc += y;
        }
        const float before = carry + inc - x;
        if (s <= khi) s_tr[s] = ex2_approx(-LOG2E_F * before);
        const bool dead = s <= khi && ex2_approx(-LOG2E_F * (before + x)) < 1e-4f;          // volume_renderer.cu:128-136
        const unsigned mk = __ballot_sync(0xffffffffu, dead);
        if (mk) { alive = c0 + __ffs(mk); break; }
        carry += __shfl_sync(0xffffffffu, inc, 31);
    }
    __syncwarp();
    return alive;
}

template <bool BWD>
__global__ void __launch_bounds__(OC_THREADS) occ_render_kernel(const OccParams P)
{
    extern __shared__ __align__(16) unsigned char oc_smem[];
    const int A = P.A, T = P.T, R = A * A;
    float* tab = reinterpret_cast<float*>(oc_smem);             // [7][A]
    float* s_t = tab + 7 * A;                                   // [T] sample radii
    float* s_w = s_t + T;                                       // [T] 1 / (r^2 + 1e-8)
    float* rows = s_w + T;                                      // [OC_WARPS][3][T]
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float* s_den = rows + (size_t)warp * 3 * T;
    float* s_ws = s_den + T;
    float* s_x = s_ws + T;                                      // forward: the warp's histogram row; backward: transmittance -> dL/dW weights
    for (int s = threadIdx.x; s < T; s += OC_THREADS) {
        const float r = P.t_samples[s];
        s_t[s] = r;
        s_w[s] = 1.0f / (r * r + 1e-8f);                        // cuda_autograd.py:308
    }
    for (int it = blockIdx.x; it < P.n_sp * P.ray_blocks; it += gridDim.x) {
        const int spi = it / P.ray_blocks, rb = it % P.ray_blocks;
        const int Pg = P.sp0 + spi;
        __syncthreads();
        const ScanPoint S = load_scan_point(P.tiles, A, Pg, tab);
        for (int s = lane; s < 3 * T; s += 32) s_den[s] = 0.f;
        __syncthreads();
        if (!S.valid) continue;
        const int rpb = (R + P.ray_blocks - 1) / P.ray_blocks;
        const int r_end = min(R, (rb + 1) * rpb);
        const float* gHrow = BWD ? P.gH + (size_t)Pg * T : nullptr;
        for (int ray = rb * rpb + warp; ray < r_end; ray += OC_WARPS) {
            const int i = ray / A, j = ray % A;
            const float sth = tab[i], dx = __fmul_rn(sth, tab[4 * A + j]), dy = __fmul_rn(sth, tab[3 * A + j]), dz = tab[A + i];
            const int* row = P.lists + ((size_t)spi * R + ray) * (P.cap + 1);
            const int n = min(row[0], P.cap);
            if (n == 0) continue;
            int klo, khi;
            ray_accumulate(P, S, row, n, dx, dy, dz, s_t, s_den, s_ws, lane, klo, khi);
            if (khi < klo) continue;
            if (!BWD) {
                // transmittance, then  acc[s] += sin(theta) T_s W_s  for the samples the reference outputs
                float carry = 0.f;
                for (int c0 = klo; c0 <= khi; c0 += 32) {
                    const int s = c0 + lane;
                    const float x = s <= khi ? s_den[s] * P.cdt : 0.f;
                    float inc = x;
#pragma unroll
                    for (int off = 1; off < 32; off <<= 1) {
                        const float y = __shfl_up_sync(0xffffffffu, inc, off);
                        if (lane >= off) inc += y;
                    }
                    const float before = P.occ ? carry + inc - x : 0.f;
                    const bool dead = P.occ && s <= khi && ex2_approx(-LOG2E_F * (before + x)) < 1e-4f;
                    const unsigned mk = __ballot_sync(0xffffffffu, dead);
                    const int last = mk ? __ffs(mk) - 1 : 31;                         // lanes up to `last` are output
                    if (s <= khi && lane <= last) s_x[s] = fmaf(sth * ex2_approx(-LOG2E_F * before), s_ws[s], s_x[s]);
                    if (mk) break;
                    carry += __shfl_sync(0xffffffffu, inc, 31);
                }
                __syncwarp();
                for (int s = klo + lane; s <= khi; s += 32) { s_den[s] = 0.f; s_ws[s] = 0.f; }
                __syncwarp();
                continue;
            }
            // ---------------- backward of this ray
            const int alive = ray_transmittance(P, s_den, s_x, klo, khi, lane);
            // per-sample weights: gW = dL/dW_s = g_s T_s ; x_s = g_s rho_s (= dL/dT_s T_s) ; gD_s = -c dT sum_{s' > s} x_s'   (alive samples only)
            const float wray = sth * S.hscale;
            for (int s = klo + lane; s <= khi; s += 32) {
                const bool a = s < alive;
                const float gr = a ? gHrow[s] * s_w[s] * wray : 0.f;
                const float tr = s_x[s];
                s_x[s] = a ? gr * tr : 0.f;                     // gW
                s_ws[s] = a ? gr * tr * s_ws[s] : 0.f;          // x
            }
            __syncwarp();
            {
                float carry = 0.f;
                for (int c1 = khi; c1 >= klo; c1 -= 32) {       // suffix scan, chunks from the far end
                    const int s = c1 - lane;
                    const float own = s >= klo ? s_ws[s] : 0.f;
                    float x = own;
#pragma unroll
                    for (int off = 1; off < 32; off <<= 1) {
                        const float y = __shfl_up_sync(0xffffffffu, x, off);
                        if (lane >= off) x += y;
                    }
                    if (s >= klo) s_den[s] = (P.occ && s < alive) ? -P.cdt * (x - own + carry) : 0.f;        // gD
                    carry += __shfl_sync(0xffffffffu, x, 31);
                }
            }
            __syncwarp();
            for (int base = 0; base < n; base += 32) {
                LaneG G;
                lane_setup<true>(P, S, row, base + lane, n, dx, dy, dz, G);
                const int m = min(32, n - base);
                float m0 = 0.f, m1 = 0.f, m2 = 0.f, ma = 0.f;      // this lane's Gaussian: S0, S1, S2, dL/dalbedo
                for (int q = 0; q < m; ++q) {
                    const int ka = __shfl_sync(0xffffffffu, G.ka, q), kb = min(__shfl_sync(0xffffffffu, G.kb, q), alive - 1);
                    if (ka > kb) continue;
                    const float rs = __shfl_sync(0xffffffffu, G.Q.rs, q), gam = __shfl_sync(0xffffffffu, G.Q.gam, q);
                    const float al = __shfl_sync(0xffffffffu, G.Q.al, q), op = __shfl_sync(0xffffffffu, G.Q.op, q);
                    const float alb = __shfl_sync(0xffffffffu, G.Q.alb, q);
                    float S0 = 0.f, S1 = 0.f, S2 = 0.f, Sa = 0.f;
                    for (int s = ka + lane; s <= kb; s += 32) {
                        const float t = s_t[s] - rs;
                        const float e = ex2_approx(fmaf(gam, t * t, al));
                        const float v = op * e;
                        const float ex = P.occ ? ex2_approx(-LOG2E_F * v * P.cdt) : 1.0f;
                        const float dv = fmaf(s_x[s] * alb * P.cdt, ex, s_den[s]);
                        Sa = fmaf(s_x[s], P.occ ? one_minus_exp_neg(v * P.cdt) : v * P.cdt, Sa);
                        const float we = dv * e, wet = we * t;
                        S0 += we;
                        S1 += wet;
                        S2 = fmaf(wet, t, S2);
                    }
                    S0 = warp_sum(S0); S1 = warp_sum(S1); S2 = warp_sum(S2); Sa = warp_sum(Sa);
                    if (lane == q) { m0 = S0; m1 = S1; m2 = S2; ma = Sa; }
                }
                if (G.g >= 0 && (m0 != 0.f || m1 != 0.f || m2 != 0.f || ma != 0.f)) {
                    const float4* r4 = reinterpret_cast<const float4*>(P.records + (size_t)G.g * NLOS_REC_FLOATS);
                    const float4 v0 = r4[0], v1 = r4[1], v2 = r4[2], v3 = r4[3];
                    const float op = G.Q.op, rs = G.Q.rs;
                    const float ux = m0 * G.p3[0] + m1 * G.b3[0], uy = m0 * G.p3[1] + m1 * G.b3[1], uz = m0 * G.p3[2] + m1 * G.b3[2];
                    const float gax = -op * ux, gay = -op * uy, gaz = -op * uz;                                     // dL/da
                    const float gbx = -op * fmaf(rs, ux, m1 * G.p3[0] + m2 * G.b3[0]), gby = -op * fmaf(rs, uy, m1 * G.p3[1] + m2 * G.b3[1]),
                                gbz = -op * fmaf(rs, uz, m1 * G.p3[2] + m2 * G.b3[2]);                              // dL/db
                    const float cx = S.ox - v0.x, cy = S.oy - v0.y, cz = S.oz - v0.z;                               // a = M (o - mu), b = M d
                    float o[14];
                    o[0] = fmaf(gax, cx, gbx * dx); o[1] = fmaf(gax, cy, gbx * dy); o[2] = fmaf(gax, cz, gbx * dz);
                    o[3] = fmaf(gay, cx, gby * dx); o[4] = fmaf(gay, cy, gby * dy); o[5] = fmaf(gay, cz, gby * dz);
                    o[6] = fmaf(gaz, cx, gbz * dx); o[7] = fmaf(gaz, cy, gbz * dy); o[8] = fmaf(gaz, cz, gbz * dz);
                    o[9] = -(v1.x * gax + v1.w * gay + v2.z * gaz);                                                 // dL/dmu = -M^T dL/da
                    o[10] = -(v1.y * gax + v2.x * gay + v2.w * gaz);
                    o[11] = -(v1.z * gax + v2.y * gay + v3.x * gaz);
                    o[12] = m0;                                                                                     // dL/dop
                    o[13] = ma;                                                                                     // dL/dalbedo (degree 0)
                    if (P.sh_degree > 0) {
                        float dshv[16];
                        const float dpre = G.Q.alb > 0.f ? ma : 0.f;
                        o[13] = 0.f;
                        sh_pair_backward(P.sh + (size_t)G.g * P.sh_dim, P.sh_degree, -cx, -cy, -cz, sqrtf(cx * cx + cy * cy + cz * cz), dpre, dshv, o[9], o[10],
                                         o[11]);
                        for (int k = 0; k < 16; ++k)
                            if (dshv[k] != 0.f) atomicAdd(P.dsh + (size_t)G.g * 16 + k, dshv[k]);
                    }
                    float* dst = P.dacc + (size_t)G.g * NLOS_GRAD_FLOATS;
#pragma unroll
                    for (int k = 0; k < 14; ++k)
                        if (o[k] != 0.f) atomicAdd(dst + k, o[k]);
                }
            }
            __syncwarp();
            for (int s = klo + lane; s <= khi; s += 32) { s_den[s] = 0.f; s_ws[s] = 0.f; s_x[s] = 0.f; }
            __syncwarp();
        }
        if (!BWD) {
            // flush the CTA's histogram rows:  H[p, s] += hscale / (r_s^2 + 1e-8) * sum_warps acc[s]
            __syncthreads();
            for (int s = threadIdx.x; s < T; s += OC_THREADS) {
                float a = 0.f;
#pragma unroll
                for (int w = 0; w < OC_WARPS; ++w) a += rows[(size_t)w * 3 * T + 2 * T + s];
                if (a != 0.f) atomicAdd(P.H + (size_t)Pg * T + s, a * S.hscale * s_w[s]);
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------------------------------------------
extern "C" long long nlosb200_occluded_list_ints(int n_scan_points, int A, int cap) { return (long long)n_scan_points * A * A * (cap + 1); }

static int occ_check(const char* who, int A, int T, int N, int cap, int sp0, int n_sp, int n_tiles)
{
    NLOS_REQUIRE(A >= 1 && A <= 64, "%s: the batched occlusion renderer holds one 256-bit bitmap per lattice ray in shared memory: A=%d > 64", who, A);
    NLOS_REQUIRE(T >= 1 && T <= 2048 && N >= 1 && cap >= 1, "%s: bad sizes (T=%d in [1,2048], N=%d, cap=%d)", who, T, N, cap);
    NLOS_REQUIRE(sp0 >= 0 && n_sp >= 1 && sp0 + n_sp <= n_tiles * NLOS_TILE, "%s: scan points [%d, %d) outside the %d tiles", who, sp0, sp0 + n_sp, n_tiles);
    return 0;
}

extern "C" int nlosb200_occluded_lists(const float* tiles, int n_tiles, int A, const float* records, int N, int cap, int sp0, int n_sp, int* lists,
                                       void* stream)
{
    int rc = occ_check("occluded_lists", A, 1, N, cap, sp0, n_sp, n_tiles);
    if (rc) return rc;
    NLOS_REQUIRE(tiles && records && lists, "occluded_lists: null pointer");
    const size_t smem = (size_t)7 * A * 4 + (size_t)A * A * 4 + (size_t)A * A * OC_WORDS * 4;
    cudaError_t e = cudaFuncSetAttribute(occ_lists_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    NLOS_REQUIRE(e == cudaSuccess, "occluded_lists: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    occ_lists_kernel<<<n_sp, OC_THREADS, smem, (cudaStream_t)stream>>>(tiles, A, records, N, cap, sp0, n_sp, lists);
    NLOS_CHECK_LAUNCH("occ_lists_kernel");
    return 0;
}

static int occ_fill(OccParams& P, const float* tiles, int n_tiles, int A, const float* records, int N, int cap, const int* lists, int sp0, int n_sp,
                    const float* t_samples, int T, float t0, float dt, float c_deltaT, float mcut, int use_occlusion, const float* sh, int sh_dim, int sh_degree,
                    size_t& smem, int& grid)
{
    int rc = occ_check("occluded_render", A, T, N, cap, sp0, n_sp, n_tiles);
    if (rc) return rc;
    NLOS_REQUIRE(tiles && records && lists && t_samples, "occluded_render: null pointer");
    NLOS_REQUIRE(sh_degree >= 0 && sh_degree <= 3 && (sh_degree == 0 || (sh && sh_dim >= (sh_degree + 1) * (sh_degree + 1))),
                 "occluded_render: degree %d with %d SH coefficients", sh_degree, sh_dim);
    P.tiles = tiles; P.A = A; P.records = records; P.N = N; P.cap = cap; P.lists = lists; P.sp0 = sp0; P.n_sp = n_sp;
    P.t_samples = t_samples; P.T = T; P.t0 = t0; P.inv_dt = 1.0f / dt; P.cdt = c_deltaT; P.mcut2 = mcut * mcut; P.occ = use_occlusion ? 1 : 0;
    P.sh = sh; P.sh_dim = sh_dim; P.sh_degree = sh_degree;
    P.H = nullptr; P.gH = nullptr; P.dacc = nullptr; P.dsh = nullptr;
    // enough CTAs for ~4 waves of the SMs: split the rays of a scan point into blocks when there are few scan points
    int n_sm = 148, dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev);
    int rbk = (4 * n_sm + n_sp - 1) / n_sp;
    const int max_rb = (A * A + OC_WARPS - 1) / OC_WARPS;
    P.ray_blocks = rbk < 1 ? 1 : (rbk > max_rb ? max_rb : rbk);
    grid = n_sp * P.ray_blocks;
    smem = ((size_t)7 * A + 2 * (size_t)T + (size_t)OC_WARPS * 3 * T) * 4;
    NLOS_REQUIRE(smem <= 227 * 1024, "occluded_render: %zu B of shared memory (A=%d, T=%d) > 227 KB", smem, A, T);
    return 0;
}

extern "C" int nlosb200_occluded_forward(const float* tiles, int n_tiles, int A, const float* records, int N, int cap, const int* lists, int sp0,
                                         int n_sp, const float* t_samples, int T, float t0, float dt, float c_deltaT, float mcut, int use_occlusion,
                                         const float* sh, int sh_dim, int sh_degree, float* H, void* stream)
{
    OccParams P;
    size_t smem;
    int grid;
    int rc = occ_fill(P, tiles, n_tiles, A, records, N, cap, lists, sp0, n_sp, t_samples, T, t0, dt, c_deltaT, mcut, use_occlusion, sh, sh_dim, sh_degree, smem, grid);
    if (rc) return rc;
    NLOS_REQUIRE(H, "occluded_forward: null H");
    P.H = H;
    cudaError_t e = cudaFuncSetAttribute(occ_render_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    NLOS_REQUIRE(e == cudaSuccess, "occluded_forward: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    occ_render_kernel<false><<<grid, OC_THREADS, smem, (cudaStream_t)stream>>>(P);
    NLOS_CHECK_LAUNCH("occ_render_kernel<fwd>");
    return 0;
}

extern "C" int nlosb200_occluded_backward(const float* tiles, int n_tiles, int A, const float* records, int N, int cap, const int* lists, int sp0,
                                          int n_sp, const float* t_samples, int T, float t0, float dt, float c_deltaT, float mcut, int use_occlusion,
                                          const float* sh, int sh_dim, int sh_degree, const float* grad_H, float* dacc, float* dsh, void* stream)
{
    OccParams P;
    size_t smem;
    int grid;
    int rc = occ_fill(P, tiles, n_tiles, A, records, N, cap, lists, sp0, n_sp, t_samples, T, t0, dt, c_deltaT, mcut, use_occlusion, sh, sh_dim, sh_degree, smem, grid);
    if (rc) return rc;
    NLOS_REQUIRE(grad_H && dacc && (sh_degree == 0 || dsh), "occluded_backward: null pointer");
    P.gH = grad_H; P.dacc = dacc; P.dsh = dsh;
    cudaError_t e = cudaFuncSetAttribute(occ_render_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    NLOS_REQUIRE(e == cudaSuccess, "occluded_backward: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    occ_render_kernel<true><<<grid, OC_THREADS, smem, (cudaStream_t)stream>>>(P);
    NLOS_CHECK_LAUNCH("occ_render_kernel<bwd>");
    return 0;
}
