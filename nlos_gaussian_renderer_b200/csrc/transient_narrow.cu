// transient_narrow.cu -- the scan-tile kernel of the batched transient renderer for SMALL footprints: Gaussians that meet only a few
// lattice rays per scan point (BASELINE configs[2..4] and the north-star scene: 1.3 - 3.5 hit rays per (Gaussian, scan point), against 9 in
// configs[1]).  Same semantics, lists, windows, record ring and persistent work-stealing grid as transient.cu (see there for the
// reference citations: volume_renderer.cu:139-182, cuda_autograd.py:254-314, nlos_helpers.py:275-276); what differs is where the
// instructions of a (Gaussian, scan point) pair go when there are only ~24 sample-evals to amortise them over:
//   * footprint rectangle from the direction of the AABB centre and its first-order angular half-extents (2 atan2 instead of 4, no corner
//     selection), widened by a rigorous bound of the second-order remainder;
//   * the common case "at most 2 x 2 candidate rays in every lane" is decided by ONE warp vote and takes a straight-line slab test
//     (same predicate as cuda_utils.cuh:97-121, bit-exact decisions); larger footprints fall back to the 4 x 8 block test of transient.cu;
//   * forward: no register accumulators (they pay when several rays of a pair share bins): the packed {up, down} recurrence adds straight
//     into the warp's window, the exact number of steps, nothing to zero and nothing to merge;
//   * backward: as transient.cu (prefix-sum moments against the window's weights, one butterfly reduction per entry).
#include "transient_shared.cuh"

template <int MODE, bool BWD, bool SH>
__global__ void __launch_bounds__(TR_WPB * 32, TR_MINB) transient_narrow_kernel(const TransientParams P)
{
    constexpr bool FWD = !BWD;
    constexpr bool GRAD = BWD;
    constexpr float KG = -0.5f * LOG2E_F;
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
    int* s_ctl = reinterpret_cast<int*>(s_bar + 1 + TR_WPB * TR_STAGES);                                      // [2] next tile, tiles consumed
    // staging of the culling words of a ring batch (backward): [TR_WPB][TR_STAGES][TR_RING][32] u16, 16-byte aligned
    unsigned short* s_stash = reinterpret_cast<unsigned short*>((reinterpret_cast<uintptr_t>(s_ctl + 2) + 15) & ~(uintptr_t)15);

    if (P.cell_offsets[(size_t)P.n_tiles * P.cpt] > P.list_capacity) return;   // list buffer overflow (flagged by bin_window_fill): render nothing
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        mbar_init(&s_bar[0], 1);
        for (int i = 0; i < TR_WPB * TR_STAGES; ++i) mbar_init(&s_bar[1 + i], 1);
        mbar_fence_init();
    }
    float* Hw = s_H + warp * wstride + TR_GUARD * NLOS_TILE;            // bin `base` of this warp's window
    float* ring = s_ring + warp * TR_STAGES * TR_RING * NLOS_REC_FLOATS;
    uint64_t* bar = s_bar + 1 + warp * TR_STAGES;
    // ---- culling words: the forward pass leaves, per (entry, scan point), the anchor of the candidate block and the bit-exact hit mask of
    //      its <= 2 x 2 rays in P.hit_stash (i0 | j0 << 6 | mask << 12; 0xFFFF = "entry with a wider rectangle: decide again"); the backward
    //      pass of the same lists reads them instead of repeating the footprint rectangle and the slab tests
    unsigned short* const stash_w = s_stash + warp * TR_STAGES * TR_RING * NLOS_TILE;
    const bool st_rd = BWD && P.hit_stash != nullptr, st_wr = FWD && P.hit_stash != nullptr;
    for (int i = lane; i < wstride; i += 32) Hw[i - TR_GUARD * NLOS_TILE] = 0.f;
    __syncthreads();

    const float t0 = P.t0, dt = P.dt, inv_dt = P.inv_dt, dt2 = P.dt * P.dt;
    const int T = P.T;
    const float* const s_hs = s_pi + PI_HSCALE * 32;
    const float* const tabl = s_tab + lane;                               // this lane's column of the lattice tables
    unsigned rb = 0;                       // ring batches consumed by this warp: stage = rb & 1, parity = (rb >> 1) & 1
    unsigned tphase = 0;
    // ---- persistent CTA: walk the tiles cyclically from an evenly spread start; warps pull slices of the current tile's
    //      entry list from its global counter, so CTAs that run out of work join the tiles that still have some
    const int n_vt = P.n_tiles * P.n_seg;          // virtual tiles (segment, tile): one launch walks up to two entry classes
    int t_cur = (int)(((long long)blockIdx.x * n_vt) / gridDim.x), visited = 0;
    for (;;) {
        if (warp == 0) {
            int found = -1, adv = 0;
            while (visited + adv < n_vt) {
                const int off = adv + lane;
                bool has = false;
                if (visited + off < n_vt) {
                    int t = t_cur + off;
                    if (t >= n_vt) t -= n_vt;
                    const int sg = P.n_seg == 2 ? (t & 1) : 0, tt = P.n_seg == 2 ? (t >> 1) : t;
                    const long long* co = P.cell_offsets + (size_t)tt * P.cpt + P.seg_cell0[sg];
                    const long long tot = co[P.n_buckets] - co[0];
                    has = *(volatile int*)(P.work + t) < (int)((tot + P.slice_len - 1) / P.slice_len);
                }
                const unsigned m = __ballot_sync(0xffffffffu, has);
                if (m) {
                    const int f = __ffs(m) - 1;
                    found = t_cur + adv + f;
                    if (found >= n_vt) found -= n_vt;
                    adv += f + 1;
                    break;
                }
                adv += 32;
            }
            if (lane == 0) { s_ctl[0] = found; s_ctl[1] = adv; }
        }
        __syncthreads();
        const int vt = s_ctl[0];
        if (vt < 0) break;
        const int seg = P.n_seg == 2 ? (vt & 1) : 0, tile = P.n_seg == 2 ? (vt >> 1) : vt;      // segments interleaved: a CTA's share of the ring holds both classes of its tiles
        const int seg_cell0 = P.seg_cell0[seg];
        const bool one_pass = P.seg_one_pass[seg] != 0;
        float* const Hout = P.seg_H[seg];
        const float* const gHin = P.seg_gH[seg];
        visited += s_ctl[1];
        t_cur += s_ctl[1];
        if (t_cur >= n_vt) t_cur -= n_vt;
        if (threadIdx.x == 0) {
            fence_proxy_async();
            mbar_arrive_expect_tx(&s_bar[0], (uint32_t)tile_floats * 4u);
            bulk_g2s(s_tab, P.tiles + (size_t)tile * tile_floats, (uint32_t)tile_floats * 4u, &s_bar[0]);
        }
        const long long tile_beg = P.cell_offsets[(size_t)tile * P.cpt + seg_cell0];
        const long long tile_tot = P.cell_offsets[(size_t)tile * P.cpt + seg_cell0 + P.n_buckets] - tile_beg;
        const int n_slices = (int)((tile_tot + P.slice_len - 1) / P.slice_len);
        mbar_wait(&s_bar[0], tphase);   // lattice tables + scan-point info have landed
        tphase ^= 1u;

        // ---- per-lane scan point
        const float ox = s_pi[PI_OX * 32 + lane], oy = s_pi[PI_OY * 32 + lane], oz = s_pi[PI_OZ * 32 + lane];
        const int flags = __float_as_int(s_pi[PI_FLAGS * 32 + lane]);
        const bool valid = flags & 1, fast = flags & 2;

        for (;;) {
            int sl = 0;
            if (lane == 0) sl = atomicAdd(P.work + vt, 1);
            sl = __shfl_sync(0xffffffffu, sl, 0);
            if (sl >= n_slices) break;
            const long long first = (long long)sl * P.slice_len;
            const int n = (int)min((long long)P.slice_len, tile_tot - first);
            const long long beg = tile_beg + first;

            // ---- linear window [base, base + W): when it advances, the bins that fall behind are flushed (forward) and the
            //      rest is shifted down; the backward pass refills the new bins with the weights dL/dH * hscale / (r^2+1e-8)
            int base = NO_BASE;
            if (FWD) {
                for (int k = 0; k < W; ++k) Hw[k * NLOS_TILE + lane] = 0.f;
                __syncwarp();
            }
            auto flush = [&](int cnt) {                       // forward: H[p,k] += acc * hscale_p / (r_k^2 + 1e-8) * flush_scale  (cuda_autograd.py:308-314)
                const float rcnt = 1.0f / (float)cnt;          // idx / cnt by a multiply: (idx + .5) / cnt stays >= .5 / cnt away from an integer, exact for idx < 2^16
                for (int idx = lane; idx < cnt * NLOS_TILE; idx += 32) {
                    const int pp = (int)(((float)idx + 0.5f) * rcnt), kk = idx - pp * cnt, k = base + kk;
                    const float v = Hw[kk * NLOS_TILE + pp];
                    if (v != 0.f && k < T) {
                        const float rk = fmaf((float)k, dt, t0);
                        atomicAdd(Hout + ((size_t)tile * NLOS_TILE + pp) * T + k, v * (s_hs[pp] * P.flush_scale) / (one_pass ? 1.0f : rk * rk + 1e-8f));
                    }
                }
            };
            auto refill = [&](int nbase, int kk_from, int cnt) {       // backward: W[k][p] = dL/dH[p,k] * hscale_p / (r_k^2 + 1e-8) * flush_scale
                const float rcnt = 1.0f / (float)cnt;          // (as in flush)
                for (int idx = lane; idx < cnt * NLOS_TILE; idx += 32) {
                    const int pp = (int)(((float)idx + 0.5f) * rcnt), kk = kk_from + (idx - pp * cnt), k = nbase + kk;
                    float v = 0.f;
                    if (k >= 0 && k < T) {
                        const float rk = fmaf((float)k, dt, t0);
                        v = gHin[((size_t)tile * NLOS_TILE + pp) * T + k] * (s_hs[pp] * P.flush_scale) / (one_pass ? 1.0f : rk * rk + 1e-8f);
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

            // ---- TMA ring over the list slice
            const int n_batches = (n + TR_RING - 1) / TR_RING;
            unsigned ent_next = 0, ent_cur = 0, bins_next = 0, bins_cur = 0;
            auto issue = [&](int b, unsigned r) {
                const int stage = (int)(r & 1u);
                const int cnt = min(TR_RING, n - b * TR_RING);
                // (no proxy fence: the stage is only ever READ through the generic proxy, and those reads were consumed before the
                //  __syncwarp below; fence.proxy.async is MEMBAR.ALL.CTA + FENCE.VIEW.ASYNC, i.e. it would wait for this warp's
                //  outstanding gradient / histogram REDs every 4 entries -- 15 % of the backward's stall samples)
                if (lane == 0) mbar_arrive_expect_tx(&bar[stage], (uint32_t)cnt * (NLOS_REC_FLOATS * 4u + (st_rd ? NLOS_TILE * 2u : 0u)));
                __syncwarp();
                if (st_rd && lane == 0)      // the batch's entries are consecutive list positions: one copy of cnt x 64 B
                    bulk_g2s(stash_w + stage * TR_RING * NLOS_TILE, P.hit_stash + (size_t)(beg + b * TR_RING) * NLOS_TILE, (uint32_t)cnt * NLOS_TILE * 2u,
                             &bar[stage]);
                ent_next = 0;
                if (lane < cnt) {
                    ent_next = P.list[beg + b * TR_RING + lane];
                    bins_next = P.list_bins[beg + b * TR_RING + lane];
                    bulk_g2s(ring + (stage * TR_RING + lane) * NLOS_REC_FLOATS,
                             P.records + (size_t)(ent_next & ((1u << NLOS_ENTRY_SHIFT) - 1u)) * NLOS_REC_FLOATS, NLOS_REC_FLOATS * 4u, &bar[stage]);
                }
            };
            issue(0, rb);
            for (int b = 0; b < n_batches; ++b, ++rb) {
                ent_cur = ent_next;
                bins_cur = bins_next;
                if (b + 1 < n_batches) issue(b + 1, rb + 1);   // (overwrites ent_next; ent_cur keeps this batch's entries)
                mbar_wait(&bar[rb & 1u], (rb >> 1) & 1u);
                const int cnt = min(TR_RING, n - b * TR_RING);
                for (int q = 0; q < cnt; ++q) {
                    const unsigned ent = __shfl_sync(0xffffffffu, ent_cur, q);
                    const int esub = (int)(ent >> NLOS_ENTRY_SHIFT);
                    const float4* rec = reinterpret_cast<const float4*>(ring + ((rb & 1u) * TR_RING + q) * NLOS_REC_FLOATS);
                    const float4 v0 = rec[0], v1 = rec[1], v2 = rec[2], v3 = rec[3], v4 = rec[4], v5 = rec[5], v6 = rec[6];
                    const float M00 = v1.x, M01 = v1.y, M02 = v1.z, M10 = v1.w, M11 = v2.x, M12 = v2.y, M20 = v2.z, M21 = v2.w, M22 = v3.x;
                    const float cx = v0.x - ox, cy = v0.y - oy, cz = v0.z - oz;                             // mu - o
                    // ---- the tile-level shell [klo, khi] of this Gaussian comes with the list entry (binning.cu:shell_bins)
                    int s_lo, s_hi;
                    bool split, lo_exact, hi_exact, force_slow;
                    const float d2 = fmaf(cx, cx, fmaf(cy, cy, cz * cz));
                    {
                        const unsigned bw = __shfl_sync(0xffffffffu, bins_cur, q);
                        const int klo_t = (int)(bw & 0xffffu), khi_t = (int)(bw >> 16);
                        s_lo = klo_t + esub * L;                                                         // this entry's sub-interval
                        s_hi = min(khi_t, s_lo + L - 1);
                        split = khi_t - klo_t + 1 > L;                                                   // the Gaussian has more than one entry
                        lo_exact = split && esub > 0;                                                    // a sibling entry ends right below s_lo
                        hi_exact = split && s_hi < khi_t;                                                // a sibling entry starts right above s_hi
                        force_slow = split && (L & 3) != 0;                                              // (entry length not a multiple of 4: all direct)
                    }
                    if (s_lo > s_hi) continue;                                                           // (cannot happen for a listed entry)
                    {
                        const int nbase = s_lo & ~(BS - 1);                                              // bucket is a power of two
                        if (nbase != base) advance(nbase);
                    }
                    const float fRlo = (float)s_lo, fRhi = (float)s_hi;
                    const float r_in = fmaf(fRlo - 0.5f, dt, t0), r_out = fmaf(fRhi + 0.5f, dt, t0);    // bin k covers [t0+(k-.5)dt, t0+(k+.5)dt]
                    // accumulators of the backward pass for this pair
                    float gax = 0.f, gay = 0.f, gaz = 0.f;                                                  // dL/da
                    float g00 = 0.f, g01 = 0.f, g02 = 0.f, g10 = 0.f, g11 = 0.f, g12 = 0.f, g20 = 0.f, g21 = 0.f, g22 = 0.f;   // sum dL/db (x) d
                    float gwa = 0.f;                                                                        // dL/d(op*alb)
                    // numerical backward: the same sums with ray 0 / ray 1 of an iteration in the two halves of packed registers
                    f32x2 pga[3] = {0ull, 0ull, 0ull}, pg[9] = {0ull, 0ull, 0ull, 0ull, 0ull, 0ull, 0ull, 0ull, 0ull}, pgwa = 0ull;
                    // ---- the pair (this lane's scan point): radial reject against the entry's bins, AABB relative to the scan point, a = M (o - mu)
                    // D = |mu - o| to ~1 ulp from ONE MUFU.RSQ + a Newton step (no IEEE-sqrt slow path); 1 / d2 for the footprint below is ir^2
                    const float ir = rsqrt_approx(fmaxf(d2, 1e-30f));
                    float D = d2 * ir;
                    D = fmaf(fmaf(-D, D, d2), 0.5f * ir, D);
                    float alb = v4.w;                                                                       // degree-0 albedo of the record
                    if (SH)                                                                                 // view-dependent: one albedo per (Gaussian, scan point)
                        alb = sh_pair_albedo(P.sh + (size_t)(ent & ((1u << NLOS_ENTRY_SHIFT) - 1u)) * P.sh_dim, P.sh_degree, cx, cy, cz, D);
                    const float opalb = v0.w * alb;
                    const bool act = valid && opalb > 0.f && (D + v5.x >= r_in) && (D - v5.x <= r_out);
                    const float lox = v3.y - ox, loy = v3.z - oy, loz = v3.w - oz;
                    const float hix = v4.x - ox, hiy = v4.y - oy, hiz = v4.z - oz;
                    const float ax = -(M00 * cx + M01 * cy + M02 * cz), ay = -(M10 * cx + M11 * cy + M12 * cz),
                                az = -(M20 * cx + M21 * cy + M22 * cz);
                    // ---- conservative footprint rectangle in lattice index space.  Every point of the AABB is mu - o + delta with
                    //      |delta_i| <= e_i (record: half extents, |e| = Rb), so its polar / azimuthal angle lies within
                    //        |grad theta| . e + R2   of theta(mu - o)     (same for phi),
                    //      grad theta = (cz cx / rho, cz cy / rho, -rho) / d2 ,  grad phi = (-cy, cx, 0) / rho^2 ,  rho^2 = cx^2 + cy^2 ,
                    //      and the second-order remainder R2 <= Rb^2 / (2 rho_min^2): the Hessians of both angles have norm <= 1 / rho^2
                    //      (eigenvalues +-1/r^2 and cot(theta)/r^2 for theta, +-1/rho^2 for phi) and rho >= rho_min = rho - Rb over the box.
                    //      Twice that bound is used; boxes with Rb > rho / 2 (never for a relay wall in the plane y = 0) take every ray.
                    int i0 = 0, i1 = -1, j0 = 0, j1 = -1;
                    unsigned pm = 0;            // culling word: the hit mask of the 2 x 2 block anchored at (i0, j0)
                    bool use_st = false;
                    if (st_rd) {
                        const unsigned stw = stash_w[((rb & 1u) * TR_RING + q) * NLOS_TILE + lane];
                        use_st = __all_sync(0xffffffffu, stw != 0xFFFFu);                       // (the marker is written by whole warps)
                        if (use_st) { i0 = (int)(stw & 63u); j0 = (int)((stw >> 6) & 63u); i1 = i0 + 1; j1 = j0 + 1; pm = stw >> 12; }
                    }
                    if (!use_st && act) {
                        i1 = A - 1; j1 = A - 1;
                        const float rho2 = fmaf(cx, cx, cy * cy), Rb = v6.w;
                        const float irho = rsqrt_approx(fmaxf(rho2, 1e-30f)), rho = rho2 * irho, rmin = rho - Rb;
                        if (fast && rmin > 0.5f * rho) {
                            const float th0 = s_pi[PI_TH0 * 32 + lane], idth = s_pi[PI_IDTH * 32 + lane];
                            // Rb^2 / rmin^2 = x^2 / (1 - x)^2, x = Rb / rho < 1/2, bounded by x^2 (1 + 2x + 12x^2): no reciprocal on the chain
                            const float xr = Rb * irho;
                            const float mg = xr * xr * fmaf(xr, fmaf(12.0f, xr, 2.0f), 1.0f);
                            const float acx = fabsf(cx), acy = fabsf(cy);
                            const float thc = atan2_fast_inv(rho, cz, irho);
                            const float eth = fmaf(fmaf(fabsf(cz) * irho, fmaf(acx, v6.x, acy * v6.y), rho * v6.z), ir * ir * 1.0001f, mg);
                            i0 = lattice_lo((thc - eth - th0) * idth - 0.02f, A);
                            i1 = lattice_hi((thc + eth - th0) * idth + 0.02f, A - 1);
                            if (loy > 0.f) {
                                const float ph0 = s_pi[PI_PH0 * 32 + lane], idph = s_pi[PI_IDPH * 32 + lane];
                                const float phc = atan2_fast(cy, cx);
                                const float eph = fmaf(fmaf(acy, v6.x, acx * v6.y), irho * irho * 1.0001f, mg);
                                j0 = lattice_lo((phc - eph - ph0) * idph - 0.02f, A);
                                j1 = lattice_hi((phc + eph - ph0) * idph + 0.02f, A - 1);
                            }
                        }
                    }
                    // ---- the pair's centre bin: up bins kc + s (s < n_up), down bins kc - 1 - s (s < n_dn).  For a Gaussian split
                    //      over several entries kc is aligned to s_lo modulo 4, so that the entry's boundaries fall on the
                    //      4-step blocks of the backward loop (L % 4 == 0: exact clipping at no cost in the hot loop)
                    int kc = rint_small(fminf(fmaxf((D - t0) * inv_dt, fRlo), fRhi + 1.0f));      // (0 <= value <= T: exact rint without F2I)
                    if (split) {
                        kc = s_lo + ((kc - s_lo + 2) & ~3);
                        if (kc > s_hi + 1) kc -= 4;
                        if (kc < s_lo) kc = s_lo;
                    }
                    const int n_up = s_hi - kc + 1, n_dn = kc - s_lo;
                    const int n_lim = max(n_up, n_dn);
                    const float fkc = (float)kc;
                    const float dD = fmaf(fkc, dt, t0) - D;                      // t(kc) - D
                    float* const hpc = Hw + (kc - base) * NLOS_TILE + lane;      // bin kc ; bin kc+s at hpc[32 s], bin kc-1-s at hpc[-32 (s+1)]
                    // backward: step (multiple of 4) from which a direction is switched off -- exact at the boundaries shared with a
                    // sibling entry, otherwise up to 3 bins past the entry (guard bins / bins beyond the Gaussian's shell)
                    const int zu = hi_exact ? n_up : ((n_up + TR_GUARD) & ~3), zd = lo_exact ? n_dn : ((n_dn + TR_GUARD) & ~3);
                    const int zmin = min(zu, zd);

                    // ---- ray -> Gaussian in r, re-centred on the sphere through the mean:  a' = a + D b = M (o + D d - mu)
                    //      m2(r) = |a' + (r - D) b|^2 = m2min + C (r - r*)^2 ,  r* = D + rsp
                    struct Ray {
                        float rsp;     // r* - D
                        float gam;     // -0.5 log2(e) C
                        float al;      // -0.5 log2(e) m2min
                        float sth;     // sin(theta): quadrature weight of the ray
                        float hh;      // half width of the mcut cut in bins (NaN if the ray misses the cut)
                        float tc;      // t(kc) - r*
                        bool ok;       // the ray intersects the cut
                    };
                    auto ray_setup = [&](const float* th, const float* ph, Ray& R) {
                        const float sth = th[TH_SIN], cth = th[TH_COS], sph = ph[PH_SIN], cph = ph[PH_COS];
                        const float dx = sth * cph, dy = sth * sph;
                        const float bx = fmaf(M00, dx, fmaf(M01, dy, M02 * cth)), by = fmaf(M10, dx, fmaf(M11, dy, M12 * cth)),
                                    bz = fmaf(M20, dx, fmaf(M21, dy, M22 * cth));                       // b = M d
                        const float px = fmaf(D, bx, ax), py = fmaf(D, by, ay), pz = fmaf(D, bz, az);   // a'
                        const float C = fmaf(bx, bx, fmaf(by, by, bz * bz));
                        const float sp = fmaf(px, bx, fmaf(py, by, pz * bz));
                        const float n2 = fmaf(px, px, fmaf(py, py, pz * pz));
                        const float iC = rcp_approx(C);
                        R.rsp = -sp * iC;
                        const float m2 = fmaf(R.rsp, sp, n2);                                           // m2min = |a'|^2 - (a'.b)^2 / C
                        const float h2 = (P.mcut2 - m2) * iC;
                        R.ok = h2 > 0.f;
                        R.hh = sqrt_approx(h2) * inv_dt;
                        R.gam = KG * C;
                        R.al = KG * m2;
                        R.sth = sth;
                        R.tc = dD - R.rsp;
                    };
                    // the same for two rays at once, one ray in each half of the packed fp32x2 operations (FFMA2/FMUL2)
                    auto ray_setup2 = [&](const float* tha, const float* pha, const float* thb_, const float* phb_, Ray& Ra, Ray& Rb) {
                        const f32x2 sth = pack2(tha[TH_SIN], thb_[TH_SIN]), cth = pack2(tha[TH_COS], thb_[TH_COS]);
                        const f32x2 sph = pack2(pha[PH_SIN], phb_[PH_SIN]), cph = pack2(pha[PH_COS], phb_[PH_COS]);
                        const f32x2 dx = mul2(sth, cph), dy = mul2(sth, sph);
                        const f32x2 bx = fma2(pack2(M00, M00), dx, fma2(pack2(M01, M01), dy, mul2(pack2(M02, M02), cth)));
                        const f32x2 by = fma2(pack2(M10, M10), dx, fma2(pack2(M11, M11), dy, mul2(pack2(M12, M12), cth)));
                        const f32x2 bz = fma2(pack2(M20, M20), dx, fma2(pack2(M21, M21), dy, mul2(pack2(M22, M22), cth)));
                        const f32x2 Dv = pack2(D, D);
                        const f32x2 px = fma2(Dv, bx, pack2(ax, ax)), py = fma2(Dv, by, pack2(ay, ay)), pz = fma2(Dv, bz, pack2(az, az));
                        const f32x2 C = fma2(bx, bx, fma2(by, by, mul2(bz, bz)));
                        const f32x2 sp = fma2(px, bx, fma2(py, by, mul2(pz, bz)));
                        const f32x2 n2 = fma2(px, px, fma2(py, py, mul2(pz, pz)));
                        float Ca, Cb;
                        unpack2(C, Ca, Cb);
                        const f32x2 iC = pack2(rcp_approx(Ca), rcp_approx(Cb));
                        const f32x2 neg1 = pack2(-1.0f, -1.0f);
                        const f32x2 rsp = mul2(mul2(sp, neg1), iC);
                        const f32x2 m2 = fma2(rsp, sp, n2);
                        const f32x2 h2 = mul2(fma2(m2, neg1, pack2(P.mcut2, P.mcut2)), iC);
                        const f32x2 gam = mul2(C, pack2(KG, KG)), al = mul2(m2, pack2(KG, KG));
                        const f32x2 tc = fma2(rsp, neg1, pack2(dD, dD));
                        float h2a, h2b;
                        unpack2(h2, h2a, h2b);
                        Ra.ok = h2a > 0.f; Rb.ok = h2b > 0.f;
                        Ra.hh = sqrt_approx(h2a) * inv_dt; Rb.hh = sqrt_approx(h2b) * inv_dt;
                        unpack2(rsp, Ra.rsp, Rb.rsp);
                        unpack2(gam, Ra.gam, Rb.gam);
                        unpack2(al, Ra.al, Rb.al);
                        unpack2(tc, Ra.tc, Rb.tc);
                        unpack2(sth, Ra.sth, Rb.sth);
                    };
                    // ---- chain rule of one ray from its moments S_n = sum_k w_k e_k tau_k^n  (tau = r_k - r*)
                    auto ray_chain = [&](const float* th, const float* ph, float rsp, float S0, float S1, float S2) {
                        const float sth = th[TH_SIN], cth = th[TH_COS], sph = ph[PH_SIN], cph = ph[PH_COS];
                        const float dx = sth * cph, dy = sth * sph, dz = cth;
                        const float bx = fmaf(M00, dx, fmaf(M01, dy, M02 * dz)), by = fmaf(M10, dx, fmaf(M11, dy, M12 * dz)),
                                    bz = fmaf(M20, dx, fmaf(M21, dy, M22 * dz));
                        const float rsu = D + rsp;
                        const float px = fmaf(rsp, bx, fmaf(D, bx, ax)), py = fmaf(rsp, by, fmaf(D, by, ay)), pz = fmaf(rsp, bz, fmaf(D, bz, az));   // perp
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
                    // the chain rule of the two rays of an iteration in packed fp32x2 (a ray outside the cut arrives with S = 0)
                    auto ray_chain2 = [&](const float* tha, const float* pha, const float* thb_, const float* phb_, f32x2 rsp, f32x2 S0, f32x2 S1,
                                          f32x2 S2) {
                        const f32x2 sth = pack2(tha[TH_SIN], thb_[TH_SIN]), cth = pack2(tha[TH_COS], thb_[TH_COS]);
                        const f32x2 sph = pack2(pha[PH_SIN], phb_[PH_SIN]), cph = pack2(pha[PH_COS], phb_[PH_COS]);
                        const f32x2 dx = mul2(sth, cph), dy = mul2(sth, sph);
                        const f32x2 bx = fma2(pack2(M00, M00), dx, fma2(pack2(M01, M01), dy, mul2(pack2(M02, M02), cth)));
                        const f32x2 by = fma2(pack2(M10, M10), dx, fma2(pack2(M11, M11), dy, mul2(pack2(M12, M12), cth)));
                        const f32x2 bz = fma2(pack2(M20, M20), dx, fma2(pack2(M21, M21), dy, mul2(pack2(M22, M22), cth)));
                        const f32x2 Dv = pack2(D, D);
                        const f32x2 px = fma2(rsp, bx, fma2(Dv, bx, pack2(ax, ax))), py = fma2(rsp, by, fma2(Dv, by, pack2(ay, ay))),
                                    pz = fma2(rsp, bz, fma2(Dv, bz, pack2(az, az)));                     // perp
                        const f32x2 ux = fma2(S1, bx, mul2(S0, px)), uy = fma2(S1, by, mul2(S0, py)), uz = fma2(S1, bz, mul2(S0, pz));
                        const f32x2 rsu = add2(Dv, rsp);
                        const f32x2 vx = fma2(rsu, ux, fma2(S2, bx, mul2(S1, px))), vy = fma2(rsu, uy, fma2(S2, by, mul2(S1, py))),
                                    vz = fma2(rsu, uz, fma2(S2, bz, mul2(S1, pz)));
                        const f32x2 nw = mul2(sth, pack2(-opalb, -opalb));
                        pgwa = fma2(sth, S0, pgwa);
                        pga[0] = fma2(nw, ux, pga[0]); pga[1] = fma2(nw, uy, pga[1]); pga[2] = fma2(nw, uz, pga[2]);
                        const f32x2 bxw = mul2(nw, vx), byw = mul2(nw, vy), bzw = mul2(nw, vz);
                        pg[0] = fma2(bxw, dx, pg[0]); pg[1] = fma2(bxw, dy, pg[1]); pg[2] = fma2(bxw, cth, pg[2]);
                        pg[3] = fma2(byw, dx, pg[3]); pg[4] = fma2(byw, dy, pg[4]); pg[5] = fma2(byw, cth, pg[5]);
                        pg[6] = fma2(bzw, dx, pg[6]); pg[7] = fma2(bzw, dy, pg[7]); pg[8] = fma2(bzw, cth, pg[8]);
                    };
                    // ---- seeds of one stream at the centre bin.  tc = t(kc) + node offset - r*
                    //      E = {e(kc), e(kc-1)} * weight ; G = {e(kc+1)/e(kc), e(kc-2)/e(kc-1)} ; q = ratio of consecutive ratios
                    auto seed = [&](float gam, float al, float tc, float weight, f32x2& E, f32x2& G, float& qv) {
                        const float gd = gam * dt;
                        const float gd2 = gd * dt, t2 = 2.0f * gd * tc;
                        const float e0 = ex2_approx(fmaf(gam * tc, tc, al)) * weight;
                        const float gu = ex2_approx(gd2 + t2);
                        const float gm = ex2_approx(gd2 - t2);
                        qv = gu * gm;
                        E = pack2(e0, e0 * gm);
                        G = pack2(gu, gm * qv);
                    };
                    // ---- two streams, nst steps.  forward: the packed {up, down} sums of the two streams go straight into the window
                    //      (exactly nst steps; 4 at a time while both directions stay inside the entry, so that the 8 loads are in flight
                    //      together); backward: prefix-sum moments against the window's weights, chain rule per stream.
                    //      (ONE instance of this loop per kernel: the hot code must stay inside the instruction cache.)
                    auto run2 = [&](f32x2 E0, f32x2 G0, float q0, f32x2 E1, f32x2 G1, float q1, int nst, float tc0, float tc1, float& S0a,
                                    float& S1a, float& S2a, float& S0b, float& S1b, float& S2b) {
                        const f32x2 Q0 = pack2(q0, q0), Q1 = pack2(q1, q1);
                        if (!BWD) {
                            float* pu = hpc;
                            float* pd = hpc - NLOS_TILE;
                            // (a step count that can be rounded up to a multiple of 4 inside the entry runs in whole blocks: the 1 - 3 extra bins per
                            //  direction lie beyond the cut and add < 8e-7 of the ray's peak, as in transient.cu, and the lanes stay together)
                            const int lim = min(n_up, n_dn), nst4 = (nst + 3) & ~3;
                            if (nst4 <= lim) nst = nst4;
                            const int nb = min(nst, lim);
                            int s = 0;
#pragma unroll 1
                            for (; s + 4 <= nb; s += 4) {
                                float u[4], d[4];
#pragma unroll
                                for (int jx = 0; jx < 4; ++jx) { u[jx] = pu[jx * NLOS_TILE]; d[jx] = pd[-jx * NLOS_TILE]; }
#pragma unroll
                                for (int jx = 0; jx < 4; ++jx) {
                                    float lo, hi;
                                    unpack2(add2(E0, E1), lo, hi);
                                    u[jx] += lo; d[jx] += hi;
                                    acc_mul2(E0, G0); acc_mul2(E1, G1);
                                    acc_mul2(G0, Q0); acc_mul2(G1, Q1);
                                }
#pragma unroll
                                for (int jx = 0; jx < 4; ++jx) { pu[jx * NLOS_TILE] = u[jx]; pd[-jx * NLOS_TILE] = d[jx]; }
                                pu += 4 * NLOS_TILE; pd -= 4 * NLOS_TILE;
                            }
#pragma unroll 1
                            for (; s < nst; ++s) {
                                float lo, hi;
                                unpack2(add2(E0, E1), lo, hi);
                                if (s < n_up) *pu += lo;
                                if (s < n_dn) *pd += hi;
                                pu += NLOS_TILE; pd -= NLOS_TILE;
                                acc_mul2(E0, G0); acc_mul2(E1, G1);
                                acc_mul2(G0, Q0); acc_mul2(G1, Q1);
                            }
                        } else {
                            f32x2 A0 = 0ull, B0 = 0ull, C0 = 0ull, A1 = 0ull, B1 = 0ull, C1 = 0ull;
                            const float* pu = hpc;
                            const float* pd = hpc - NLOS_TILE;
                            int s = 0;
                            auto block4 = [&]() {
#pragma unroll
                                for (int jx = 0; jx < 4; ++jx) {
                                    const f32x2 w2 = pack2(pu[jx * NLOS_TILE], pd[-jx * NLOS_TILE]);
                                    acc_fma2(A0, w2, E0); acc_fma2(A1, w2, E1);
                                    acc_add2(B0, A0); acc_add2(B1, A1);
                                    acc_add2(C0, B0); acc_add2(C1, B1);
                                    acc_mul2(E0, G0); acc_mul2(E1, G1);
                                    acc_mul2(G0, Q0); acc_mul2(G1, Q1);
                                }
                                pu += 4 * NLOS_TILE; pd -= 4 * NLOS_TILE;
                            };
                            // phase 1: both directions inside the entry (zmin is a multiple of 4) -- the bare loop
                            const int n1 = min(nst, zmin);
#pragma unroll 1
                            for (; s < n1; s += 4) block4();
                            // phase 2 (rare): one direction has left the entry and is switched off
#pragma unroll 1
                            for (; s < nst; s += 4) {
                                float a, b;
                                unpack2(E0, a, b); E0 = pack2(s >= zu ? 0.f : a, s >= zd ? 0.f : b);
                                unpack2(E1, a, b); E1 = pack2(s >= zu ? 0.f : a, s >= zd ? 0.f : b);
                                if (s >= zu) pu = hpc;                          // keep the (ignored) reads inside the window
                                if (s >= zd) pd = hpc;
                                block4();
                            }
                            stream_moments(A0, B0, C0, (float)s, tc0, dt, S0a, S1a, S2a);
                            stream_moments(A1, B1, C1, (float)s, tc1, dt, S0b, S1b, S2b);
                        }
                    };

                    // ---- candidate rays.  The loop over candidate blocks is WARP-UNIFORM (vote), the lanes without a block idle in it.
                    //      fast2 (one vote): every lane's rectangle is at most 2 x 2 -> ONE straight-line 2 x 2 slab test anchored at (i0, j0);
                    //      otherwise 4 x 8 blocks as transient.cu.  Slab test = cuda_utils.cuh:97-121 with every t multiplied by
                    //      sin(theta) > 0 (the x and y slabs then depend on the column only, the z slab on the row only).  That form rounds
                    //      differently from the reference's (b - o) * (1 / (d + 1e-8)): each slab value carries a relative error <= k =
                    //      1e-6 + 1.5e-8 max(1/sin theta) max|1/cos phi, 1/sin phi, tan theta| (roundings + the 1e-8 of the reference's reciprocal,
                    //      which grows with 1/|d_axis|), so the two forms can only disagree where |tmax - tmin| <= 2 k |tmax| / (1 - k): those
                    //      candidates (band 2.5 k |tmax|; every candidate of a block with k > 0.05) are decided again by the reference's own
                    //      expression (ray_aabb_x): bit-exact hit masks.
                    const bool fast2 = use_st || __all_sync(0xffffffffu, !act || (i1 - i0 <= 1 && j1 - j0 <= 1));
                    const float kband = 1.5e-8f * s_pi[PI_ISTMAX * 32 + lane];      // band of the slab test per unit of 1/|d_axis|
                    int ib = i0, jb = j0;
                    bool more = use_st ? (act && pm != 0u) : (act && i0 <= i1 && j0 <= j1);
                    unsigned stash_out = fast2 ? 0u : 0xFFFFu;
                    while (__any_sync(0xffffffffu, more)) {
                        const float* const thb = tabl + TAB_TH(A, more ? ib : 0);
                        const float* const phb = tabl + TAB_PH(A, more ? jb : 0);
                        unsigned mask = 0, border = 0;
                        if (more) {
                            if (pm) {                                   // (pm != 0 only with use_st; testing pm keeps ptxas from rotating the backward loop's accumulators)
                                mask = (pm & 3u) | ((pm & 12u) << 6);                      // rays (0,0) (0,1) (1,0) (1,1) -> bits 0, 1, 8, 9
                            } else if (fast2) {
                                const bool r1 = i1 > ib, c1 = j1 > jb;
                                const float icp0 = phb[PH_ICOS], isp0 = phb[PH_ISIN], icp1 = phb[128 + PH_ICOS], isp1 = phb[128 + PH_ISIN];
                                const float tth0 = thb[TH_TAN], tth1 = thb[96 + TH_TAN];
                                const float kblk = fmaf(fmaxf(fmaxf(fmaxf(fabsf(icp0), fabsf(isp0)), fmaxf(fabsf(icp1), fabsf(isp1))), fmaxf(fabsf(tth0), fabsf(tth1))),
                                                        kband, 1e-6f);
                                const bool allb = kblk > 0.05f;
                                const float kk = 2.5f * kblk;
                                float xymn[2], xymx[2], zmn[2], zmx[2];
                                {
                                    const float tx0 = lox * icp0, tx1 = hix * icp0, ty0 = loy * isp0, ty1 = hiy * isp0;
                                    xymn[0] = fmaxf(fminf(tx0, tx1), fminf(ty0, ty1)); xymx[0] = fminf(fmaxf(tx0, tx1), fmaxf(ty0, ty1));
                                }
                                {
                                    const float tx0 = lox * icp1, tx1 = hix * icp1, ty0 = loy * isp1, ty1 = hiy * isp1;
                                    xymn[1] = fmaxf(fminf(tx0, tx1), fminf(ty0, ty1)); xymx[1] = fminf(fmaxf(tx0, tx1), fmaxf(ty0, ty1));
                                }
                                { const float tz0 = loz * tth0, tz1 = hiz * tth0; zmn[0] = fminf(tz0, tz1); zmx[0] = fmaxf(tz0, tz1); }
                                { const float tz0 = loz * tth1, tz1 = hiz * tth1; zmn[1] = fminf(tz0, tz1); zmx[1] = fmaxf(tz0, tz1); }
#pragma unroll
                                for (int ii = 0; ii < 2; ++ii)
#pragma unroll
                                    for (int jj = 0; jj < 2; ++jj) {
                                        const float tmin = fmaxf(xymn[jj], zmn[ii]), tmax = fminf(xymx[jj], zmx[ii]);
                                        const float diff = tmax - tmin;
                                        const bool in = (ii == 0 || r1) && (jj == 0 || c1);
                                        const bool hit = in && diff >= 0.f && tmax >= 0.f;
                                        const bool bd = in && (allb || fabsf(diff) <= kk * fabsf(tmax));
                                        mask |= (hit ? 1u : 0u) << (ii * 8 + jj);
                                        border |= (bd ? 1u : 0u) << (ii * 8 + jj);
                                    }
                            } else {
                                const int ni = min(4, i1 - ib + 1), nj = min(8, j1 - jb + 1);
                                // (columns in two halves of 4: footprints are rarely wider, the second half is skipped by a branch)
#pragma unroll
                                for (int jh = 0; jh < 8; jh += 4) {
                                    if (jh < nj) {
                                        float xymn[4], xymx[4];
                                        float kmx = 0.f;
#pragma unroll
                                        for (int jj = 0; jj < 4; ++jj) {
                                            xymn[jj] = INFINITY; xymx[jj] = -INFINITY;
                                            if (jh + jj < nj) {
                                                const float icp = phb[(jh + jj) * 128 + PH_ICOS], isp = phb[(jh + jj) * 128 + PH_ISIN];
                                                const float tx0 = lox * icp, tx1 = hix * icp, ty0 = loy * isp, ty1 = hiy * isp;
                                                xymn[jj] = fmaxf(fminf(tx0, tx1), fminf(ty0, ty1));
                                                xymx[jj] = fminf(fmaxf(tx0, tx1), fmaxf(ty0, ty1));
                                                kmx = fmaxf(kmx, fmaxf(fabsf(icp), fabsf(isp)));
                                            }
                                        }
#pragma unroll
                                        for (int ii = 0; ii < 4; ++ii)
                                            if (ii < ni) kmx = fmaxf(kmx, fabsf(thb[ii * 96 + TH_TAN]));
                                        const float kblk = fmaf(kmx, kband, 1e-6f), kk = 2.5f * kblk;
                                        const bool allb = kblk > 0.05f;
#pragma unroll
                                        for (int ii = 0; ii < 4; ++ii) {
                                            if (ii < ni) {
                                                const float tth = thb[ii * 96 + TH_TAN];
                                                const float tz0 = loz * tth, tz1 = hiz * tth;
                                                const float zmn = fminf(tz0, tz1), zmx = fmaxf(tz0, tz1);
#pragma unroll
                                                for (int jj = 0; jj < 4; ++jj) {
                                                    const float tmin = fmaxf(xymn[jj], zmn);
                                                    const float tmax = fminf(xymx[jj], zmx);
                                                    const float diff = tmax - tmin;
                                                    const bool in = jh + jj < nj;
                                                    const bool hit = diff >= 0.f && tmax >= 0.f;
                                                    const bool bd = in && (allb || fabsf(diff) <= kk * fabsf(tmax));
                                                    mask |= (hit ? 1u : 0u) << (ii * 8 + jh + jj);
                                                    border |= (bd ? 1u : 0u) << (ii * 8 + jh + jj);
                                                }
                                            }
                                        }
                                    }
                                }
                            }
                            // ---- borderline candidates: the reference's own expression decides (cuda_utils.cuh:97-121, d = (sin th cos ph, sin th sin ph, cos th))
                            while (border) {
                                const int bb = __ffs(border) - 1;
                                border &= border - 1;
                                const float* const thp = thb + (bb >> 3) * 96;
                                const float* const php = phb + (bb & 7) * 128;
                                const float box[6] = {v3.y, v3.z, v3.w, v4.x, v4.y, v4.z};
                                const bool h = ray_aabb_x(ox, oy, oz, __fmul_rn(thp[TH_SIN], php[PH_COS]), __fmul_rn(thp[TH_SIN], php[PH_SIN]), thp[TH_COS], box);
                                mask = h ? (mask | (1u << bb)) : (mask & ~(1u << bb));
                            }
                            if (st_wr && fast2) stash_out = (unsigned)ib | ((unsigned)jb << 6) | (((mask & 3u) | ((mask >> 6) & 12u)) << 12);
                            if (P.hit_dump) {      // (tests only) dense membership [scan point][ray][Gaussian], compared bit for bit with the reference's filter lists
                                for (unsigned m = mask; m; m &= m - 1) {
                                    const int bb = __ffs(m) - 1;
                                    const size_t ray = (size_t)(ib + (bb >> 3)) * A + jb + (bb & 7);
                                    P.hit_dump[(((size_t)tile * NLOS_TILE + lane) * A * A + ray) * P.N + (ent & ((1u << NLOS_ENTRY_SHIFT) - 1u))] = 1;
                                }
                            }
                        }
                            // ---- 2. consume the hits.  Rays the recurrence cannot carry (seed not representable, narrower than a bin)
                            //         are deferred to ONE slow loop below
                            unsigned slow_full = force_slow ? mask : 0u;
                            if (force_slow) mask = 0u;
                            while (mask) {
                                const int b0 = __ffs(mask) - 1;
                                mask &= mask - 1;
                                const float* const th0p = thb + (b0 >> 3) * 96;
                                const float* const ph0p = phb + (b0 & 7) * 128;
                                float S0a = 0.f, S1a = 0.f, S2a = 0.f, S0b = 0.f, S1b = 0.f, S2b = 0.f;
                                // ---- two rays per iteration.  A missing second ray, or a ray outside the cut, rides along as a zero-weight
                                //      copy of its partner (finite seeds, no contribution)
                                const bool has1 = mask != 0;
                                const int b1 = has1 ? __ffs(mask) - 1 : b0;
                                mask &= mask - 1;
                                const float* const th1p = thb + (b1 >> 3) * 96;
                                const float* const ph1p = phb + (b1 & 7) * 128;
                                Ray R0, R1;
                                ray_setup2(th0p, ph0p, th1p, ph1p, R0, R1);
                                const bool ok0 = R0.ok, ok1 = has1 && R1.ok;
                                if (!(ok0 || ok1)) continue;
                                if (!ok0) { R0.gam = R1.gam; R0.al = R1.al; R0.tc = R1.tc; R0.hh = R1.hh; }
                                if (!ok1) { R1.gam = R0.gam; R1.al = R0.al; R1.tc = R0.tc; R1.hh = R0.hh; }
                                const unsigned bits = (ok0 ? 1u << b0 : 0u) | (ok1 ? 1u << b1 : 0u);
                                const float off = (MODE == MODE_ANA) ? 0.5f : 0.0f;                  // analytic: the nodes sit up to half a bin off
                                const float at0 = fabsf(R0.tc) + off * dt, at1 = fabsf(R1.tc) + off * dt;
                                const float hx = fmaxf(fmaf(at0, inv_dt, R0.hh), fmaf(at1, inv_dt, R1.hh));   // steps - 1 that reach the cut
                                const float Lmin = fminf(fmaf(R0.gam * at0, at0, R0.al), fmaf(R1.gam * at1, at1, R1.al));
                                const float gpm = -fminf(R0.gam, R1.gam) * dt2;                      // 0.72 / (sigma in bins)^2 of the narrower ray
                                // seeds representable, recurrence well conditioned (analytic: quadrature accurate)
                                if (!(Lmin > -100.f && gpm < (MODE == MODE_ANA ? 2.0f : 6.0f))) { slow_full |= bits; continue; }
                                const int nst = min((int)hx + 1, n_lim);
                                const float w0 = ok0 ? (BWD ? 1.0f : opalb * R0.sth) : 0.f, w1 = ok1 ? (BWD ? 1.0f : opalb * R1.sth) : 0.f;
                                if (MODE == MODE_ANA) {
                                    // ---- analytic-exact: the bin integral is the Gaussian convolved with the unit box of the bin.  The box is
                                    //      replaced by a mixture of Gaussians with the same low moments, so every node is ONE stream of the same
                                    //      recurrence on a Gaussian whose variance is inflated by s2 (precision C -> C / u, u = 1 + C s2,
                                    //      amplitude u^-1/2):
                                    //        kind 0 (both rays wider than 2.5 bins): 1 node, s2 = dt^2/12 (moments to order 3, <= 2.6e-5 of the peak);
                                    //                the two rays of the pair are the two streams of ONE pass
                                    //        kind 1 (wider than 0.8 bins): 2 nodes +-x, x^4 = dt^4/240, s2 = dt^2/12 - x^2 (moments to order 5: the
                                    //                accuracy of GL3, <= 2.5e-5 of the peak, at the cost of GL2).  Passes: {A+, B+}, {A-, B-}
                                    //        kind 2: plain GL4 (s2 = 0).  Passes: {A+, B+}, {A-, B-} for each of the two node magnitudes.
                                    //      Without a second ray the two streams of a pass are the mirrored nodes {A+, A-} (half the passes).
                                    const int kind = (one_pass || gpm <= TR_GPM_K0) ? 0 : (gpm <= TR_GPM_K1 ? 1 : 2);      // (one-pass entries: see analytic_filter_kernel)
                                    const bool two = ok0 && ok1;
                                    const float s2k = kind == 0 ? dt2 * (1.0f / 12.0f) : (kind == 1 ? dt2 * 0.018783611f : 0.f);
                                    float iu0, iu1;
                                    {
                                        // x = C s2 <= 0.03: 1/u and log2(u) by their series (errors < 1e-6), both rays at once (packed)
                                        const float s2c = s2k * (1.0f / KG);                              // (a constant multiply, not an IEEE division)
                                        const f32x2 one = pack2(1.0f, 1.0f);
                                        const f32x2 gm2 = pack2(R0.gam, R1.gam);
                                        const f32x2 nx = mul2(gm2, pack2(-s2c, -s2c));                                   // -x
                                        const f32x2 iu = fma2(nx, fma2(nx, add2(one, nx), one), one);                    // 1 - x + x^2 - x^3
                                        const f32x2 lg = fma2(nx, fma2(nx, pack2(1.0f / 3.0f, 1.0f / 3.0f), pack2(0.5f, 0.5f)), one);   // ln(u) / x
                                        const f32x2 al2 = fma2(mul2(nx, pack2(0.5f * LOG2E_F, 0.5f * LOG2E_F)), lg, pack2(R0.al, R1.al));
                                        const f32x2 gi = mul2(gm2, iu);
                                        unpack2(iu, iu0, iu1);
                                        unpack2(al2, R0.al, R1.al);
                                        unpack2(gi, R0.gam, R1.gam);
                                    }
                                    // stream 0 is ray A, stream 1 ray B: after the copy step above an invalid ray IS its partner (same gam, al, tc),
                                    // only the weights differ.  With two rays the passes walk the nodes (+x, -x, ..); with one ray stream 1 is the
                                    // same ray at the mirrored node (kind 0: weight 0)
                                    const Ray& RA = R0;
                                    const float wA = (ok0 ? w0 : w1) * dt;
                                    const float gB = R1.gam, aB = R1.al, tB = R1.tc;
                                    const float wB = two ? w1 * dt : (kind == 0 ? 0.f : wA);
                                    const int npass = kind == 0 ? 1 : (two ? 2 * kind : kind);
                                    float SA0 = 0.f, SA1 = 0.f, SA2 = 0.f, SB0 = 0.f, SB1 = 0.f, SB2 = 0.f;
#pragma unroll 1
                                    for (int ps = 0; ps < npass; ++ps) {
                                        const int nd = two ? (ps >> 1) : ps;                                     // node of this pass
                                        float xn = kind == 0 ? 0.f : (kind == 1 ? 0.25406644f : (nd ? 0.43056815579702626f : 0.16999052179242816f));
                                        const float wn = kind == 0 ? 1.0f : (kind == 1 ? 0.5f : (nd ? 0.17392742256872692f : 0.32607257743127305f));
                                        xn *= (two && (ps & 1)) ? -dt : dt;
                                        const float t0s = RA.tc + xn, t1s = two ? tB + xn : tB - xn;
                                        f32x2 E0, G0, E1, G1;
                                        float q0, q1;
                                        seed(RA.gam, RA.al, t0s, wA * wn, E0, G0, q0);
                                        seed(gB, aB, t1s, wB * wn, E1, G1, q1);
                                        float m10 = 0.f, m11 = 0.f, m12 = 0.f;
                                        run2(E0, G0, q0, E1, G1, q1, nst, t0s, t1s, SA0, SA1, SA2, m10, m11, m12);
                                        if (BWD) {
                                            if (two) { SB0 += m10; SB1 += m11; SB2 += m12; } else { SA0 += m10; SA1 += m11; SA2 += m12; }
                                        }
                                    }
                                    if (BWD) {
                                        // chain rule through the inflation:  d ln e' = -d(m2min)/2 - tau^2 dC / (2 u^2) + C tau dr* / u - s2 dC / (2 u),
                                        // i.e. the moments the plain chain rule expects are  S0, S1 / u, S2 / u^2 + S0 s2 / u
                                        const float iuA = iu0, iuB = iu1;                              // (equal when a ray is invalid)
                                        SA2 = fmaf(SA2 * iuA, iuA, SA0 * (s2k * iuA)); SA1 *= iuA;
                                        SB2 = fmaf(SB2 * iuB, iuB, SB0 * (s2k * iuB)); SB1 *= iuB;
                                        // (ray A is R1 when R0 missed the cut: its moments then belong to the second half)
                                        if (ok0) ray_chain2(th0p, ph0p, th1p, ph1p, pack2(R0.rsp, R1.rsp), pack2(SA0, SB0), pack2(SA1, SB1), pack2(SA2, SB2));
                                        else ray_chain2(th0p, ph0p, th1p, ph1p, pack2(R0.rsp, R1.rsp), pack2(0.f, SA0), pack2(0.f, SA1), pack2(0.f, SA2));
                                    }
                                } else {
                                    // ---- numerical mode: the two rays are the two streams
                                    f32x2 E0, G0, E1, G1;
                                    float q0, q1;
                                    seed(R0.gam, R0.al, R0.tc, w0, E0, G0, q0);
                                    seed(R1.gam, R1.al, R1.tc, w1, E1, G1, q1);
                                    run2(E0, G0, q0, E1, G1, q1, nst, R0.tc, R1.tc, S0a, S1a, S2a, S0b, S1b, S2b);
                                    if (BWD)
                                        ray_chain2(th0p, ph0p, th1p, ph1p, pack2(R0.rsp, R1.rsp), pack2(S0a, S0b), pack2(S1a, S1b), pack2(S2a, S2b));
                                }
                            }
                            // ---- 3. slow loop: direct evaluation of whole rays -- numerical: 1 MUFU per bin; analytic: closed-form erf per bin edge
                            unsigned slow = slow_full;
                            while (slow) {
                                const int b0 = __ffs(slow) - 1;
                                slow &= slow - 1;
                                const float* const th0p = thb + (b0 >> 3) * 96;
                                const float* const ph0p = phb + (b0 & 7) * 128;
                                Ray R;
                                ray_setup(th0p, ph0p, R);
                                if (!R.ok) continue;
                                const float xc = fkc - R.tc * inv_dt;                                           // (r* - t0) / dt
                                const float off = (MODE == MODE_ANA) ? 0.5f : 0.0f;                              // bins are [k-.5, k+.5] for the integral
                                const int kl = (int)fmaxf(ceilf(xc - R.hh - off), fRlo), kh = (int)fminf(floorf(xc + R.hh + off), fRhi);
                                if (kl > kh) continue;
                                const float wt = opalb * R.sth;
                                const float rs = D + R.rsp;
                                float S0 = 0.f, S1 = 0.f, S2 = 0.f;
                                {
                                    const int ka = kl, kb = kh;
                                    float* hp = Hw + (ka - base) * NLOS_TILE + lane;
                                    if (MODE == MODE_NUM) {
                                        float t = fmaf((float)ka, dt, t0) - rs;
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
                                    } else {
                                        // integral of exp(-C tau^2 / 2) e^{-m2min/2} over bin k = pref [erf(sq tau1) - erf(sq tau0)]  
                                        const float C = R.gam * (-2.0f / LOG2E_F);
                                        const float sq = sqrtf(0.5f * C), iC = 1.0f / C;
                                        const float pk = ex2_approx(R.al);                                       // e^{-m2min/2}
                                        const float pref = SQRT_PI_2 * rsqrt_approx(C) * pk;
                                        float tau = fmaf((float)ka - 0.5f, dt, t0) - rs;                         // lower edge of bin ka
                                        float Ep = erf_fast(sq * tau);
                                        float fp = BWD ? pk * ex2_approx(R.gam * tau * tau) : 0.f;
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
                                }
                                if (BWD) ray_chain(th0p, ph0p, R.rsp, S0, S1, S2);
                            }
                        if (more) {
                            if (fast2) {
                                more = false;
                            } else {
                                jb += 8;
                                if (jb > j1) { jb = j0; ib += 4; more = ib <= i1; }
                            }
                        }
                    }
                    if (st_wr) P.hit_stash[(size_t)(beg + b * TR_RING + q) * NLOS_TILE + lane] = (unsigned short)stash_out;
                    if (GRAD) {
                        {
                            gax += lo_of(pga[0]) + hi_of(pga[0]); gay += lo_of(pga[1]) + hi_of(pga[1]); gaz += lo_of(pga[2]) + hi_of(pga[2]);
                            g00 += lo_of(pg[0]) + hi_of(pg[0]); g01 += lo_of(pg[1]) + hi_of(pg[1]); g02 += lo_of(pg[2]) + hi_of(pg[2]);
                            g10 += lo_of(pg[3]) + hi_of(pg[3]); g11 += lo_of(pg[4]) + hi_of(pg[4]); g12 += lo_of(pg[5]) + hi_of(pg[5]);
                            g20 += lo_of(pg[6]) + hi_of(pg[6]); g21 += lo_of(pg[7]) + hi_of(pg[7]); g22 += lo_of(pg[8]) + hi_of(pg[8]);
                            gwa += lo_of(pgwa) + hi_of(pgwa);
                        }
                        // a = M (o - mu) = -M c :  dL/dM += dL/da (x) (o - mu) ;  dL/dmu = -M^T dL/da
                        float v[16];
                        v[0] = fmaf(-gax, cx, g00); v[1] = fmaf(-gax, cy, g01); v[2] = fmaf(-gax, cz, g02);
                        v[3] = fmaf(-gay, cx, g10); v[4] = fmaf(-gay, cy, g11); v[5] = fmaf(-gay, cz, g12);
                        v[6] = fmaf(-gaz, cx, g20); v[7] = fmaf(-gaz, cy, g21); v[8] = fmaf(-gaz, cz, g22);
                        v[9] = -(M00 * gax + M10 * gay + M20 * gaz);
                        v[10] = -(M01 * gax + M11 * gay + M21 * gaz);
                        v[11] = -(M02 * gax + M12 * gay + M22 * gaz);
                        v[12] = gwa * alb;                                       // dL/dop      (gwa = dL/d(op * albedo) of this pair)
                        v[13] = gwa * v0.w;                                      // dL/dalbedo  (degree 0: one albedo per Gaussian)
                        v[14] = 0.f; v[15] = 0.f;
                        const int g = (int)(ent & ((1u << NLOS_ENTRY_SHIFT) - 1u));
                        const int slot = lane >> 1;
                        if (SH) {
                            // view-dependent albedo: dL/d(SH coefficients) = sum over the scan points of dpre Y_k(dir), and the direction's own
                            // dependence on the mean goes into dL/dmu (slots 9..11)
                            float dshv[16];
                            const float dpre = alb > 0.f ? v[13] : 0.f;
                            v[13] = 0.f;
                            float gmx = 0.f, gmy = 0.f, gmz = 0.f;          // (scalars: v[] must not have its address taken, it lives in registers)
                            sh_pair_backward(P.sh + (size_t)g * P.sh_dim, P.sh_degree, cx, cy, cz, D, dpre, dshv, gmx, gmy, gmz);
                            v[9] += gmx; v[10] += gmy; v[11] += gmz;
                            const float sval = warp_reduce16(dshv, lane);
                            if (!(lane & 1) && sval != 0.f) atomicAdd(P.dsh + (size_t)tile * P.part_stride + (size_t)g * 16 + slot, sval);
                        }
                        // slots: 0..8 dM, 9..11 dmu, 12 dop, 13 dalb (degree 0)
                        const float val = warp_reduce16(v, lane);
                        if (!(lane & 1) && slot < 14 && val != 0.f) atomicAdd(P.dacc + (size_t)tile * P.part_stride + (size_t)g * NLOS_GRAD_FLOATS + slot, val);
                    }
                }
                __syncwarp();
            }
            if (FWD && base != NO_BASE) {
                __syncwarp();
                flush(W);
                __syncwarp();
            }
        }
        __syncthreads();   // every warp is done with this tile's tables (and with s_ctl)
    }

}

// ---- launch (called by the C-ABI entry points in transient.cu)
template <int MODE, bool BWD, bool SH = false>
static int launch_narrow(const TransientParams& P, cudaStream_t s, const char* name)
{
    if (!SH && P.sh_degree > 0) return launch_narrow<MODE, BWD, true>(P, s, name);     // view-dependent albedo: its own instantiation
    int n_sm = transient_num_sms();
    cudaError_t em = cudaMemsetAsync(P.work, 0, (size_t)P.n_tiles * P.n_seg * sizeof(int), s);
    NLOS_REQUIRE(em == cudaSuccess, "%s: memset: %s", name, cudaGetErrorString(em));
    size_t smem = transient_smem_bytes(P.A, P.window);
    NLOS_REQUIRE(smem <= 227 * 1024, "%s: %zu B of shared memory needed (A=%d, window=%d) > 227 KB", name, smem, P.A, P.window);
    cudaError_t e = cudaFuncSetAttribute(transient_narrow_kernel<MODE, BWD, SH>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    NLOS_REQUIRE(e == cudaSuccess, "%s: cudaFuncSetAttribute: %s", name, cudaGetErrorString(e));
    transient_narrow_kernel<MODE, BWD, SH><<<n_sm, TR_WPB * 32, smem, s>>>(P);
    NLOS_CHECK_LAUNCH(name);
    return 0;
}

int launch_transient_narrow(const TransientParams& P, int mode, bool bwd, cudaStream_t s)
{
    if (!bwd)
        return mode == MODE_NUM ? launch_narrow<MODE_NUM, false>(P, s, "transient_narrow_kernel<num,fwd>")
                                : launch_narrow<MODE_ANA, false>(P, s, "transient_narrow_kernel<ana,fwd>");
    return mode == MODE_NUM ? launch_narrow<MODE_NUM, true>(P, s, "transient_narrow_kernel<num,bwd>")
                            : launch_narrow<MODE_ANA, true>(P, s, "transient_narrow_kernel<ana,bwd>");
}
