/* nlosb200.h -- C ABI of libnlosb200.so, the B200 (sm_100a) transient renderer.
 *
 * Drop-in boundary for the reference's pybind module `nlos_gaussian_renderer._C`
 * (/root/reference/submodules/cuda_renderer/src/bindings.cpp:26-35) and for the
 * torch.autograd.Function that wraps it (gaussian_model/cuda_autograd.py:27-191).
 *
 * Conventions: every pointer is a DEVICE pointer to contiguous fp32 / int32 data unless
 * stated; `stream` is a cudaStream_t passed as void*; calls are stream-ordered and never
 * synchronise the host; return 0 on success, <0 on error with the message available from
 * nlosb200_last_error() (thread local).  Inputs are borrowed, outputs are caller-allocated.
 */
#ifndef NLOSB200_H
#define NLOSB200_H
#ifdef __cplusplus
extern "C" {
#endif

const char* nlosb200_last_error(void);
int nlosb200_version(void);
int nlosb200_record_floats(void);            /* 32: floats per preprocessed Gaussian record          */
int nlosb200_grad_floats(void);              /* 16: floats per Gaussian in the packed gradient buffer */
long long nlosb200_tile_floats(int A);       /* floats per scan tile block for an A x A lattice       */
int nlosb200_transient_warps_per_block(void);
int nlosb200_bin_num_chunks(int N);

/* ---- per-Gaussian activation / packing -------------------------------------------------------
 * Replaces compute_gaussian_bboxes_kernel (include/bbox_compute.cuh:76-120) and the per-sample
 * re-activation inside volume_render_kernel (src/volume_renderer.cu:86-111).
 * records [N,32] (128-B aligned), shell [N,4] = (mean, mcut*max scale; negated when the smallest scale is >= time_wide_min > 0: the
 * "time-wide" flag of the window binning), bboxes [N,6]; each nullable. */
int nlosb200_preprocess(const float* means, const float* log_scales, const float* rotations, const float* opacities,
                        const float* features, int N, int sh_dim, float scaling_modifier, float sigma_bbox, float mcut,
                        float time_wide_min, float* records, float* shell, float* bboxes, void* stream);
/* dacc [N,16] = {dL/dM(9), dL/dmu(3), dL/dop, dL/dalb0, 0, 0} -> raw parameter gradients
 * (the backward the reference leaves as zeros, gaussian_model/cuda_autograd.py:139-172). */
int nlosb200_preprocess_backward(const float* log_scales, const float* rotations, const float* opacities, const float* features,
                                 int N, int sh_dim, float scaling_modifier, const float* dacc, float* d_means,
                                 float* d_log_scales, float* d_rotations, float* d_opacities, float* d_features, int accumulate,
                                 void* stream);
/* the same with view-dependent albedo (sh_degree 1..3): d_features[:, :(sh_degree+1)^2] come from dsh [N,16] (nlosb200_transient_backward) */
int nlosb200_preprocess_backward_sh(const float* log_scales, const float* rotations, const float* opacities, const float* features,
                                    int N, int sh_dim, int sh_degree, float scaling_modifier, const float* dacc, const float* dsh, float* d_means,
                                    float* d_log_scales, float* d_rotations, float* d_opacities, float* d_features, int accumulate,
                                    void* stream);
/* GaussianModel.get_bboxes (gaussian_model/gaussian_model.py:140-178) / bbox kernel: [N,6] min xyz, max xyz */
int nlosb200_compute_bboxes(const float* means, const float* log_scales, const float* rotations, int N, float scaling_modifier,
                            float sigma, float* bboxes, void* stream);

/* ---- scan tiles: 32 consecutive scan points per tile, lattice of cuda_autograd.py:254-268 ------
 * scan_o [S,3]; sin_th/cos_th/sin_ph/cos_ph [S,A]; lattice [S,4] = (theta0, dtheta, phi0, dphi) of
 * the linspace; hscale [S] = dtheta*dphi*vp_y^2 (cuda_autograd.py:311-314, nlos_helpers.py:275-276);
 * valid int32 [S] nullable (0 = padding lane of a ragged tile);
 * tiles out: ceil(S/32) * nlosb200_tile_floats(A) floats, 128-B aligned. */
int nlosb200_build_scan_tiles(const float* scan_o, const float* sin_th, const float* cos_th, const float* sin_ph,
                              const float* cos_ph, const float* lattice, const float* hscale, const int* valid, int S, int A,
                              float* tiles, void* stream);

/* ---- window binning (replaces filter_gaussians_kernel, src/ray_aabb.cu:10-102, as the culling structure of the batched
 * renderer).  Per (tile, Gaussian) the conservative radial shell is an inclusive bin interval [klo,khi]; the renderer's
 * accumulator is a sliding window of `window` bins (multiple of 8) advancing in steps of `bucket` bins, so a Gaussian is
 * listed once per tile unless its interval is longer than L = window-bucket (then entry e covers [klo+eL, klo+(e+1)L-1]).
 * Entry (g,e) lives in cell (tile, (klo+eL)/bucket); list value = g | e<<26; per cell ascending g.
 * n_buckets = ceil(T/bucket) <= 64.  Entry classes: with n_classes = 2 the cells of a tile are class-major (n_cells = n_tiles * n_classes *
 * n_buckets, cell = tile * 2 n_buckets + cls * n_buckets + bucket); an entry is class 0 ("narrow") when its Gaussian is listed once in the tile
 * and (6 / mcut) rad <= narrow_rays * pitch * min-distance, pitch = the smallest lattice step of the tile's scan points -- i.e. its 3-sigma AABB
 * spans at most narrow_rays lattice pitches -- else class 1 (narrow_rays = 0: no footprint classes).  time_wide = 1 adds an independent class
 * bit: the shell carries the time-wide flag (nlosb200_preprocess) and the interval keeps 3 bins from both ends of the time axis.  Class index =
 * footprint bit + n_fp * time bit; n_classes = n_fp * n_tw must be passed consistently (1, 2 or 4).
 * chunk_counts int32 [n_cells * nlosb200_bin_num_chunks(N)], cell_counts int32 [n_cells], cell_offsets int64 [n_cells+1],
 * tile_items int32 [n_tiles+1], header int64 [4] = {total list entries, CTA work items, overflow flag, -}. */
int nlosb200_bin_window_plan(const float* tiles, int n_tiles, int A, const float* shell, int N, float t0, float dt, int T,
                             int window, int bucket, int n_classes, float narrow_rays, float mcut, int time_wide, int slice_len, int warps_per_block,
                             int* chunk_counts, int* cell_counts, long long* cell_offsets, int* tile_items, long long* header, void* stream);
int nlosb200_bin_window_fill(const float* tiles, int n_tiles, int A, const float* shell, int N, float t0, float dt, int T,
                             int window, int bucket, int n_classes, float narrow_rays, float mcut, int time_wide, const int* chunk_counts_scanned,
                             const long long* cell_offsets, int* list, int* list_bins, long long list_capacity, long long* header,
                             void* stream);

/* ---- batched transient H[n_tiles*32, T] and its backward.  mode 0 = numerical (volume_renderer.cu
 * no-occlusion branch :139-182), 1 = analytic-exact per-bin integral.
 * bin k sits at radius t0 + k*dt.  list_bins[i] = klo | khi << 16 is the tile-level shell of entry i as written by
 * nlosb200_bin_window_fill (T <= 65535); list_capacity = entries the two list buffers hold (a launch whose lists do not fit renders
 * nothing: the overflow flag of bin_window_fill tells).  The cells of a tile are class-major: n_classes * n_buckets per tile, a launch
 * walks class `cls` (nlosb200_bin_window_plan).  variant 0 = register-accumulator kernel (wide footprints, transient.cu), 1 = direct-to-window
 * kernel (footprints of a few rays, transient_narrow.cu); both render any entry, they differ in speed only.  accumulate = 0: H / dacc /
 * counters are zeroed by the callee first; 1: the launch adds to them (further classes).  A launch may walk a SECOND class in the same grid
 * (cls2 >= 0, else -1) with its own one_pass flag and output rows H2 / input rows grad_H2 (may equal H / grad_H): the persistent CTAs steal
 * from both classes' lists, so nothing idles between them.  `work` is int32 scratch [2*n_tiles] (zeroed by the callee): the per-(class, tile)
 * slice counters the persistent CTAs pull work from; slice_len = list entries a warp takes at a time.
 * View-dependent albedo: sh_degree 1..3 with sh = the raw SH coefficients [N, sh_dim], sh_dim >= (sh_degree+1)^2 (fewer is an error: the
 * reference reads out of bounds, volume_renderer.cu:110): albedo(g, p) = max(SH(dir(mu_g - o_p)) + 0.5, 0) per (Gaussian, scan point)
 * (volume_renderer.cu:107-111, spherical_harmonics.cuh:20-80); the backward then also fills dsh [N,16] = dL/d(SH coefficients) and adds the
 * direction's dependence on the mean to dL/dmu.  sh_degree 0: sh / dsh may be NULL, the degree-0 albedo of the record is used. */
int nlosb200_transient_forward(const float* tiles, int n_tiles, int A, const float* records, int N, const int* list,
                               const int* list_bins, long long list_capacity, const long long* cell_offsets, int n_classes, int cls,
                               int* work, float t0, float dt, int T, int window, int bucket, int slice_len, float c_deltaT, float mcut,
                               int mode, int variant, int one_pass, int accumulate, const float* sh, int sh_dim, int sh_degree, float* H,
                               int cls2, int one_pass2, float* H2, unsigned short* hit_stash, void* stream);
int nlosb200_transient_backward(const float* tiles, int n_tiles, int A, const float* records, int N, const int* list,
                                const int* list_bins, long long list_capacity, const long long* cell_offsets, int n_classes, int cls,
                                int* work, float t0, float dt, int T, int window, int bucket, int slice_len, float c_deltaT, float mcut,
                                int mode, int variant, int one_pass, int accumulate, const float* sh, int sh_dim, int sh_degree,
                                const float* grad_H, int cls2, int one_pass2, const float* grad_H2, float* dacc, float* dsh,
                                int per_tile_partials, const unsigned short* hit_stash, void* stream);
/* nlosb200_transient_forward for callers that may not allocate (code under CUDA graph capture, serving loops, harnesses that forbid
 * cudaMalloc).  It has NO counterpart in the reference.  Same arguments and results as
 * nlosb200_transient_forward without `work` and `hit_stash`: the slice counters live in a static device array inside the library
 * (nlosb200_transient_static_work_ints() = 65536 int32 = 256 KB per GPU that loads the library, held until exit), so the call performs no
 * cudaMalloc / cudaFree and calls no allocating routine.  Limit: n_tiles * (2 if cls2 >= 0 else 1) <= 65536, otherwise an error is returned
 * and nothing is launched.  Every call on a device shares the array (the counters a launch uses are zeroed first, stream ordered): calls
 * that overlap on the GPU (different streams or host threads) race -- serialise them, or use nlosb200_transient_forward with own buffers. */
int nlosb200_transient_static_work_ints(void);
int nlosb200_transient_forward_static(const float* tiles, int n_tiles, int A, const float* records, int N, const int* list,
                                      const int* list_bins, long long list_capacity, const long long* cell_offsets, int n_classes, int cls,
                                      float t0, float dt, int T, int window, int bucket, int slice_len, float c_deltaT, float mcut,
                                      int mode, int variant, int one_pass, int accumulate, const float* sh, int sh_dim, int sh_degree, float* H,
                                      int cls2, int one_pass2, float* H2, void* stream);
/* Culling words (variant 1 only, A <= 64; NULL = off): hit_stash [list_capacity][32] uint16, 16-byte aligned.  A forward launch writes, for every
 * (list entry, scan point) it renders, the anchor of the candidate block and the bit-exact hit mask of its <= 2 x 2 rays (i0 | j0 << 6 | mask << 12;
 * 0xFFFF marks an entry with a wider rectangle); a backward launch over the SAME lists and records reads them instead of repeating the footprint
 * rectangle and the slab tests (-15 % of its instructions on cfg4).  The caller passes the pointer to the backward only after a forward of the same
 * prepare() has filled it. */
/* One-pass analytic plan (mode 1, launches over the TIME-WIDE entry classes of the binning: smallest scale >= 1.3 bins): one_pass = 1 renders
 * every ray with the 1-node mixture and leaves the 1/(r^2+1e-8) weight out -- H is then a buffer Ha of unweighted rows, and
 *   forward : nlosb200_analytic_filter(Ha, H, rows, T, t0, dt, 0)   adds  w F(Ha)  to H   (F = the fixed 7-tap filter that turns 1-node rows into
 *             bin integrals to <= 1.6e-5 of a ray's peak; w = 1/(r^2+1e-8));
 *   backward: nlosb200_analytic_filter(gH, gHa, rows, T, t0, dt, 1) gives  gHa = F(w gH), the grad_H of the one_pass = 1 backward launches. */
int nlosb200_analytic_filter(const float* in, float* out, long long rows, int T, float t0, float dt, int backward, void* stream);
/* Deterministic mode (bit-identical results run to run): the float REDs of the render kernels are the only order-dependent operations.
 * Forward: launch with slice_len >= the longest tile list, so that ONE warp walks a tile's list in order and is the only writer of the tile's
 * rows of H.  Backward: additionally per_tile_partials = 1 with dacc (and dsh) = [n_tiles][N][16] partial buffers -- every tile accumulates into
 * its own copy -- followed by nlosb200_reduce_tile_partials, which sums them in tile order into [N][16]. */
int nlosb200_reduce_tile_partials(const float* partial, int n_tiles, long long n, float* out, void* stream);
/* counters uint64 [4]: hit (g,ray) pairs, W_s, sample-evals evaluated, W_b */
int nlosb200_transient_count(const float* tiles, int n_tiles, int A, const float* records, int N, const int* list,
                             const int* list_bins, long long list_capacity, const long long* cell_offsets, int n_classes, int cls,
                             int* work, float t0, float dt, int T, int window, int bucket, int slice_len, float mcut, int accumulate,
                             unsigned long long* counters, void* stream);

/* tests only: while `hits` is non-NULL every render launch also marks the (scan point, ray, Gaussian) triples that pass its slab test in the
 * dense byte array hits [n_tiles*32][A*A][N] (caller zeroes) -- compared bit for bit with the reference's per-ray filter lists
 * (src/ray_aabb.cu:10-61).  Pass NULL to switch the dump off. */
int nlosb200_debug_hit_dump(unsigned char* hits);

/* ==== one-scan-point operators: the reference's pybind module nlos_gaussian_renderer._C ===========================
 * (submodules/cuda_renderer/src/bindings.cpp:26-35).  A Python/torch caller allocates the outputs and passes
 * data_ptr()s; see nlos_gaussian_renderer_b200/_C.py for the exact mirror of the three pybind signatures. */

/* _C.filter_gaussians_per_ray (include/ray_aabb.h:7-13, src/ray_aabb.cu:10-102): per ray the ascending ids of the
 * Gaussians whose AABB passes the slab test of cuda_utils.cuh:97-121, cut at `cap` (reference: 256).
 * bboxes [N,6] 16-B aligned; out int32 [n_rays, cap+1], column 0 = count, unused slots = -1. */
int nlosb200_filter_gaussians_per_ray(const float* ray_origins, const float* ray_directions, const float* bboxes, int n_rays,
                                      int N, int cap, int* out, void* stream);
/* per-scan-point albedo max(SH(dir(mean-cam))+0.5,0) into records[:,19] (volume_renderer.cu:107-111; CUDA sign convention of
 * spherical_harmonics.cuh:20-80).  sh_dim < (degree+1)^2 is an error (the reference reads out of bounds). */
int nlosb200_compat_albedo(const float* means, const float* features, const float* camera_pos, int N, int sh_dim,
                           int sh_degree, float* records, void* stream);
/* backward of nlosb200_compat_albedo for active_sh_degree >= 1: dacc[g][13] = dL/d(albedo_g) (nlosb200_render_rays_backward) ->
 * d_features [N, sh_dim] (overwritten) and d_means [N,3] += the dependence of the view direction on the mean.  Call after
 * nlosb200_preprocess_backward. */
int nlosb200_compat_albedo_backward(const float* means, const float* features, const float* camera_pos, int N, int sh_dim,
                                    int sh_degree, const float* dacc, float* d_means, float* d_features, void* stream);
/* _C.render_rays (include/volume_renderer.h:8-24, src/volume_renderer.cu:16-305) after nlosb200_preprocess (records, bboxes with
 * sigma 3.0 as volume_renderer.cu:232,244), nlosb200_compat_albedo and nlosb200_filter_gaussians_per_ray:
 * rho_density, density, transmittance each [n_rays, T]. */
int nlosb200_render_rays(const float* ray_origins, const float* ray_directions, const float* t_samples, const float* records,
                         const int* filter, int cap, int n_rays, int T, int N, float c, float deltaT, int use_occlusion,
                         float* rho_density, float* density, float* transmittance, void* stream);
/* backward of render_rays -- CUDARenderFunction.backward (gaussian_model/cuda_autograd.py:110-191 returns zeros).
 * grad_* [n_rays,T] nullable; rho_density/transmittance = saved forward outputs (needed with occlusion);
 * dacc [N,16] (zeroed by the callee) -> nlosb200_preprocess_backward. */
int nlosb200_render_rays_backward(const float* ray_origins, const float* ray_directions, const float* t_samples,
                                  const float* records, const int* filter, int cap, int n_rays, int T, int N, float c,
                                  float deltaT, int use_occlusion, const float* rho_density, const float* transmittance,
                                  const float* grad_rho, const float* grad_density, const float* grad_transmittance, float* dacc,
                                  void* stream);
/* _C.render_rays_analytic (src/bindings.cpp:6-24, src/volume_renderer_analytic.cu:23-241): one scalar per ray;
 * max_sections <= 128 (MAX_SECTIONS_PER_RAY). */
int nlosb200_render_rays_analytic(const float* ray_origins, const float* ray_directions, float t_min, float t_max,
                                  const int* filter, int cap, const float* means, const float* log_scales,
                                  const float* rotations, const float* opacities, const float* features, const float* camera_pos,
                                  int n_rays, int N, int sh_dim, int sh_degree, float scaling_modifier, float sigma_threshold,
                                  int max_sections, float* out, void* stream);

/* ==== batched transient WITH occlusion (per-ray semantics of the reference: the first `cap` = 256 Gaussians in index order per ray,
 * src/ray_aabb.cu:10-61; shared transmittance with the early stop at 1e-4, src/volume_renderer.cu:80-137; attenuation + angular sum,
 * gaussian_model/cuda_autograd.py:302-314; x vp_y^2, nlos_helpers.py:275-276) for the scan points [sp0, sp0 + n_sp) of the tile-major grid
 * (row P of H = tile P/32, lane P%32).  Nothing of size [rays, T] is materialised.
 *   lists int32 [n_sp][A*A][cap+1] (nlosb200_occluded_list_ints): column 0 = count, then ASCENDING Gaussian ids = the reference's per-ray lists;
 *   t_samples [T] = the radii (uniform: t0 + k dt); c_deltaT = c * deltaT; mcut: samples with Mahalanobis distance > mcut are skipped;
 *   use_occlusion 0 renders the no-occlusion branch (:139-182) on the same capped lists; sh / sh_dim / sh_degree as nlosb200_transient_forward.
 * forward ADDS to H [n_tiles*32, T]; backward ADDS to dacc [N,16] (and dsh [N,16] when sh_degree > 0): the caller zeroes them.  A <= 64, T <= 2048. */
long long nlosb200_occluded_list_ints(int n_scan_points, int A, int cap);
int nlosb200_occluded_lists(const float* tiles, int n_tiles, int A, const float* records, int N, int cap, int sp0, int n_sp, int* lists,
                            void* stream);
int nlosb200_occluded_forward(const float* tiles, int n_tiles, int A, const float* records, int N, int cap, const int* lists, int sp0,
                              int n_sp, const float* t_samples, int T, float t0, float dt, float c_deltaT, float mcut, int use_occlusion,
                              const float* sh, int sh_dim, int sh_degree, float* H, void* stream);
int nlosb200_occluded_backward(const float* tiles, int n_tiles, int A, const float* records, int N, int cap, const int* lists, int sp0,
                               int n_sp, const float* t_samples, int T, float t0, float dt, float c_deltaT, float mcut, int use_occlusion,
                               const float* sh, int sh_dim, int sh_degree, const float* grad_H, float* dacc, float* dsh, void* stream);

/* ==== fused squared-error loss of the batched fitting step (nlos_helpers.py:280-346 compute_loss, per scan point in the reference; its
 * batched variant nlos_helpers.py:348-351 is an empty stub): loss += inv_numel * sum (H - target)^2 ; grad_H = 2 inv_numel (H - target)
 * on the tile-major layout of nlosb200_transient_forward.  n % 4 == 0, 16-byte aligned; loss is ONE double, accumulated (caller zeroes);
 * grad_H may be NULL. */
int nlosb200_mse_loss_grad(const float* H, const float* target, long long n, float inv_numel, float* grad_H, double* loss, void* stream);

/* ==== fused multi-tensor Adam step of the Gaussian parameter groups (GaussianModel.training_setup, gaussian_model/gaussian_model.py:223-249:
 * one torch.optim.Adam over mu / f_dc / f_rest / opacity / scaling / rotation with per-group lr, eps = 1e-15, no weight decay; stepped at
 * main.py:161,213).  params / grads / exp_avg / exp_avg_sq are HOST arrays of n_tensors (<= 8) DEVICE pointers, sizes / lrs HOST arrays;
 * step = 1, 2, ... selects the bias corrections.  Same operation order as torch.optim.Adam's single-tensor path. */
int nlosb200_adam_step(int n_tensors, float* const* params, const float* const* grads, float* const* exp_avg, float* const* exp_avg_sq,
                       const long long* sizes, const float* lrs, float beta1, float beta2, float eps, int step, void* stream);

/* ==== MCMC relocation helper (GaussianModel._update_params, gaussian_model/gaussian_model.py:471-483, calls compute_relocation_cuda -- a name
 * the reference never defines: the helper of the public 3DGS-MCMC code, not vendored; PARITY UNPINNED, restated from the publication, eq. 9):
 *   opacity_new = 1 - (1 - opacity_old)^(1/N) ;  scale_new = scale_old * opacity_old / sum_{i=1..N} sum_{k<i} C(i-1,k) (-1)^k opacity_new^(k+1) / sqrt(k+1)
 * counts N [P] int32 (clamped to [1, n_max-1]), binoms [n_max, n_max] float with C(n,k) at [n*n_max+k], n_max <= 64; activated values in
 * and out: opacity [P], scale [P,3]. */
int nlosb200_compute_relocation(const float* opacity_old, const float* scale_old, const int* counts, const float* binoms, int n_max, int P,
                                float* opacity_new, float* scale_new, void* stream);

/* ==== dense evaluation of the mixture at arbitrary points: density = sum_g pdf_g sigmoid(o_g), rho_density = sum_g pdf_g sigmoid(o_g) albedo_g over
 * ALL Gaussians (GaussianModel.estimate_rho_w_no_occlusion(out_separately=True), gaussian_model/gaussian_model.py:346-364, as used by the volume
 * evaluation nlos_helpers.py:40-69 gaussian2volume; pdf in the CUDA renderer's arithmetic, include/cuda_utils.cuh:124-151).  records from
 * nlosb200_preprocess (+ nlosb200_compat_albedo for a camera-dependent albedo).  Outputs [n_points], overwritten. */
int nlosb200_evaluate_points(const float* points, long long n_points, const float* records, int N, float* density, float* rho_density,
                             void* stream);
/* records for nlosb200_evaluate_points in the DENSE PyTorch semantics gaussian2volume actually runs: scale = exp(exp(_scaling) *
 * scaling_modifier) (gaussian_model.py:265), R (x - mu) instead of R^T (x - mu) (:281-282), no epsilon, degree-0 albedo.  Only mean,
 * sigmoid(opacity), M and albedo are filled. */
int nlosb200_preprocess_dense(const float* means, const float* log_scales, const float* rotations, const float* opacities,
                              const float* features, int N, int sh_dim, float scaling_modifier, float* records, void* stream);

/* ==== scale-initialisation helper: simple-knn's distCUDA2 (gaussian_model/gaussian_model.py:7-11,202-203; the submodule is pinned at
 * 60f461f in .SUBMODULES.json but absent from the tree).  out[i] = mean of the squared distances from point i to its 3 nearest OTHER
 * points (fewer if N < 4; the sum is still divided by 3).  Exact k-NN on a uniform hash grid, no host synchronisation.
 * workspace: nlosb200_knn_workspace_bytes(N) bytes of device scratch, 16-byte aligned. */
long long nlosb200_knn_workspace_bytes(int N);
int nlosb200_knn_dist2(const float* points, int N, float* out, void* workspace, long long workspace_bytes, void* stream);

#ifdef __cplusplus
}
#endif
#endif
