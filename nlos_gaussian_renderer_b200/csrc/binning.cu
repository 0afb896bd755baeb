// binning.cu -- window binning of Gaussians per scan tile for the batched transient renderer.
//
// Replaces the reference's brute-force per-ray AABB filter as the culling structure of the batched
// renderer (/root/reference/submodules/cuda_renderer/src/ray_aabb.cu:10-102 is O(N_rays * N) and caps
// every ray at 256 Gaussians).  In confocal NLOS every scan point's angular lattice covers the whole
// hidden volume, so the only effective cull is radial: per (tile, Gaussian) the conservative shell
//   [ sqrt(min_p |mu-o_p|^2) - rad , sqrt(max_p |mu-o_p|^2) + rad ]   (p over the tile's scan points)
// is converted to an inclusive bin interval [klo, khi].  The render kernel accumulates into a sliding window of
// `window` bins that advances in steps of `bucket` bins, so a Gaussian is listed ONCE per tile (entry e = 0) unless
// its interval is longer than L = window - bucket, in which case entry e covers [klo + e L, klo + (e+1) L - 1].
// Entry (g, e) lives in cell (tile, (klo + e L) / bucket); per cell the entries are ascending in g; the list value
// is g | e << 26.  Definition and bit-exact operation order: oracle/nlos_oracle.c:orc_shell_bins and
// oracle/__init__.py:window_lists.
//
// Entry classes (n_classes = 2): the cells of a tile are class-major, cell = tile * 2 n_buckets + cls * n_buckets + bucket.  An entry is
// class 0 ("narrow": rendered by transient_narrow.cu) when the Gaussian is listed once in the tile (interval <= L) and its 3-sigma AABB
// spans at most `narrow_rays` lattice pitches of the tile:  (6 / mcut) rad <= narrow_rays * pitch * sqrt(min_p |mu - o_p|^2) , with
// rad = mcut max(scale) from the shell and pitch = the smallest angular step (dtheta, dphi) of the tile's scan points; otherwise class 1.
// Same single-rounding operation order in oracle/__init__.py:entry_classes.  A second, independent class bit (time_wide = 1): the Gaussian's
// smallest scale is at least 1.3 bins (flag = sign bit of the shell radius, set by nlosb200_preprocess), it is listed once in the tile and
// its interval keeps 3 bins from both ends of the time axis -- every ray of such an entry is a ONE-node ray of the analytic-exact mode (transient.cu, one-pass plan).
// Class index = footprint bit + n_fp * time bit, n_classes = n_fp * n_tw in {1, 2, 4}.
//
// Passes (all stream ordered, no host sync inside):
//   count : per (tile, chunk of 1024 Gaussians) histogram of cell membership
//   scan  : per cell exclusive scan over chunks; then over cells (int64 offsets), and the render work
//           decomposition (per tile: slices of slice_len entries, CTA items of warps_per_block slices)
//   fill  : ordered (ballot-ranked) scatter of entries into the cell lists
#include "common.cuh"
#include "../../include/nlosb200.h"

using namespace nlos;

#define BIN_CHUNK 1024
#define BIN_THREADS 256
#define BIN_GPT (BIN_CHUNK / BIN_THREADS)
#define BIN_MAX_RANGES 64
#define BIN_MAX_CLASSES 4

struct TilePoints {
    float ox[NLOS_TILE], oy[NLOS_TILE], oz[NLOS_TILE];
    float step[NLOS_TILE];     // min(dtheta, dphi) of the scan point's lattice (0: no usable lattice)
    int valid[NLOS_TILE];
};

__device__ __forceinline__ void load_tile_points(TilePoints& tp, const float* __restrict__ tiles, int tile, int A)
{
    size_t tile_floats = (size_t)(NLOS_NTAB * A + NLOS_PINFO_ROWS) * NLOS_TILE;
    const float* pi = tiles + tile * tile_floats + (size_t)NLOS_NTAB * A * NLOS_TILE;
    if (threadIdx.x < NLOS_TILE) {
        int l = threadIdx.x;
        tp.ox[l] = pi[PI_OX * NLOS_TILE + l];
        tp.oy[l] = pi[PI_OY * NLOS_TILE + l];
        tp.oz[l] = pi[PI_OZ * NLOS_TILE + l];
        tp.valid[l] = __float_as_int(pi[PI_FLAGS * NLOS_TILE + l]) & 1;
        tp.step[l] = (__float_as_int(pi[PI_FLAGS * NLOS_TILE + l]) & 2) ? fminf(pi[PI_DTH * NLOS_TILE + l], pi[PI_DPH * NLOS_TILE + l]) : 0.f;
    }
}

// oracle/nlos_oracle.c:orc_shell_bins, same single-rounding operation order.  klo > khi = empty.
// cls: 0 = narrow entry, 1 = wide (header comment); narrow_k = narrow_rays (0: a single class, everything is class 0), rad_k = 6 / mcut.
__device__ __forceinline__ void shell_bins(const TilePoints& tp, float4 sh, float t0, float inv_dt, int T, int L, float narrow_k, float rad_k,
                                           int& klo, int& khi, int& cls)
{
    const bool tw_flag = sh.w < 0.f;           // time-wide Gaussian (preprocess); rad_k < 0 says the flag takes part in the class index
    sh.w = fabsf(sh.w);
    float d2min = INFINITY, d2max = 0.0f, pitch = INFINITY;
#pragma unroll 8
    for (int l = 0; l < NLOS_TILE; ++l) {
        if (!tp.valid[l]) continue;
        float dx = __fsub_rn(sh.x, tp.ox[l]), dy = __fsub_rn(sh.y, tp.oy[l]), dz = __fsub_rn(sh.z, tp.oz[l]);
        float d2 = __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
        d2min = fminf(d2min, d2);
        d2max = fmaxf(d2max, d2);
        pitch = fminf(pitch, tp.step[l]);
    }
    float rlo = __fsub_rn(__fsqrt_rn(d2min), sh.w), rhi = __fadd_rn(__fsqrt_rn(d2max), sh.w);
    float klo_f = floorf(__fadd_rn(__fmul_rn(__fsub_rn(rlo, t0), inv_dt), 0.5f));
    float khi_f = floorf(__fadd_rn(__fmul_rn(__fsub_rn(rhi, t0), inv_dt), 0.5f));
    klo = klo_f < 0.0f ? 0 : (klo_f > (float)(T - 1) ? T : (int)klo_f);
    khi = khi_f < 0.0f ? -1 : (khi_f > (float)(T - 1) ? T - 1 : (int)khi_f);
    cls = 0;
    int n_fp = 1;
    if (narrow_k > 0.f) {
        bool narrow = khi - klo + 1 <= L && __fmul_rn(sh.w, fabsf(rad_k)) <= __fmul_rn(__fmul_rn(narrow_k, pitch), __fsqrt_rn(d2min));
        cls = narrow ? 0 : 1;
        n_fp = 2;
    }
    if (rad_k < 0.f && tw_flag && klo >= 3 && khi <= T - 4 && khi - klo + 1 <= L) cls += n_fp;
}

// sub-interval index e of the entry of [klo, khi] that lives in bucket r, or -1  (L > bucket => at most one)
__device__ __forceinline__ int entry_in_bucket(int klo, int khi, int r, int L, int bucket)
{
    if (klo > khi) return -1;
    int top = r * bucket + bucket - 1 - klo;          // last bin of the bucket, relative to klo
    if (top < 0) return -1;
    int e = top / L;
    int s = klo + e * L;
    return (s <= khi && s >= r * bucket) ? e : -1;
}

__global__ void __launch_bounds__(BIN_THREADS) shell_count_kernel(const float* __restrict__ tiles, int A, const float4* __restrict__ shell, int N,
                                                                  float t0, float inv_dt, int T, int L, int bucket, int n_buckets, int n_chunks,
                                                                  int n_classes, float narrow_k, float rad_k, int* __restrict__ chunk_counts)
{
    __shared__ TilePoints tp;
    __shared__ int cnt[BIN_MAX_CLASSES * BIN_MAX_RANGES];
    int tile = blockIdx.y, chunk = blockIdx.x;
    load_tile_points(tp, tiles, tile, A);
    if (threadIdx.x < BIN_MAX_CLASSES * BIN_MAX_RANGES) cnt[threadIdx.x] = 0;
    __syncthreads();
#pragma unroll
    for (int j = 0; j < BIN_GPT; ++j) {
        int g = chunk * BIN_CHUNK + j * BIN_THREADS + threadIdx.x;
        if (g < N) {
            int klo, khi, cls;
            shell_bins(tp, shell[g], t0, inv_dt, T, L, narrow_k, rad_k, klo, khi, cls);
            for (int s = klo; s <= khi; s += L) atomicAdd(&cnt[cls * n_buckets + s / bucket], 1);
        }
    }
    __syncthreads();
    const int cpt = n_classes * n_buckets;
    if (threadIdx.x < cpt) chunk_counts[((size_t)tile * cpt + threadIdx.x) * n_chunks + chunk] = cnt[threadIdx.x];
}

// per cell: exclusive scan over chunks (in place) and the cell total
__global__ void __launch_bounds__(256) scan_chunks_kernel(int* __restrict__ chunk_counts, int n_chunks, int* __restrict__ cell_counts)
{
    __shared__ int part[256];
    int cell = blockIdx.x;
    int* row = chunk_counts + (size_t)cell * n_chunks;
    int seg = (n_chunks + 255) / 256;
    int b = threadIdx.x * seg, e = min(b + seg, n_chunks);
    int s = 0;
    for (int i = b; i < e; ++i) s += row[i];
    part[threadIdx.x] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        int run = 0;
        for (int i = 0; i < 256; ++i) { int v = part[i]; part[i] = run; run += v; }
        cell_counts[cell] = run;
    }
    __syncthreads();
    int run = part[threadIdx.x];
    for (int i = b; i < e; ++i) { int v = row[i]; row[i] = run; run += v; }
}

// single CTA: cell offsets (int64), CTA-item prefix per tile, header
//   header[0] = total list entries, header[1] = number of CTA work items, header[2] = overflow flag (set by fill)
__global__ void __launch_bounds__(1024) scan_cells_kernel(const int* __restrict__ cell_counts, int n_tiles, int n_buckets /* cells per tile */, int slice_len, int wpb,
                                                          long long* __restrict__ cell_offsets, int* __restrict__ tile_items,
                                                          long long* __restrict__ header)
{
    __shared__ long long part[1024];
    __shared__ int parti[1024];
    int n_cells = n_tiles * n_buckets;
    {
        int seg = (n_cells + 1023) / 1024;
        int b = threadIdx.x * seg, e = min(b + seg, n_cells);
        long long s = 0;
        for (int i = b; i < e; ++i) s += cell_counts[i];
        part[threadIdx.x] = s;
        __syncthreads();
        if (threadIdx.x == 0) {
            long long run = 0;
            for (int i = 0; i < 1024; ++i) { long long v = part[i]; part[i] = run; run += v; }
            cell_offsets[n_cells] = run;
            header[0] = run;
            header[2] = 0;
        }
        __syncthreads();
        long long run = part[threadIdx.x];
        for (int i = b; i < e; ++i) { cell_offsets[i] = run; run += cell_counts[i]; }
    }
    {
        // a tile's entries are contiguous and ordered by bucket; slices of slice_len entries, wpb slices per CTA item
        int seg = (n_tiles + 1023) / 1024;
        int b = threadIdx.x * seg, e = min(b + seg, n_tiles);
        auto items_of = [&](int t) {
            long long tot = 0;
            for (int r = 0; r < n_buckets; ++r) tot += cell_counts[t * n_buckets + r];
            long long nsub = (tot + slice_len - 1) / slice_len;
            return (int)((nsub + wpb - 1) / wpb);
        };
        int s = 0;
        for (int t = b; t < e; ++t) s += items_of(t);
        parti[threadIdx.x] = s;
        __syncthreads();
        if (threadIdx.x == 0) {
            int run = 0;
            for (int i = 0; i < 1024; ++i) { int v = parti[i]; parti[i] = run; run += v; }
            tile_items[n_tiles] = run;
            header[1] = run;
        }
        __syncthreads();
        int run = parti[threadIdx.x];
        for (int t = b; t < e; ++t) { tile_items[t] = run; run += items_of(t); }
    }
}

__global__ void __launch_bounds__(BIN_THREADS) shell_fill_kernel(const float* __restrict__ tiles, int A, const float4* __restrict__ shell, int N,
                                                                 float t0, float inv_dt, int T, int L, int bucket, int n_buckets, int n_chunks,
                                                                 int n_classes, float narrow_k, float rad_k,
                                                                 const int* __restrict__ chunk_off, const long long* __restrict__ cell_offsets,
                                                                 unsigned* __restrict__ list, unsigned* __restrict__ list_bins, long long capacity,
                                                                 long long* __restrict__ header)
{
    // ordered fill without a loop over the cells: every entry sets the bit of its Gaussian's index inside the chunk in the bitmap of its cell
    // (shared memory), a prefix over the bitmap words gives each entry its rank among the chunk's entries of the same cell in ascending
    // Gaussian index -- the order oracle.window_lists defines
    __shared__ TilePoints tp;
    extern __shared__ __align__(16) unsigned char fill_smem[];        // bitmaps [cpt][32] words + word prefixes [cpt][32] shorts
    int tile = blockIdx.y, chunk = blockIdx.x;
    const int cpt = n_classes * n_buckets;
    unsigned (*bm)[BIN_CHUNK / 32] = reinterpret_cast<unsigned (*)[BIN_CHUNK / 32]>(fill_smem);
    unsigned short (*pre)[BIN_CHUNK / 32] = reinterpret_cast<unsigned short (*)[BIN_CHUNK / 32]>(fill_smem + (size_t)cpt * (BIN_CHUNK / 32) * 4);
    load_tile_points(tp, tiles, tile, A);
    for (int i = threadIdx.x; i < cpt * (BIN_CHUNK / 32); i += BIN_THREADS) (&bm[0][0])[i] = 0u;
    __syncthreads();
    int lo[BIN_GPT], hi[BIN_GPT], cl[BIN_GPT];
#pragma unroll
    for (int j = 0; j < BIN_GPT; ++j) {
        int g = chunk * BIN_CHUNK + j * BIN_THREADS + threadIdx.x;
        lo[j] = 1;
        hi[j] = 0;
        cl[j] = 0;
        if (g < N) shell_bins(tp, shell[g], t0, inv_dt, T, L, narrow_k, rad_k, lo[j], hi[j], cl[j]);
        const int li = j * BIN_THREADS + threadIdx.x;                     // index inside the chunk, ascending with g
        for (int s = lo[j]; s <= hi[j]; s += L) atomicOr(&bm[cl[j] * n_buckets + s / bucket][li >> 5], 1u << (li & 31));
    }
    __syncthreads();
    if (threadIdx.x < cpt) {
        unsigned run = 0;
        for (int w = 0; w < BIN_CHUNK / 32; ++w) { pre[threadIdx.x][w] = (unsigned short)run; run += __popc(bm[threadIdx.x][w]); }
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < BIN_GPT; ++j) {
        int g = chunk * BIN_CHUNK + j * BIN_THREADS + threadIdx.x;
        const int li = j * BIN_THREADS + threadIdx.x;
        int e = 0;
        for (int s = lo[j]; s <= hi[j]; s += L, ++e) {
            const int c = cl[j] * n_buckets + s / bucket;
            const size_t cell = (size_t)tile * cpt + c;
            const long long pos = cell_offsets[cell] + chunk_off[cell * n_chunks + chunk] + pre[c][li >> 5] +
                                  __popc(bm[c][li >> 5] & ((1u << (li & 31)) - 1u));
            if (pos < capacity) {
                list[pos] = (unsigned)g | ((unsigned)e << NLOS_ENTRY_SHIFT);
                if (list_bins) list_bins[pos] = (unsigned)lo[j] | ((unsigned)hi[j] << 16);      // the Gaussian's tile-level shell [klo, khi]
            } else header[2] = 1;
        }
    }
}

extern "C" int nlosb200_bin_num_chunks(int N) { return (N + BIN_CHUNK - 1) / BIN_CHUNK; }

static int check_window(const char* who, int T, int window, int bucket, int N)
{
    NLOS_REQUIRE(window >= 32 && window % 8 == 0, "%s: window=%d must be a multiple of 8, >= 32", who, window);
    NLOS_REQUIRE(bucket >= 1 && 2 * bucket <= window, "%s: bucket=%d must be in [1, window/2]", who, bucket);
    NLOS_REQUIRE((T + bucket - 1) / bucket <= BIN_MAX_RANGES, "%s: %d buckets > %d (raise the bucket size)", who, (T + bucket - 1) / bucket,
                 BIN_MAX_RANGES);
    NLOS_REQUIRE(N < (1 << NLOS_ENTRY_SHIFT), "%s: N=%d >= 2^%d", who, N, NLOS_ENTRY_SHIFT);
    NLOS_REQUIRE((T + (window - bucket) - 1) / (window - bucket) < (1 << (32 - NLOS_ENTRY_SHIFT)), "%s: too many sub-intervals", who);
    return 0;
}

// rad_k: |rad_k| = 6 / mcut (footprint class), sign = the time-wide flag of the shell takes part in the class index
static int check_classes(const char* who, int n_classes, float narrow_rays, float mcut, int time_wide, float& rad_k)
{
    NLOS_REQUIRE(narrow_rays >= 0.f && (time_wide == 0 || time_wide == 1), "%s: bad class arguments", who);
    const int want = (narrow_rays > 0.f ? 2 : 1) * (time_wide ? 2 : 1);
    NLOS_REQUIRE(n_classes == want, "%s: n_classes=%d but narrow_rays=%g, time_wide=%d give %d", who, n_classes, narrow_rays, time_wide, want);
    NLOS_REQUIRE(mcut > 0.f, "%s: mcut must be > 0", who);
    rad_k = (time_wide ? -1.0f : 1.0f) * 6.0f / mcut;
    return 0;
}

extern "C" int nlosb200_bin_window_plan(const float* tiles, int n_tiles, int A, const float* shell, int N, float t0, float dt, int T,
                                        int window, int bucket, int n_classes, float narrow_rays, float mcut, int time_wide, int slice_len, int warps_per_block,
                                        int* chunk_counts, int* cell_counts, long long* cell_offsets, int* tile_items, long long* header,
                                        void* stream)
{
    NLOS_REQUIRE(n_tiles > 0 && N > 0 && T > 0 && slice_len > 0 && warps_per_block > 0, "bin_window_plan: bad sizes");
    int rc = check_window("bin_window_plan", T, window, bucket, N);
    if (rc) return rc;
    float rad_k;
    rc = check_classes("bin_window_plan", n_classes, narrow_rays, mcut, time_wide, rad_k);
    if (rc) return rc;
    int n_buckets = (T + bucket - 1) / bucket;
    NLOS_REQUIRE(n_tiles <= 65535, "bin_window_plan: n_tiles=%d > 65535 (shard the scan grid)", n_tiles);
    NLOS_REQUIRE(tiles && shell && chunk_counts && cell_counts && cell_offsets && tile_items && header, "bin_window_plan: null pointer");
    int n_chunks = (N + BIN_CHUNK - 1) / BIN_CHUNK;
    float inv_dt = 1.0f / dt;
    cudaStream_t s = (cudaStream_t)stream;
    shell_count_kernel<<<dim3(n_chunks, n_tiles), BIN_THREADS, 0, s>>>(tiles, A, reinterpret_cast<const float4*>(shell), N, t0, inv_dt, T,
                                                                       window - bucket, bucket, n_buckets, n_chunks, n_classes, narrow_rays, rad_k,
                                                                       chunk_counts);
    NLOS_CHECK_LAUNCH("shell_count_kernel");
    scan_chunks_kernel<<<n_tiles * n_buckets * n_classes, 256, 0, s>>>(chunk_counts, n_chunks, cell_counts);
    NLOS_CHECK_LAUNCH("scan_chunks_kernel");
    scan_cells_kernel<<<1, 1024, 0, s>>>(cell_counts, n_tiles, n_buckets * n_classes, slice_len, warps_per_block, cell_offsets, tile_items, header);
    NLOS_CHECK_LAUNCH("scan_cells_kernel");
    return 0;
}

extern "C" int nlosb200_bin_window_fill(const float* tiles, int n_tiles, int A, const float* shell, int N, float t0, float dt, int T,
                                        int window, int bucket, int n_classes, float narrow_rays, float mcut, int time_wide, const int* chunk_counts_scanned,
                                        const long long* cell_offsets, int* list, int* list_bins, long long list_capacity, long long* header,
                                        void* stream)
{
    NLOS_REQUIRE(n_tiles > 0 && N > 0 && T > 0, "bin_window_fill: bad sizes");
    int rc = check_window("bin_window_fill", T, window, bucket, N);
    if (rc) return rc;
    float rad_k;
    rc = check_classes("bin_window_fill", n_classes, narrow_rays, mcut, time_wide, rad_k);
    if (rc) return rc;
    int n_buckets = (T + bucket - 1) / bucket;
    NLOS_REQUIRE(list || list_capacity == 0, "bin_window_fill: null list");
    NLOS_REQUIRE(!list_bins || T <= 65535, "bin_window_fill: list_bins packs two 16-bit bins (T=%d)", T);
    int n_chunks = (N + BIN_CHUNK - 1) / BIN_CHUNK;
    float inv_dt = 1.0f / dt;
    const size_t fill_smem = (size_t)n_classes * n_buckets * (BIN_CHUNK / 32) * 6;
    cudaError_t ea = cudaFuncSetAttribute(shell_fill_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)fill_smem);
    NLOS_REQUIRE(ea == cudaSuccess, "bin_window_fill: cudaFuncSetAttribute: %s", cudaGetErrorString(ea));
    shell_fill_kernel<<<dim3(n_chunks, n_tiles), BIN_THREADS, fill_smem, (cudaStream_t)stream>>>(
        tiles, A, reinterpret_cast<const float4*>(shell), N, t0, inv_dt, T, window - bucket, bucket, n_buckets, n_chunks, n_classes, narrow_rays,
        rad_k, chunk_counts_scanned, cell_offsets, reinterpret_cast<unsigned*>(list), reinterpret_cast<unsigned*>(list_bins), list_capacity, header);
    NLOS_CHECK_LAUNCH("shell_fill_kernel");
    return 0;
}
