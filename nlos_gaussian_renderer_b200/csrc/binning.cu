// binning.cu -- shell-intersection culling / binning of Gaussians per (scan tile x time-bin range).
//
// Replaces the reference's brute-force per-ray AABB filter as the culling structure of the batched
// renderer (/root/reference/submodules/cuda_renderer/src/ray_aabb.cu:10-102 is O(N_rays * N) and caps
// every ray at 256 Gaussians).  In confocal NLOS every scan point's angular lattice covers the whole
// hidden volume, so the only effective cull is radial: Gaussian g belongs to cell (tile, range) iff
// its conservative shell  [ sqrt(min_p |mu-o_p|^2) - rad , sqrt(max_p |mu-o_p|^2) + rad ]  seen from
// the tile's scan points intersects the range's bins.  The definition (and the bit-exact operation
// order) is oracle/nlos_oracle.c:orc_shell_intervals / orc_shell_lists; lists are ascending in g.
//
// Passes (all stream ordered, no host sync inside):
//   count : per (tile, chunk of 1024 Gaussians) histogram of range membership
//   scan  : per cell exclusive scan over chunks; then over cells (int64 offsets), and the render
//           work decomposition (sub-items of <= slice_len list entries, CTA items of warps_per_block)
//   fill  : ordered (ballot-ranked) scatter of Gaussian ids into the cell lists
#include "common.cuh"
#include "../../include/nlosb200.h"

using namespace nlos;

#define BIN_CHUNK 1024
#define BIN_THREADS 256
#define BIN_GPT (BIN_CHUNK / BIN_THREADS)
#define BIN_MAX_RANGES 64

struct TilePoints {
    float ox[NLOS_TILE], oy[NLOS_TILE], oz[NLOS_TILE];
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
    }
}

// oracle/nlos_oracle.c:orc_shell_intervals, same single-rounding operation order
__device__ __forceinline__ void shell_interval(const TilePoints& tp, float4 sh, float t0, float inv_dt, int T, int bpr, int n_ranges,
                                               int& lo, int& hi)
{
    float d2min = INFINITY, d2max = 0.0f;
#pragma unroll 8
    for (int l = 0; l < NLOS_TILE; ++l) {
        if (!tp.valid[l]) continue;
        float dx = __fsub_rn(sh.x, tp.ox[l]), dy = __fsub_rn(sh.y, tp.oy[l]), dz = __fsub_rn(sh.z, tp.oz[l]);
        float d2 = __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
        d2min = fminf(d2min, d2);
        d2max = fmaxf(d2max, d2);
    }
    float rlo = __fsub_rn(__fsqrt_rn(d2min), sh.w), rhi = __fadd_rn(__fsqrt_rn(d2max), sh.w);
    float klo_f = floorf(__fadd_rn(__fmul_rn(__fsub_rn(rlo, t0), inv_dt), 0.5f));
    float khi_f = floorf(__fadd_rn(__fmul_rn(__fsub_rn(rhi, t0), inv_dt), 0.5f));
    int klo = klo_f < 0.0f ? 0 : (klo_f > (float)(T - 1) ? T : (int)klo_f);
    int khi = khi_f < 0.0f ? -1 : (khi_f > (float)(T - 1) ? T - 1 : (int)khi_f);
    if (klo > khi) {
        lo = 1;
        hi = 0;
    } else {
        lo = klo / bpr;
        hi = khi / bpr;
        if (hi >= n_ranges) hi = n_ranges - 1;
    }
}

__global__ void __launch_bounds__(BIN_THREADS) shell_count_kernel(const float* __restrict__ tiles, int A, const float4* __restrict__ shell, int N,
                                                                  float t0, float inv_dt, int T, int bpr, int n_ranges, int n_chunks,
                                                                  int* __restrict__ chunk_counts)
{
    __shared__ TilePoints tp;
    __shared__ int cnt[BIN_MAX_RANGES];
    int tile = blockIdx.y, chunk = blockIdx.x;
    load_tile_points(tp, tiles, tile, A);
    if (threadIdx.x < BIN_MAX_RANGES) cnt[threadIdx.x] = 0;
    __syncthreads();
#pragma unroll
    for (int j = 0; j < BIN_GPT; ++j) {
        int g = chunk * BIN_CHUNK + j * BIN_THREADS + threadIdx.x;
        if (g < N) {
            int lo, hi;
            shell_interval(tp, shell[g], t0, inv_dt, T, bpr, n_ranges, lo, hi);
            for (int r = lo; r <= hi; ++r) atomicAdd(&cnt[r], 1);
        }
    }
    __syncthreads();
    if (threadIdx.x < n_ranges) chunk_counts[((size_t)tile * n_ranges + threadIdx.x) * n_chunks + chunk] = cnt[threadIdx.x];
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

// single CTA: cell offsets (int64), per-tile sub-item prefix, CTA-item prefix, header
//   header[0] = total list entries, header[1] = number of CTA work items, header[2] = overflow flag (set by fill)
__global__ void __launch_bounds__(1024) scan_cells_kernel(const int* __restrict__ cell_counts, int n_tiles, int n_ranges, int slice_len, int wpb,
                                                          long long* __restrict__ cell_offsets, int* __restrict__ cell_sub_off,
                                                          int* __restrict__ tile_items, long long* __restrict__ header)
{
    __shared__ long long part[1024];
    __shared__ int parti[1024];
    int n_cells = n_tiles * n_ranges;
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
        int seg = (n_tiles + 1023) / 1024;
        int b = threadIdx.x * seg, e = min(b + seg, n_tiles);
        int s = 0;
        for (int t = b; t < e; ++t) {
            int nsub = 0;
            for (int r = 0; r < n_ranges; ++r) {
                cell_sub_off[t * n_ranges + r] = nsub;
                nsub += (cell_counts[t * n_ranges + r] + slice_len - 1) / slice_len;
            }
            s += (nsub + wpb - 1) / wpb;
        }
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
        for (int t = b; t < e; ++t) {
            tile_items[t] = run;
            int nsub = 0;
            for (int r = 0; r < n_ranges; ++r) nsub += (cell_counts[t * n_ranges + r] + slice_len - 1) / slice_len;
            run += (nsub + wpb - 1) / wpb;
        }
    }
}

__global__ void __launch_bounds__(BIN_THREADS) shell_fill_kernel(const float* __restrict__ tiles, int A, const float4* __restrict__ shell, int N,
                                                                 float t0, float inv_dt, int T, int bpr, int n_ranges, int n_chunks,
                                                                 const int* __restrict__ chunk_off, const long long* __restrict__ cell_offsets,
                                                                 int* __restrict__ list, long long capacity, long long* __restrict__ header)
{
    __shared__ TilePoints tp;
    __shared__ int wc[BIN_GPT][BIN_THREADS / 32][BIN_MAX_RANGES];
    int tile = blockIdx.y, chunk = blockIdx.x;
    int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    load_tile_points(tp, tiles, tile, A);
    __syncthreads();
    int lo[BIN_GPT], hi[BIN_GPT];
#pragma unroll
    for (int j = 0; j < BIN_GPT; ++j) {
        int g = chunk * BIN_CHUNK + j * BIN_THREADS + threadIdx.x;
        lo[j] = 1;
        hi[j] = 0;
        if (g < N) shell_interval(tp, shell[g], t0, inv_dt, T, bpr, n_ranges, lo[j], hi[j]);
    }
#pragma unroll
    for (int j = 0; j < BIN_GPT; ++j)
        for (int r = 0; r < n_ranges; ++r) {
            unsigned b = __ballot_sync(0xffffffffu, lo[j] <= r && r <= hi[j]);
            if (lane == 0) wc[j][warp][r] = __popc(b);
        }
    __syncthreads();
    if (threadIdx.x < n_ranges) {
        int r = threadIdx.x, run = 0;
        for (int j = 0; j < BIN_GPT; ++j)
            for (int w = 0; w < BIN_THREADS / 32; ++w) { int v = wc[j][w][r]; wc[j][w][r] = run; run += v; }
    }
    __syncthreads();
    unsigned lt = (1u << lane) - 1u;
#pragma unroll
    for (int j = 0; j < BIN_GPT; ++j) {
        int g = chunk * BIN_CHUNK + j * BIN_THREADS + threadIdx.x;
        for (int r = 0; r < n_ranges; ++r) {
            bool m = lo[j] <= r && r <= hi[j];
            unsigned b = __ballot_sync(0xffffffffu, m);
            if (m) {
                size_t cell = (size_t)tile * n_ranges + r;
                long long pos = cell_offsets[cell] + chunk_off[cell * n_chunks + chunk] + wc[j][warp][r] + __popc(b & lt);
                if (pos < capacity) list[pos] = g;
                else header[2] = 1;
            }
        }
    }
}

extern "C" int nlosb200_bin_num_chunks(int N) { return (N + BIN_CHUNK - 1) / BIN_CHUNK; }

extern "C" int nlosb200_bin_shells_plan(const float* tiles, int n_tiles, int A, const float* shell, int N, float t0, float dt, int T,
                                        int bins_per_range, int slice_len, int warps_per_block, int* chunk_counts, int* cell_counts,
                                        long long* cell_offsets, int* cell_sub_off, int* tile_items, long long* header, void* stream)
{
    NLOS_REQUIRE(n_tiles > 0 && N > 0 && T > 0 && bins_per_range > 0 && slice_len > 0 && warps_per_block > 0, "bin_shells_plan: bad sizes");
    int n_ranges = (T + bins_per_range - 1) / bins_per_range;
    NLOS_REQUIRE(n_ranges <= BIN_MAX_RANGES, "bin_shells_plan: %d ranges > %d (raise bins_per_range)", n_ranges, BIN_MAX_RANGES);
    NLOS_REQUIRE(n_tiles <= 65535, "bin_shells_plan: n_tiles=%d > 65535 (shard the scan grid)", n_tiles);
    NLOS_REQUIRE(tiles && shell && chunk_counts && cell_counts && cell_offsets && cell_sub_off && tile_items && header,
                 "bin_shells_plan: null pointer");
    int n_chunks = (N + BIN_CHUNK - 1) / BIN_CHUNK;
    float inv_dt = 1.0f / dt;
    cudaStream_t s = (cudaStream_t)stream;
    shell_count_kernel<<<dim3(n_chunks, n_tiles), BIN_THREADS, 0, s>>>(tiles, A, reinterpret_cast<const float4*>(shell), N, t0, inv_dt, T,
                                                                       bins_per_range, n_ranges, n_chunks, chunk_counts);
    NLOS_CHECK_LAUNCH("shell_count_kernel");
    scan_chunks_kernel<<<n_tiles * n_ranges, 256, 0, s>>>(chunk_counts, n_chunks, cell_counts);
    NLOS_CHECK_LAUNCH("scan_chunks_kernel");
    scan_cells_kernel<<<1, 1024, 0, s>>>(cell_counts, n_tiles, n_ranges, slice_len, warps_per_block, cell_offsets, cell_sub_off, tile_items,
                                         header);
    NLOS_CHECK_LAUNCH("scan_cells_kernel");
    return 0;
}

extern "C" int nlosb200_bin_shells_fill(const float* tiles, int n_tiles, int A, const float* shell, int N, float t0, float dt, int T,
                                        int bins_per_range, const int* chunk_counts_scanned, const long long* cell_offsets, int* list,
                                        long long list_capacity, long long* header, void* stream)
{
    NLOS_REQUIRE(n_tiles > 0 && N > 0 && T > 0 && bins_per_range > 0, "bin_shells_fill: bad sizes");
    int n_ranges = (T + bins_per_range - 1) / bins_per_range;
    NLOS_REQUIRE(n_ranges <= BIN_MAX_RANGES, "bin_shells_fill: too many ranges");
    NLOS_REQUIRE(list || list_capacity == 0, "bin_shells_fill: null list");
    int n_chunks = (N + BIN_CHUNK - 1) / BIN_CHUNK;
    float inv_dt = 1.0f / dt;
    shell_fill_kernel<<<dim3(n_chunks, n_tiles), BIN_THREADS, 0, (cudaStream_t)stream>>>(
        tiles, A, reinterpret_cast<const float4*>(shell), N, t0, inv_dt, T, bins_per_range, n_ranges, n_chunks, chunk_counts_scanned,
        cell_offsets, list, list_capacity, header);
    NLOS_CHECK_LAUNCH("shell_fill_kernel");
    return 0;
}
