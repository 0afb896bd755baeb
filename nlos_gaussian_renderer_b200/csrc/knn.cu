// knn.cu -- mean squared distance to the 3 nearest OTHER points: the scale initialisation helper the reference takes
// from the simple-knn submodule (distCUDA2; /root/reference/gaussian_model/gaussian_model.py:7-11,202-203; submodule
// pinned at 60f461f in .SUBMODULES.json but absent from the tree, so parity is pinned to the brute-force restatement
// oracle/nlos_oracle.c:orc_knn_dist2 -- same squared-distance expression, same summation order: bit-exact).
//
// B200 design: exact k-NN on a uniform hash grid instead of simple-knn's Morton boxes -- O(N) work, all HBM-bound passes:
//   bbox (warp reductions + ordered-int atomics) -> cell histogram -> exclusive scan -> counting-sort scatter (float4 records:
//   coalesced 16-B loads in the query) -> one thread per point walks Chebyshev rings of cells until the third-best distance
//   is provably final.  No host synchronisation: the grid geometry is derived on the device from the bounding box.
#include "common.cuh"
#include "../../include/nlosb200.h"

using namespace nlos;

#define KNN_THREADS 256
#define KNN_SCAN_BLOCK 1024

struct KnnGrid {
    float lox, loy, loz, h, inv_h;
    int gx, gy, gz;
};

__device__ __forceinline__ unsigned f2ord(float f)
{
    unsigned u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float ord2f(unsigned u)
{
    return __uint_as_float((u & 0x80000000u) ? (u & 0x7fffffffu) : ~u);
}

// bbox[0..2] = ordered min, bbox[3..5] = ordered max
__global__ void __launch_bounds__(KNN_THREADS) knn_bbox_kernel(const float* __restrict__ pts, int N, unsigned* __restrict__ bbox)
{
    float mn[3] = {INFINITY, INFINITY, INFINITY}, mx[3] = {-INFINITY, -INFINITY, -INFINITY};
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < N; i += gridDim.x * blockDim.x)
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            const float v = pts[3 * (size_t)i + a];
            mn[a] = fminf(mn[a], v);
            mx[a] = fmaxf(mx[a], v);
        }
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        const unsigned lo = __reduce_min_sync(0xffffffffu, f2ord(mn[a])), hi = __reduce_max_sync(0xffffffffu, f2ord(mx[a]));
        if ((threadIdx.x & 31) == 0) {
            atomicMin(bbox + a, lo);
            atomicMax(bbox + 3 + a, hi);
        }
    }
}

// grid geometry from the bounding box: g_long cells along the longest axis (same code in every kernel -> same cells)
__device__ __forceinline__ KnnGrid knn_grid(const unsigned* __restrict__ bbox, int g_long)
{
    KnnGrid G;
    G.lox = ord2f(bbox[0]); G.loy = ord2f(bbox[1]); G.loz = ord2f(bbox[2]);
    const float ex = ord2f(bbox[3]) - G.lox, ey = ord2f(bbox[4]) - G.loy, ez = ord2f(bbox[5]) - G.loz;
    const float longest = fmaxf(fmaxf(ex, ey), fmaxf(ez, 1e-30f));
    G.h = longest / (float)g_long * 1.0001f;
    G.inv_h = 1.0f / G.h;
    G.gx = min(g_long, (int)(ex * G.inv_h) + 1);
    G.gy = min(g_long, (int)(ey * G.inv_h) + 1);
    G.gz = min(g_long, (int)(ez * G.inv_h) + 1);
    return G;
}
__device__ __forceinline__ void knn_cell(const KnnGrid& G, float x, float y, float z, int& cx, int& cy, int& cz)
{
    cx = min(max((int)((x - G.lox) * G.inv_h), 0), G.gx - 1);
    cy = min(max((int)((y - G.loy) * G.inv_h), 0), G.gy - 1);
    cz = min(max((int)((z - G.loz) * G.inv_h), 0), G.gz - 1);
}

__global__ void __launch_bounds__(KNN_THREADS) knn_count_kernel(const float* __restrict__ pts, int N, const unsigned* __restrict__ bbox, int g_long,
                                                                int* __restrict__ cell_of, int* __restrict__ cell_count)
{
    const KnnGrid G = knn_grid(bbox, g_long);
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    int cx, cy, cz;
    knn_cell(G, pts[3 * (size_t)i], pts[3 * (size_t)i + 1], pts[3 * (size_t)i + 2], cx, cy, cz);
    const int c = (cz * G.gy + cy) * G.gx + cx;
    cell_of[i] = c;
    atomicAdd(cell_count + c, 1);
}

// exclusive scan over n ints in three passes (block sums -> scan of the block sums by one CTA -> add back)
__global__ void __launch_bounds__(KNN_SCAN_BLOCK) knn_scan1_kernel(const int* __restrict__ in, int n, int* __restrict__ out, int* __restrict__ block_sums)
{
    __shared__ int ws[32];
    const int i = blockIdx.x * KNN_SCAN_BLOCK + threadIdx.x;
    const int v = i < n ? in[i] : 0;
    int x = v;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int y = __shfl_up_sync(0xffffffffu, x, o);
        if (lane >= o) x += y;
    }
    if (lane == 31) ws[warp] = x;
    __syncthreads();
    if (warp == 0) {
        int s = ws[lane];
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int y = __shfl_up_sync(0xffffffffu, s, o);
            if (lane >= o) s += y;
        }
        ws[lane] = s;
    }
    __syncthreads();
    const int incl = x + (warp ? ws[warp - 1] : 0);
    if (i < n) out[i] = incl - v;
    if (threadIdx.x == KNN_SCAN_BLOCK - 1) block_sums[blockIdx.x] = incl;
}
__global__ void __launch_bounds__(1024) knn_scan2_kernel(int* __restrict__ block_sums, int nb)
{
    __shared__ int carry;
    __shared__ int ws[32];
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int b0 = 0; b0 < nb; b0 += 1024) {
        const int i = b0 + threadIdx.x;
        const int v = i < nb ? block_sums[i] : 0;
        int x = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int y = __shfl_up_sync(0xffffffffu, x, o);
            if (lane >= o) x += y;
        }
        if (lane == 31) ws[warp] = x;
        __syncthreads();
        if (warp == 0) {
            int s = ws[lane];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int y = __shfl_up_sync(0xffffffffu, s, o);
                if (lane >= o) s += y;
            }
            ws[lane] = s;
        }
        __syncthreads();
        const int incl = x + (warp ? ws[warp - 1] : 0) + carry;
        if (i < nb) block_sums[i] = incl - v;
        __syncthreads();
        if (threadIdx.x == 1023) carry = incl;
        __syncthreads();
    }
}
__global__ void __launch_bounds__(KNN_SCAN_BLOCK) knn_scan3_kernel(int* __restrict__ out, int n, const int* __restrict__ block_sums)
{
    const int i = blockIdx.x * KNN_SCAN_BLOCK + threadIdx.x;
    if (i < n) out[i] += block_sums[blockIdx.x];
}

__global__ void __launch_bounds__(KNN_THREADS) knn_scatter_kernel(const float* __restrict__ pts, int N, const int* __restrict__ cell_of,
                                                                  const int* __restrict__ cell_start, int* __restrict__ cursor,
                                                                  float4* __restrict__ sorted)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    const int c = cell_of[i];
    const int pos = cell_start[c] + atomicAdd(cursor + c, 1);
    sorted[pos] = make_float4(pts[3 * (size_t)i], pts[3 * (size_t)i + 1], pts[3 * (size_t)i + 2], __int_as_float(i));
}

__global__ void __launch_bounds__(KNN_THREADS) knn_query_kernel(const float4* __restrict__ sorted, int N, const unsigned* __restrict__ bbox, int g_long,
                                                                const int* __restrict__ cell_start, const int* __restrict__ cell_count,
                                                                float* __restrict__ out)
{
    const KnnGrid G = knn_grid(bbox, g_long);
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= N) return;
    const float4 p = sorted[t];
    const int self = __float_as_int(p.w);
    int cx, cy, cz;
    knn_cell(G, p.x, p.y, p.z, cx, cy, cz);
    // distance from p to the nearest face of its own cell (>= 0 up to rounding): every point outside the cube of Chebyshev
    // radius r around the cell is farther than r h + margin
    const float fx = p.x - (G.lox + cx * G.h), fy = p.y - (G.loy + cy * G.h), fz = p.z - (G.loz + cz * G.h);
    const float margin = fmaxf(fminf(fminf(fminf(fx, G.h - fx), fminf(fy, G.h - fy)), fminf(fz, G.h - fz)), 0.f);
    float b0 = INFINITY, b1 = INFINITY, b2 = INFINITY;
    const int rmax = max(max(G.gx, G.gy), G.gz);
    for (int r = 0; r <= rmax; ++r) {
        const int z0 = max(cz - r, 0), z1 = min(cz + r, G.gz - 1), y0 = max(cy - r, 0), y1 = min(cy + r, G.gy - 1);
        const int x0 = max(cx - r, 0), x1 = min(cx + r, G.gx - 1);
        for (int z = z0; z <= z1; ++z)
            for (int y = y0; y <= y1; ++y) {
                const bool shell_row = (abs(z - cz) == r) || (abs(y - cy) == r);
                // rows on the shell's faces are scanned whole; inner rows only contribute their two end cells
                for (int x = x0; x <= x1; x += (shell_row || r == 0) ? 1 : max(x1 - x0, 1)) {
                    if (!shell_row && abs(x - cx) != r) continue;
                    const int c = (z * G.gy + y) * G.gx + x;
                    const int beg = cell_start[c], n = cell_count[c];
                    for (int k = 0; k < n; ++k) {
                        const float4 q = sorted[beg + k];
                        if (__float_as_int(q.w) == self) continue;
                        const float dx = __fsub_rn(q.x, p.x), dy = __fsub_rn(q.y, p.y), dz = __fsub_rn(q.z, p.z);
                        const float d = __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
                        if (d < b0) { b2 = b1; b1 = b0; b0 = d; }
                        else if (d < b1) { b2 = b1; b1 = d; }
                        else if (d < b2) { b2 = d; }
                    }
                }
            }
        const float reach = ((float)r * G.h + margin) * 0.9999f;
        if (b2 <= reach * reach) break;
        if (x0 == 0 && y0 == 0 && z0 == 0 && x1 == G.gx - 1 && y1 == G.gy - 1 && z1 == G.gz - 1) break;   // whole grid scanned
    }
    const int k = N - 1 < 3 ? N - 1 : 3;
    float s = 0.0f;
    if (k > 0) s = __fadd_rn(s, b0);
    if (k > 1) s = __fadd_rn(s, b1);
    if (k > 2) s = __fadd_rn(s, b2);
    out[self] = k > 0 ? __fdiv_rn(s, 3.0f) : 0.0f;
}

static int knn_g_long(int N)
{
    int g = (int)floor(cbrt((double)N / 2.0));
    return g < 1 ? 1 : (g > 256 ? 256 : g);
}
static size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }

extern "C" long long nlosb200_knn_workspace_bytes(int N)
{
    if (N <= 0) return 256;
    const size_t g = (size_t)knn_g_long(N), cells = g * g * g;
    const size_t nb = (cells + KNN_SCAN_BLOCK - 1) / KNN_SCAN_BLOCK;
    return (long long)(align256(8 * sizeof(unsigned)) + align256((size_t)N * sizeof(int)) + 3 * align256(cells * sizeof(int)) +
                       align256(nb * sizeof(int)) + align256((size_t)N * sizeof(float4)));
}

extern "C" int nlosb200_knn_dist2(const float* points, int N, float* out, void* workspace, long long workspace_bytes, void* stream)
{
    NLOS_REQUIRE(N >= 0, "knn_dist2: N=%d", N);
    if (N == 0) return 0;
    NLOS_REQUIRE(points && out && workspace, "knn_dist2: null pointer");
    NLOS_REQUIRE(workspace_bytes >= nlosb200_knn_workspace_bytes(N), "knn_dist2: workspace of %lld B < %lld B", workspace_bytes,
                 nlosb200_knn_workspace_bytes(N));
    NLOS_REQUIRE(((uintptr_t)workspace % 16) == 0, "knn_dist2: workspace must be 16-byte aligned");
    cudaStream_t s = (cudaStream_t)stream;
    const int g = knn_g_long(N);
    const size_t cells = (size_t)g * g * g, nb = (cells + KNN_SCAN_BLOCK - 1) / KNN_SCAN_BLOCK;
    char* w = (char*)workspace;
    unsigned* bbox = (unsigned*)w;                w += align256(8 * sizeof(unsigned));
    int* cell_of = (int*)w;                       w += align256((size_t)N * sizeof(int));
    int* cell_count = (int*)w;                    w += align256(cells * sizeof(int));
    int* cell_start = (int*)w;                    w += align256(cells * sizeof(int));
    int* cursor = (int*)w;                        w += align256(cells * sizeof(int));
    int* block_sums = (int*)w;                    w += align256(nb * sizeof(int));
    float4* sorted = (float4*)w;
    // bbox: ordered-int +inf for the minima, -inf (0) for the maxima
    cudaError_t e = cudaMemsetAsync(bbox, 0xff, 3 * sizeof(unsigned), s);
    if (e == cudaSuccess) e = cudaMemsetAsync(bbox + 3, 0, 3 * sizeof(unsigned), s);
    if (e == cudaSuccess) e = cudaMemsetAsync(cell_count, 0, cells * sizeof(int), s);
    if (e == cudaSuccess) e = cudaMemsetAsync(cursor, 0, cells * sizeof(int), s);
    NLOS_REQUIRE(e == cudaSuccess, "knn_dist2: memset: %s", cudaGetErrorString(e));
    const int nblk = (N + KNN_THREADS - 1) / KNN_THREADS;
    knn_bbox_kernel<<<min(nblk, 148 * 8), KNN_THREADS, 0, s>>>(points, N, bbox);
    NLOS_CHECK_LAUNCH("knn_bbox_kernel");
    knn_count_kernel<<<nblk, KNN_THREADS, 0, s>>>(points, N, bbox, g, cell_of, cell_count);
    NLOS_CHECK_LAUNCH("knn_count_kernel");
    knn_scan1_kernel<<<(unsigned)nb, KNN_SCAN_BLOCK, 0, s>>>(cell_count, (int)cells, cell_start, block_sums);
    NLOS_CHECK_LAUNCH("knn_scan1_kernel");
    knn_scan2_kernel<<<1, 1024, 0, s>>>(block_sums, (int)nb);
    NLOS_CHECK_LAUNCH("knn_scan2_kernel");
    knn_scan3_kernel<<<(unsigned)nb, KNN_SCAN_BLOCK, 0, s>>>(cell_start, (int)cells, block_sums);
    NLOS_CHECK_LAUNCH("knn_scan3_kernel");
    knn_scatter_kernel<<<nblk, KNN_THREADS, 0, s>>>(points, N, cell_of, cell_start, cursor, sorted);
    NLOS_CHECK_LAUNCH("knn_scatter_kernel");
    knn_query_kernel<<<nblk, KNN_THREADS, 0, s>>>(sorted, N, bbox, g, cell_start, cell_count, out);
    NLOS_CHECK_LAUNCH("knn_query_kernel");
    return 0;
}
