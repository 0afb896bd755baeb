// loss.cu -- fused squared-error loss and its gradient on the tile-major transient.
//
// Replaces, for the batched fitting step, what the reference does per scan point in Python
// (/root/reference/nlos_helpers.py:280-346 compute_loss: loss = sum((pred_histogram - measured)^2) over the bins of ONE scan
// point per step, accumulated over the scan points by the caller, nlos_helpers.py:323-325; the batched variant is an empty stub,
// nlos_helpers.py:348-351): one pass over H[n_rows, T] produces  loss += inv_numel * sum (H - target)^2  and
// dL/dH = 2 inv_numel (H - target)  in the layout the backward kernel consumes -- no untile / tile round trip, one read of H
// and target, one write of dL/dH (HBM-bound: 12 B per bin).
#include "common.cuh"
#include "../../include/nlosb200.h"

using namespace nlos;

__global__ void __launch_bounds__(256) mse_loss_grad_kernel(const float4* __restrict__ H, const float4* __restrict__ target, long long n4,
                                                            float inv_numel, float4* __restrict__ gH, double* __restrict__ loss)
{
    float acc = 0.f;
    const float g2 = 2.0f * inv_numel;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
        const float4 h = H[i], t = target[i];
        const float dx = h.x - t.x, dy = h.y - t.y, dz = h.z - t.z, dw = h.w - t.w;
        acc += (dx * dx + dy * dy) + (dz * dz + dw * dw);
        if (gH) gH[i] = make_float4(g2 * dx, g2 * dy, g2 * dz, g2 * dw);
    }
    acc = warp_sum(acc);
    __shared__ float ws[8];
    if ((threadIdx.x & 31) == 0) ws[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x < 8) {
        float v = ws[threadIdx.x];
        v += __shfl_xor_sync(0xffu, v, 4);
        v += __shfl_xor_sync(0xffu, v, 2);
        v += __shfl_xor_sync(0xffu, v, 1);
        if (threadIdx.x == 0) atomicAdd(loss, (double)v * (double)inv_numel);
    }
}

// H, target, grad_H: [n] fp32 (n % 4 == 0, 16-byte aligned; padding rows of the tile-major layout must hold equal values in H and
// target, e.g. zeros).  loss: one double, ACCUMULATED (caller zeroes).  grad_H may be NULL (loss only).
extern "C" int nlosb200_mse_loss_grad(const float* H, const float* target, long long n, float inv_numel, float* grad_H, double* loss, void* stream)
{
    NLOS_REQUIRE(n >= 0 && (n & 3) == 0, "mse_loss_grad: n=%lld must be a non-negative multiple of 4", n);
    if (n == 0) return 0;
    NLOS_REQUIRE(H && target && loss, "mse_loss_grad: null pointer");
    NLOS_REQUIRE(((uintptr_t)H % 16) == 0 && ((uintptr_t)target % 16) == 0 && ((uintptr_t)grad_H % 16) == 0, "mse_loss_grad: 16-byte alignment");
    const long long n4 = n / 4;
    const int blocks = (int)((n4 + 255) / 256 < 148 * 8 ? (n4 + 255) / 256 : 148 * 8);
    mse_loss_grad_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(reinterpret_cast<const float4*>(H), reinterpret_cast<const float4*>(target), n4,
                                                                    inv_numel, reinterpret_cast<float4*>(grad_H), loss);
    NLOS_CHECK_LAUNCH("mse_loss_grad_kernel");
    return 0;
}
