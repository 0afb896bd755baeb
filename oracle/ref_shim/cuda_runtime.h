/* Host shim so that the reference's device headers (submodules/cuda_renderer/include/*.cuh)
 * and its __global__ kernel bodies compile as plain C++ with g++.  TEST INFRASTRUCTURE ONLY
 * (used by oracle/build_ref.py to build oracle/_ref/libnlos_ref.so). */
#pragma once
#include <cmath>
#include <cstdint>
#include <algorithm>
#ifndef M_PI
#define M_PI 3.14159265358979323846
#endif
#define __device__
#define __host__
#define __global__
#define __forceinline__ inline
#define __restrict__
struct float3 { float x, y, z; };
struct float4 { float x, y, z, w; };
static inline float3 make_float3(float x, float y, float z) { float3 r; r.x = x; r.y = y; r.z = z; return r; }
static inline float4 make_float4(float x, float y, float z, float w) { float4 r; r.x = x; r.y = y; r.z = z; r.w = w; return r; }
struct shim_dim3 { unsigned x, y, z; };
static thread_local shim_dim3 threadIdx = {0, 0, 0};
static thread_local shim_dim3 blockIdx = {0, 0, 0};
static thread_local shim_dim3 blockDim = {1, 1, 1};
#define TORCH_CHECK(...)
