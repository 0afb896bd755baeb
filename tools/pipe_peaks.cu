// pipe_peaks.cu -- microbenchmark of the two ceilings the transient kernel's roofline uses: FP32 FMA and MUFU.EX2 issue rates,
// plus the packed fp32x2 forms (FFMA2 / FMUL2 / FADD2) the kernel is written in.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o pipe_peaks tools/pipe_peaks.cu && ./pipe_peaks
// Every thread runs ILP independent dependency chains; grid = 148 SMs x 4 CTAs x 512 threads (all 64 warps per SM resident).
#include <cstdio>
#include <cuda_runtime.h>

#define ILP 8
#define ITERS 4096

template <int OP>
__global__ void __launch_bounds__(512) burn(float* out, float seed)
{
    float x[ILP];
    unsigned long long p[ILP];
#pragma unroll
    for (int i = 0; i < ILP; ++i) {
        x[i] = seed + threadIdx.x * 1e-6f + i;
        p[i] = ((unsigned long long)__float_as_uint(x[i]) << 32) | __float_as_uint(x[i] + 0.5f);
    }
    const float a = 0.999f, b = 1e-3f;
    const unsigned long long pa = ((unsigned long long)__float_as_uint(a) << 32) | __float_as_uint(a);
    const unsigned long long pb = ((unsigned long long)__float_as_uint(b) << 32) | __float_as_uint(b);
#pragma unroll 1
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < ILP; ++i) {
            if (OP == 0) x[i] = fmaf(x[i], a, b);                                                            // FFMA
            if (OP == 1) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(x[i]));                             // MUFU.EX2
            if (OP == 2) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(p[i]) : "l"(pa), "l"(pb));        // FFMA2
            if (OP == 3) asm volatile("mul.rn.f32x2 %0, %0, %1;" : "+l"(p[i]) : "l"(pa));                     // FMUL2
            if (OP == 4) asm volatile("add.rn.f32x2 %0, %0, %1;" : "+l"(p[i]) : "l"(pb));                     // FADD2
            if (OP == 5) { x[i] = fmaf(x[i], a, b); if ((i & 3) == 0) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(x[i])); }   // 4 FFMA : 1 MUFU
        }
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < ILP; ++i) s += x[i] + __uint_as_float((unsigned)(p[i] >> 32)) + __uint_as_float((unsigned)p[i]);
    if (s == 12345.678f) out[0] = s;
}

template <int OP>
static void run(const char* name, double per_thread_iter, const char* unit, double clk_ghz, int sms, double per_inst_scale)
{
    float* out;
    cudaMalloc(&out, 4);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int grid = sms * 4;
    burn<OP><<<grid, 512>>>(out, 1.0f);
    cudaDeviceSynchronize();
    float best = 1e30f;
    for (int r = 0; r < 5; ++r) {
        cudaEventRecord(e0);
        burn<OP><<<grid, 512>>>(out, 1.0f);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms;
        cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    const double threads = (double)grid * 512;
    const double ops = threads * ITERS * per_thread_iter;                 // counted operations
    const double rate = ops / (best * 1e-3);
    const double warp_inst_per_clk_sm = threads / 32.0 * ITERS * ILP * per_inst_scale / (best * 1e-3) / (clk_ghz * 1e9) / sms;
    printf("%-28s %8.3f ms  %10.3f %s   %.2f warp-inst/clk/SM at %.3f GHz\n", name, best, rate / 1e12, unit, warp_inst_per_clk_sm, clk_ghz);
    cudaFree(out);
}

int main()
{
    cudaDeviceProp pr;
    cudaGetDeviceProperties(&pr, 0);
    int clk_khz = 0;
    cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0);
    const double ghz = clk_khz * 1e-6;
    printf("%s: %d SMs, max SM clock %.3f GHz (rates below are per second; warp-inst/clk assumes the max clock)\n", pr.name, pr.multiProcessorCount, ghz);
    const int sms = pr.multiProcessorCount;
    run<0>("FFMA (2 flop)", ILP * 2.0, "TFLOP/s", ghz, sms, 1.0);
    run<2>("FFMA2 (4 flop)", ILP * 4.0, "TFLOP/s", ghz, sms, 1.0);
    run<3>("FMUL2 (2 flop)", ILP * 2.0, "TFLOP/s", ghz, sms, 1.0);
    run<4>("FADD2 (2 flop)", ILP * 2.0, "TFLOP/s", ghz, sms, 1.0);
    run<1>("MUFU.EX2", ILP * 1.0, "T ex2/s", ghz, sms, 1.0);
    run<5>("4 FFMA : 1 MUFU (flop)", ILP * 2.0, "TFLOP/s", ghz, sms, 1.25);
    return 0;
}
