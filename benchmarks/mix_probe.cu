// Measurement aid (NOT part of the product ABI): can the fp64 vector pipe (DFMA) and the fp64
// tensor path (DMMA, mma.sync.m8n8k4.f64) of one SM run at the same time?  Times three kernels
// with the same number of warps: DFMA only, DMMA only, and both interleaved in every warp.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o benchmarks/variants/mix_probe benchmarks/mix_probe.cu
#include <cuda_runtime.h>
#include <stdio.h>

template <int kFma, int kMma>
__global__ void __launch_bounds__(256) mix_probe(double* out, int iters, double a, double b) {
    double v[8], c[8][2];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        v[k] = threadIdx.x * 1e-9 + k * 1e-3;
        c[k][0] = c[k][1] = threadIdx.x * 1e-9 + k;
    }
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int u = 0; u < 4; ++u) {
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                if (kFma)
#pragma unroll
                    for (int r = 0; r < kFma; ++r) v[k] = fma(v[k], a, b);
                if (kMma)
#pragma unroll
                    for (int r = 0; r < kMma; ++r)
                        asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                                     : "+d"(c[k][0]), "+d"(c[k][1])
                                     : "d"(a), "d"(b));
            }
        }
    }
    double s = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) s += v[k] + c[k][0] + c[k][1];
    if (s == 123.456) out[0] = s;
}

template <int kFma, int kMma>
static void run(const char* name, double* out, int sms) {
    const int iters = 2048, threads = 256, blocks = sms * 8;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    float best = 1e30f;
    for (int r = 0; r < 4; ++r) {
        cudaEventRecord(e0);
        mix_probe<kFma, kMma><<<blocks, threads>>>(out, iters, 0.999999, 1e-7);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms = 0.f;
        cudaEventElapsedTime(&ms, e0, e1);
        if (r > 0 && ms < best) best = ms;
    }
    const double warps = (double)blocks * threads / 32, steps = (double)iters * 32;
    const double fma_tf = 2.0 * 32 * kFma * steps * warps / (best * 1e-3) / 1e12;
    const double mma_tf = 512.0 * kMma * steps * warps / (best * 1e-3) / 1e12;
    printf("%-28s %8.3f ms   DFMA %6.2f TF/s   DMMA %6.2f TF/s   sum %6.2f TF/s\n", name, best, fma_tf, mma_tf,
           fma_tf + mma_tf);
}

int main() {
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    double* out = nullptr;
    cudaMalloc(&out, 8);
    run<1, 0>("DFMA only", out, sms);
    run<0, 1>("DMMA only", out, sms);
    run<1, 1>("1 DFMA : 1 DMMA per step", out, sms);
    run<4, 1>("4 DFMA : 1 DMMA per step", out, sms);
    run<8, 1>("8 DFMA : 1 DMMA per step", out, sms);
    run<1, 2>("1 DFMA : 2 DMMA per step", out, sms);
    cudaFree(out);
    return 0;
}
