// Measurement aid (NOT part of the product ABI): fp64 FMA peak of the device, the
// roofline denominator for the emulator kernel (MEASURED_PEAKS.json has no fp64
// figure).  Each thread runs `iters` rounds of 8 independent dependent-FMA chains.
#include <cuda_runtime.h>
#include <stdint.h>

__global__ void __launch_bounds__(256) fp64_fma_probe(double* out, int iters, double a, double b) {
    double v0 = threadIdx.x * 1e-9, v1 = v0 + 1e-3, v2 = v0 + 2e-3, v3 = v0 + 3e-3;
    double v4 = v0 + 4e-3, v5 = v0 + 5e-3, v6 = v0 + 6e-3, v7 = v0 + 7e-3;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int k = 0; k < 16; ++k) {
            v0 = fma(v0, a, b); v1 = fma(v1, a, b); v2 = fma(v2, a, b); v3 = fma(v3, a, b);
            v4 = fma(v4, a, b); v5 = fma(v5, a, b); v6 = fma(v6, a, b); v7 = fma(v7, a, b);
        }
    }
    double s = v0 + v1 + v2 + v3 + v4 + v5 + v6 + v7;
    if (s == 123.456) out[0] = s;     // never true; keeps the chains alive
}

// Same, but every FMA reads THREE register operands (the emulator loop's shape: G += w * zh, a += zh * x)
// instead of one register and two constants: tells whether operand delivery from the register file lowers
// the reachable rate.  v_k = v_k * p_k + q_k on 8 chains, p/q rotated so that nothing folds into constants.
__global__ void __launch_bounds__(256) fp64_fma3_probe(double* out, int iters, const double* __restrict__ coef) {
    double v[8], p[8], q[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        v[k] = threadIdx.x * 1e-9 + k * 1e-3;
        p[k] = coef[k] - threadIdx.x * 1e-12;            // thread-dependent: must live in vector registers
        q[k] = coef[8 + k] + threadIdx.x * 1e-15;
    }
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int u = 0; u < 16; ++u)
#pragma unroll
            for (int k = 0; k < 8; ++k) v[k] = fma(v[k], p[(k + u) & 7], q[(k + 3 * u) & 7]);
    }
    double s = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) s += v[k];
    if (s == 123.456) out[0] = s;
}

// The emulator loop's OPERAND MIX without its dependent chain: per row and pair of pixels 20 dot-product FMAs
// (row value from a broadcast LDS, pixel state and accumulator from registers: the first pixel's FMA reads three
// registers, the second reuses the row value), 16 two-operand fp64 instructions standing in for the
// exponential, 20 accumulations (weight reused, row value and sum from registers) and 2 adds; 16 warps per SM,
// so latency is hidden and what is measured is how fast the register file and the pipe deliver this mix.
__global__ void __launch_bounds__(128, 4) fp64_loop_shape_probe(double* out, int rows, const double* __restrict__ tab) {
    __shared__ __align__(16) double s_tab[64 * 12];
    for (int k = threadIdx.x; k < 64 * 12; k += 128) s_tab[k] = tab[k];
    __syncthreads();
    double x0[10], x1[10], g0[10], g1[10], m0 = 0, m1 = 0;
#pragma unroll
    for (int d = 0; d < 10; ++d) {
        x0[d] = 1e-3 * (threadIdx.x + d);
        x1[d] = 2e-3 * (threadIdx.x + d) + 1e-4;
        g0[d] = g1[d] = 0;
    }
    for (int j = 0; j < rows; ++j) {
        const double* row = s_tab + (j & 63) * 12;
        double z[10];
#pragma unroll
        for (int d = 0; d < 10; d += 2) {
            double2 v = *reinterpret_cast<const double2*>(row + d);
            z[d] = v.x;
            z[d + 1] = v.y;
        }
        double a0 = row[10], a1 = row[10];
#pragma unroll
        for (int d = 0; d < 10; ++d) {
            a0 = fma(z[d], x0[d], a0);
            a1 = fma(z[d], x1[d], a1);
        }
        double p0 = a0, p1 = a1;
#pragma unroll
        for (int u = 0; u < 8; ++u) {                  // 8 two-operand fp64 instructions per pixel (3 adds, 3 FMAs with
            p0 = fma(p0, 0.999, 1e-3);                 // constant coefficients, a multiply, an add in the real loop)
            p1 = fma(p1, 0.999, 1e-3);
        }
#pragma unroll
        for (int d = 0; d < 10; ++d) {
            g0[d] = fma(p0, z[d], g0[d]);
            g1[d] = fma(p1, z[d], g1[d]);
        }
        m0 += p0;
        m1 += p1;
    }
    double s = m0 + m1;
#pragma unroll
    for (int d = 0; d < 10; ++d) s += g0[d] + g1[d];
    if (s == 123.456) out[0] = s;
}

// fp64 tensor-core path: mma.sync m8n8k4 (DMMA), 8 independent accumulator chains per warp
__global__ void __launch_bounds__(256) fp64_dmma_probe(double* out, int iters, double a, double b) {
    double c[8][2];
#pragma unroll
    for (int k = 0; k < 8; ++k) c[k][0] = c[k][1] = threadIdx.x * 1e-9 + k;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int u = 0; u < 4; ++u)
#pragma unroll
            for (int k = 0; k < 8; ++k)
                asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                             : "+d"(c[k][0]), "+d"(c[k][1])
                             : "d"(a), "d"(b));
    }
    double s = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) s += c[k][0] + c[k][1];
    if (s == 123.456) out[0] = s;
}

// Planar streaming pattern of the solve kernel: every thread reads R planes and writes W planes
// (plane stride ld doubles), nothing else.  Tells how much of the copy bandwidth this access
// pattern can reach, independent of the arithmetic.
__global__ void __launch_bounds__(128) planar_stream_probe(const double* __restrict__ in, double* __restrict__ out,
                                                           long long n, long long ld, int R, int W) {
    long long i = (long long)blockIdx.x * 128 + threadIdx.x;
    if (i >= n) return;
    double acc = 0.0;
    for (int r = 0; r < R; r += 4) {
        double a = __ldcs(in + (long long)r * ld + i);
        double b = r + 1 < R ? __ldcs(in + (long long)(r + 1) * ld + i) : 0.0;
        double c = r + 2 < R ? __ldcs(in + (long long)(r + 2) * ld + i) : 0.0;
        double d = r + 3 < R ? __ldcs(in + (long long)(r + 3) * ld + i) : 0.0;
        acc += (a + b) + (c + d);
    }
    for (int w = 0; w < W; ++w) __stcs(out + (long long)w * ld + i, acc + w);
}

// Dependent-issue latency of DFMA: one CTA per SM, `warps` warps per scheduler, CH independent
// chains per thread; clock64 around a long run of dependent FMAs.  cycles / (steps) = how often a
// warp can issue the next FMA of ONE chain set, i.e. max(latency, CH * 2 * warps-sharing-the-pipe).
template <int CH>
__global__ void dfma_latency_probe(double* out, long long* cycles, int iters, double a, double b) {
    double v[CH];
#pragma unroll
    for (int c = 0; c < CH; ++c) v[c] = threadIdx.x * 1e-9 + c * 1e-3;
    __syncthreads();
    long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int k = 0; k < 32; ++k)
#pragma unroll
            for (int c = 0; c < CH; ++c) v[c] = fma(v[c], a, b);
    }
    long long t1 = clock64();
    double s = 0;
#pragma unroll
    for (int c = 0; c < CH; ++c) s += v[c];
    if (s == 123.456) out[0] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) cycles[0] = t1 - t0;
}

extern "C" {
// cycles between successive dependent steps of a thread's CH chains with `warps` warps per scheduler
double kafka_probe_dfma_step_cycles(int chains, int warps) {
    double* out = nullptr;
    long long* cyc = nullptr;
    if (cudaMalloc(&out, 8) != cudaSuccess || cudaMalloc(&cyc, 8) != cudaSuccess) return -1.0;
    const int iters = 2000, threads = 128 * warps;
    for (int rep = 0; rep < 2; ++rep) {
        switch (chains) {
            case 1: dfma_latency_probe<1><<<1, threads>>>(out, cyc, iters, 0.999999, 1e-7); break;
            case 2: dfma_latency_probe<2><<<1, threads>>>(out, cyc, iters, 0.999999, 1e-7); break;
            case 3: dfma_latency_probe<3><<<1, threads>>>(out, cyc, iters, 0.999999, 1e-7); break;
            case 4: dfma_latency_probe<4><<<1, threads>>>(out, cyc, iters, 0.999999, 1e-7); break;
            case 6: dfma_latency_probe<6><<<1, threads>>>(out, cyc, iters, 0.999999, 1e-7); break;
            case 8: dfma_latency_probe<8><<<1, threads>>>(out, cyc, iters, 0.999999, 1e-7); break;
            default: return -2.0;
        }
        if (cudaDeviceSynchronize() != cudaSuccess) return -1.0;
    }
    long long h = 0;
    cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
    cudaFree(out);
    cudaFree(cyc);
    return (double)h / ((double)iters * 32);
}

double kafka_probe_planar_stream_gbs(long long n, int R, int W, int reps) {
    double *in = nullptr, *out = nullptr;
    long long ld = n;
    if (cudaMalloc(&in, (size_t)R * ld * 8) != cudaSuccess) return -1.0;
    if (cudaMalloc(&out, (size_t)W * ld * 8) != cudaSuccess) { cudaFree(in); return -1.0; }
    cudaMemset(in, 0, (size_t)R * ld * 8);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    double best = 0.0;
    for (int r = 0; r < reps + 1; ++r) {
        cudaEventRecord(e0);
        planar_stream_probe<<<(unsigned)((n + 127) / 128), 128>>>(in, out, n, ld, R, W);
        cudaEventRecord(e1);
        if (cudaEventSynchronize(e1) != cudaSuccess) { best = -1.0; break; }
        float ms = 0.f;
        cudaEventElapsedTime(&ms, e0, e1);
        double gbs = (double)(R + W) * n * 8 / (ms * 1e-3) / 1e9;
        if (r > 0 && gbs > best) best = gbs;
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(in);
    cudaFree(out);
    return best;
}

double kafka_probe_dmma_tflops(int reps) {
    int dev = 0, sms = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    double* out = nullptr;
    if (cudaMalloc(&out, 8) != cudaSuccess) return -1.0;
    const int iters = 2048, threads = 256, blocks = sms * 8;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    double best = 0.0;
    for (int r = 0; r < reps + 1; ++r) {
        cudaEventRecord(e0);
        fp64_dmma_probe<<<blocks, threads>>>(out, iters, 0.999999, 1e-7);
        cudaEventRecord(e1);
        if (cudaEventSynchronize(e1) != cudaSuccess) { best = -1.0; break; }
        float ms = 0.f;
        cudaEventElapsedTime(&ms, e0, e1);
        double flops = 512.0 * 32 * (double)iters * (threads / 32) * blocks;
        double tf = flops / (ms * 1e-3) / 1e12;
        if (r > 0 && tf > best) best = tf;
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(out);
    return best;
}

// executed TFLOP/s (2 x 58 fp64 instructions per row and pixel pair) of the loop-shaped probe, best of `reps`
double kafka_probe_fp64_loop_shape_tflops(int reps) {
    int dev = 0, sms = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    double *out = nullptr, *tab = nullptr;
    if (cudaMalloc(&out, 8) != cudaSuccess || cudaMalloc(&tab, 64 * 12 * 8) != cudaSuccess) return -1.0;
    double h[64 * 12];
    for (int k = 0; k < 64 * 12; ++k) h[k] = 1e-3 * ((k * 37) % 101) - 0.05;
    cudaMemcpy(tab, h, sizeof(h), cudaMemcpyHostToDevice);
    const int rows = 20000, threads = 128, blocks = sms * 16;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    double best = 0.0;
    for (int r = 0; r < reps + 1; ++r) {
        cudaEventRecord(e0);
        fp64_loop_shape_probe<<<blocks, threads>>>(out, rows, tab);
        cudaEventRecord(e1);
        if (cudaEventSynchronize(e1) != cudaSuccess) { best = -1.0; break; }
        float ms = 0.f;
        cudaEventElapsedTime(&ms, e0, e1);
        double tf = 2.0 * 58 * (double)rows * threads * blocks / (ms * 1e-3) / 1e12;
        if (r > 0 && tf > best) best = tf;
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(out);
    cudaFree(tab);
    return best;
}

// DFMA with three register operands, TFLOP/s, best of `reps`
double kafka_probe_fp64_fma3_tflops(int reps) {
    int dev = 0, sms = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    double *out = nullptr, *coef = nullptr;
    if (cudaMalloc(&out, 8) != cudaSuccess || cudaMalloc(&coef, 128) != cudaSuccess) return -1.0;
    double h[16];
    for (int k = 0; k < 16; ++k) h[k] = k < 8 ? 0.999999 - 1e-9 * k : 1e-7 * (k - 7);
    cudaMemcpy(coef, h, sizeof(h), cudaMemcpyHostToDevice);
    const int iters = 4096, threads = 256, blocks = sms * 8;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    double best = 0.0;
    for (int r = 0; r < reps + 1; ++r) {
        cudaEventRecord(e0);
        fp64_fma3_probe<<<blocks, threads>>>(out, iters, coef);
        cudaEventRecord(e1);
        if (cudaEventSynchronize(e1) != cudaSuccess) { best = -1.0; break; }
        float ms = 0.f;
        cudaEventElapsedTime(&ms, e0, e1);
        double tf = 2.0 * 8 * 16 * (double)iters * threads * blocks / (ms * 1e-3) / 1e12;
        if (r > 0 && tf > best) best = tf;
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(out);
    cudaFree(coef);
    return best;
}

// returns achieved TFLOP/s (2 flops per FMA), best of `reps`
double kafka_probe_fp64_tflops(int reps) {
    int dev = 0, sms = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    double* out = nullptr;
    if (cudaMalloc(&out, 8) != cudaSuccess) return -1.0;
    const int iters = 4096, threads = 256, blocks = sms * 8;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    double best = 0.0;
    for (int r = 0; r < reps + 1; ++r) {
        cudaEventRecord(e0);
        fp64_fma_probe<<<blocks, threads>>>(out, iters, 0.999999, 1e-7);
        cudaEventRecord(e1);
        if (cudaEventSynchronize(e1) != cudaSuccess) { best = -1.0; break; }
        float ms = 0.f;
        cudaEventElapsedTime(&ms, e0, e1);
        double flops = 2.0 * 8 * 16 * (double)iters * threads * blocks;
        double tf = flops / (ms * 1e-3) / 1e12;
        if (r > 0 && tf > best) best = tf;
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(out);
    return best;
}
}
