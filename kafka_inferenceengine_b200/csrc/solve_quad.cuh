// Four lanes per pixel: the update of solve_bands_kernel for the largest blocks (n_params 13-16).
// One thread per pixel needs 91-136 packed entries in registers (237+ registers: 8 warps per SM,
// or spills), two lanes per pixel (solve_pair.cuh) still 254 registers; with 8 warps per SM both
// wait on the factorisation's dependent chain and stay at 60-83 % of the HBM peak.  Here a pixel
// is shared by the four lanes of a quad -- 40 packed entries per lane at 16 parameters, 128
// registers, 16 warps per SM -- so that twice as many independent chains hide that latency.
//
// Lane g (0..3) owns the rows r = 4k + g of the packed lower triangle, k = 0..K-1, K = ceil(P/4);
// row k of a lane sits at a[2k(k+1) .. 2k(k+1) + 4k + 3] (columns 0 .. 4k+3: the slots right of the
// diagonal are unused and may hold anything after the band loop; nothing reads them).  P is padded
// to a multiple of four with decoupled unit rows.  What crosses rows goes through shuffles inside
// the quad, entry by entry in the order of the one-thread kernel:
//   * right-looking Cholesky: the pivot and then every entry of column j are broadcast from their
//     owners while all four lanes update their own rows;
//   * forward substitution: y_j broadcast from its owner; backward: every lane sums over its own
//     rows and two __shfl_xor add up the quad.
// Band values are staged per PIXEL by cp.async (the four lanes copy every fourth value) and read by
// all four after __syncwarp.  Band kinds: per-pixel arrays and one shared linearisation; an image
// with LINEAR or WCM bands at these sizes takes the two-lane kernel.  256 threads = 64 pixels per CTA, two
// CTAs per SM; scratch and the deterministic reduction are those of solve_bands_kernel.
#pragma once

namespace kafka {

// Measured on B200 (profiles/r02_bench_c5_quad_vs_pair.json, 1e7 px): 34-45 % of the HBM peak at 13-16 parameters
// against 58-69 % for the two-lane kernel, so it is NOT the default anywhere; KAFKA_SOLVE_QUAD selects it.
#ifndef KAFKA_QUAD_FROM_PARAMS
#define KAFKA_QUAD_FROM_PARAMS 99
#endif
constexpr int kQuadFromParams = KAFKA_QUAD_FROM_PARAMS;
constexpr int kQuadLanes = 4;
constexpr int kQuadPix = kSolveThreads / 2;                  // 64 pixels = 256 threads per CTA, two CTAs per SM (see kPairPix)
constexpr int kQuadThreads = kQuadLanes * kQuadPix;

__host__ __device__ constexpr int quad_off(int k) { return 2 * k * (k + 1); }

__device__ __forceinline__ double quad_bcast(double v, unsigned lane) { return __shfl_sync(0xffffffffu, v, lane); }
__device__ __forceinline__ double quad_sum(double v) {
    v += __shfl_xor_sync(0xffffffffu, v, 1);
    v += __shfl_xor_sync(0xffffffffu, v, 2);
    return v;
}

template <int P, bool kAllIdentity = false>
__global__ void __launch_bounds__(kQuadThreads, 2)
solve_bands_quad_kernel(const __grid_constant__ SolveParams prm, int n_dim_max) {
    constexpr int G = kQuadLanes;
    constexpr int K = (P + G - 1) / G;
    constexpr int PE = G * K;
    constexpr int LA = quad_off(K);
    extern __shared__ __align__(16) double stages[];               // kAsyncStages x (3 + n_dim_max) x kQuadPix
    __shared__ double s_part[2][kQuadThreads / 32];
    __shared__ __align__(16) double s_uni[KAFKA_MAX_BANDS][P + 2];
    const int g = threadIdx.x & (G - 1);
    const int slot = threadIdx.x >> 2;
    const unsigned quad_base = (threadIdx.x & 31u) & ~3u;
    const int64_t tile0 = (int64_t)blockIdx.x * kQuadPix;
    const int64_t i = tile0 + slot;
    const int64_t ld = prm.ld;
    const int stage_doubles = (3 + n_dim_max) * kQuadPix;
    const bool live = i < prm.n_pix;
    const bool in_ld = i < ld;
    double ssq = 0.0, bad = 0.0;

    // value v of a band sits at stage[v * kQuadPix + slot]: 0 = w, 1 = y, 2 = H0, 3 + d = Jacobian plane d;
    // lane g copies the values v = g, g + 4, g + 8, ...  Selects and running pointers, no loop over v with
    // a per-lane trip count (that compiles to a jump table the lanes of a quad diverge on).
    auto issue = [&](int b) {
        if (b < prm.n_bands && in_ld) {
            const SolveBandDev& bd = prm.bands[b];
            double* st = stages + (b % kAsyncStages) * stage_doubles + slot;
            const bool arrays = kAllIdentity || bd.kind == KAFKA_BAND_ARRAYS;
            const double* first = g == 0 ? bd.w : (g == 1 ? bd.y : (g == 2 ? bd.h0 : bd.jac));
            if (arrays || g < 2) cp_async8(st + g * kQuadPix, first + i);
            if (arrays) {
                const double* src = bd.jac + (int64_t)(g + 1) * ld + i;            // planes g+1, g+5, ...
                double* dst = st + (4 + g) * kQuadPix;
                const int n_planes = kAllIdentity ? P : bd.n_planes;
                const int n_mine = (n_planes - g + 2) / 4;                         // planes g+1+4t < n_planes
#pragma unroll
                for (int t = 0; t < KAFKA_MAX_DIM / 4; ++t, src += 4 * ld, dst += 4 * kQuadPix)
                    if (t < n_mine) cp_async8(dst, src);
            }
        }
        cp_async_commit();
    };
#pragma unroll
    for (int b = 0; b < kAsyncStages - 1; ++b) issue(b);

    bool any_uniform = false;
    if (!kAllIdentity)
        for (int b = 0; b < prm.n_bands; ++b) any_uniform |= prm.bands[b].kind == KAFKA_BAND_UNIFORM;
    if (any_uniform) {
        for (int k = threadIdx.x; k < prm.n_bands * (P + 1); k += kQuadThreads) {
            const int b = k / (P + 1), p = k - b * (P + 1);
            const SolveBandDev& bd = prm.bands[b];
            double v = 0.0;
            if (bd.kind == KAFKA_BAND_UNIFORM) v = p == P ? bd.h0[0] : (bd.src[p] >= 0 ? bd.jac[bd.src[p]] : 0.0);
            s_uni[b][p] = v;
        }
        __syncthreads();
    }

    const int64_t il = live ? i : tile0;
    const bool bcast = (prm.flags & KAFKA_PRIOR_BROADCAST) != 0;
    double a[LA], rown[K], xpo[K];
    {   // packed entry c = r(r+1)/2 + q of row r = 4k + g: consecutive rows of a lane start
        // 16k + 4g + 10 planes apart, so the address is a running pointer
        const int64_t step = bcast ? 1 : ld;
        const double* rowp = prm.pinv_f + (bcast ? 0 : il) + (int64_t)(g * (g + 1) / 2) * step;
        const int64_t g4 = 4 * (int64_t)g * step;
#pragma unroll
        for (int k = 0; k < K; ++k) {
            const double* ep = rowp;
            const int r = G * k + g;
#pragma unroll
            for (int q = 0; q < G * k + G; ++q, ep += step) {
                const bool valid = q <= r && r < P;
                double v = (q == r) ? 1.0 : 0.0;                    // padding row: decoupled unit block
                if (valid) v = bcast ? *ep : ldg_stream(ep);
                a[quad_off(k) + q] = (q <= r) ? v : 0.0;
            }
            rowp += (int64_t)(G * G * k + G * (G + 1) / 2) * step + g4;
        }
    }
    {
    double xfo[K];                                                  // own rows of x_f: only for the right-hand side
#pragma unroll
    for (int k = 0; k < K; ++k) {
        const int r = G * k + g;
        const bool in = r < P;
        xpo[k] = in ? ldg_stream(prm.x_prev + (int64_t)r * ld + il) : 0.0;
        xfo[k] = in ? __ldg(prm.x_f + (int64_t)r * ld + il) : 0.0;
        rown[k] = 0.0;
    }
    // rhs = P_f^-1 x_f, one column at a time: the row part needs x_f[q] (the four lanes read the same
    // address), the column part sum_{r > q} A(r,q) x_f[r] is summed over the quad and kept by q's owner
#pragma unroll
    for (int q = 0; q < PE; ++q) {
        const int kq = q / G, gq = q % G;
        const double xq = q < P ? __ldg(prm.x_f + (int64_t)(q < P ? q : 0) * ld + il) : 0.0;
        double c = 0.0;
#pragma unroll
        for (int k = 0; k < K; ++k)
            if (k >= kq) {
                const double v = a[quad_off(k) + q];                 // 0 right of the diagonal
                rown[k] = fma(v, xq, rown[k]);                        // row r: A(r,q) x_q, q <= r
                const bool below = k > kq || g > gq;                  // strictly below the diagonal: r > q
                c = fma(below ? v : 0.0, xfo[k], c);
            }
        c = quad_sum(c);
        if (g == gq) rown[kq] += c;
    }
    }

    for (int b = 0; b < prm.n_bands; ++b) {
        issue(b + kAsyncStages - 1);
        cp_async_wait<kAsyncStages - 1>();
        __syncwarp();                                               // the other lanes' copies of band b are visible too
        const SolveBandDev& bd = prm.bands[b];
        const double* st = stages + (b % kAsyncStages) * stage_doubles + slot;
        const double w = in_ld ? st[0] : 0.0;
        const double y = in_ld ? st[kQuadPix] : 0.0;
        const bool arrays = kAllIdentity || bd.kind == KAFKA_BAND_ARRAYS;
        const bool ident = kAllIdentity || bd.identity;
        const bool on = w != 0.0;
        // J(p): Jacobian entry of state slot p for this pixel (all four lanes read the same address)
        auto jac_at = [&](int p) -> double {
            if (arrays) {
                const int pl = ident ? p : bd.src[p];
                return pl >= 0 ? st[(3 + pl) * kQuadPix] : 0.0;
            }
            return on ? s_uni[b][p] : 0.0;
        };
        const double h0 = arrays ? st[2 * kQuadPix] : (on ? s_uni[b][P] : 0.0);
        double jr[K], part = 0.0;
#pragma unroll
        for (int k = 0; k < K; ++k) {
            const int r = G * k + g;
            jr[k] = r < P ? jac_at(r < P ? r : 0) : 0.0;
            part = fma(jr[k], xpo[k], part);
        }
        const double wy = w * ((y - h0) + quad_sum(part));
        double wj[K];
#pragma unroll
        for (int k = 0; k < K; ++k) {
            wj[k] = w * jr[k];
            rown[k] = fma(wy, jr[k], rown[k]);
        }
#pragma unroll
        for (int q = 0; q < PE; ++q) {
            const double jq = q < P ? jac_at(q < P ? q : 0) : 0.0;
#pragma unroll
            for (int k = 0; k < K; ++k)
                if (G * k + G > q) a[quad_off(k) + q] = fma(wj[k], jq, a[quad_off(k) + q]);
        }
        if (prm.innovations && live && (b & 3) == g) stg_stream(prm.innovations + (int64_t)b * ld + i, y - h0);
        __syncwarp();                                               // all four lanes are done with this stage
    }
    cp_async_wait<0>();

    if (live && prm.a_out) {
        double* rowp = prm.a_out + i + (int64_t)(g * (g + 1) / 2) * ld;            // same walk as the prior load
        const int64_t g4 = 4 * (int64_t)g * ld;
#pragma unroll
        for (int k = 0; k < K; ++k) {
            double* ep = rowp;
            const int r = G * k + g;
#pragma unroll
            for (int q = 0; q < G * k + G; ++q, ep += ld)
                if (q <= r && r < P) stg_stream(ep, a[quad_off(k) + q]);
            rowp += (int64_t)(G * G * k + G * (G + 1) / 2) * ld + g4;
        }
    }

    // ---- Cholesky, right-looking; 1/L_jj is left on the diagonal ----
    bool ok = true;
#pragma unroll
    for (int j = 0; j < PE; ++j) {
        const int kj = j / G, gj = j % G;
        const double d = quad_bcast(a[quad_off(kj) + j], quad_base + gj);
        ok = ok && (d > 0.0);
        const double inv = rsqrt(d);
#pragma unroll
        for (int k = 0; k < K; ++k)
            if (k >= kj) a[quad_off(k) + j] *= inv;                 // rows > j: L(r,j); unused slots and L(j,j): no harm
        if (g == gj) a[quad_off(kj) + j] = inv;
#pragma unroll
        for (int q = 0; q < PE; ++q)
            if (q > j) {
                const double cq = quad_bcast(a[quad_off(q / G) + j], quad_base + (q % G));      // L(q,j)
#pragma unroll
                for (int k = 0; k < K; ++k)
                    if (G * k + G > q && k >= kj)                   // rows that can lie below j and have a slot q
                        a[quad_off(k) + q] = fma(-a[quad_off(k) + j], cq, a[quad_off(k) + q]);
            }
    }

    // ---- forward substitution: y_j broadcast from its owner ----
#pragma unroll
    for (int j = 0; j < PE; ++j) {
        const int kj = j / G, gj = j % G;
        const double yj = quad_bcast(rown[kj] * a[quad_off(kj) + j], quad_base + gj);
#pragma unroll
        for (int k = 0; k < K; ++k) {
            if (k > kj) rown[k] = fma(-a[quad_off(k) + j], yj, rown[k]);
            if (k == kj && gj < G - 1) rown[k] = fma(g > gj ? -a[quad_off(k) + j] : 0.0, yj, rown[k]);
        }
        if (g == gj) rown[kj] = yj;
    }
    // ---- backward substitution: each lane sums over its own rows, the quad adds up ----
#pragma unroll
    for (int jj = 0; jj < PE; ++jj) {
        const int j = PE - 1 - jj;
        const int kj = j / G, gj = j % G;
        double s = 0.0;
#pragma unroll
        for (int k = 0; k < K; ++k) {
            if (k > kj) s = fma(a[quad_off(k) + j], rown[k], s);
            if (k == kj && gj < G - 1) s = fma(g > gj ? a[quad_off(k) + j] : 0.0, g > gj ? rown[k] : 0.0, s);
        }
        s = quad_sum(s);
        const double xj = (rown[kj] - s) * a[quad_off(kj) + j];
        if (g == gj) rown[kj] = xj;
    }

    // own rows out; a failed pixel stays at the forecast
#pragma unroll
    for (int k = 0; k < K; ++k) {
        const int r = G * k + g;
        if (r < P && live) {
            const double xn = ok ? rown[k] : __ldg(prm.x_f + (int64_t)r * ld + i);
            stg_stream(prm.x_out + (int64_t)r * ld + i, xn);
            const double dlt = xn - xpo[k];
            ssq = fma(dlt, dlt, ssq);
        }
    }
    if (live && g == 0) bad = ok ? 0.0 : 1.0;
    if (prm.fwd) {
        // forward-modelled values H0 + J.(x - x_f) (solvers.py:137), band b by lane b mod 4: the
        // solution is read back from x_out, which the quad has just written
        __syncwarp();
        if (live) {
            for (int b = g; b < prm.n_bands; b += G) {
                const SolveBandDev& bd = prm.bands[b];
                const bool on = bd.w[i] != 0.0;
                double f = bd.kind == KAFKA_BAND_ARRAYS ? bd.h0[i] : (on ? bd.h0[0] : 0.0);
                for (int p = 0; p < P; ++p) {
                    const int pl = bd.src[p];
                    if (pl < 0) continue;
                    const double jv = bd.kind == KAFKA_BAND_ARRAYS ? bd.jac[(int64_t)pl * ld + i] : (on ? bd.jac[pl] : 0.0);
                    f = fma(jv, prm.x_out[(int64_t)p * ld + i] - prm.x_f[(int64_t)p * ld + i], f);
                }
                stg_stream(prm.fwd + (int64_t)b * ld + i, f);
            }
        }
    }

    ssq = warp_sum(ssq);
    bad = warp_sum(bad);
    if ((threadIdx.x & 31) == 0) {
        s_part[0][threadIdx.x >> 5] = ssq;
        s_part[1][threadIdx.x >> 5] = bad;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        double s0 = 0.0, s1 = 0.0;
#pragma unroll
        for (int k = 0; k < kQuadThreads / 32; ++k) {
            s0 += s_part[0][k];
            s1 += s_part[1][k];
        }
        const unsigned n_slots = (gridDim.x + 1) / 2;
        atomicAdd(prm.partials + (blockIdx.x >> 1), s0);
        atomicAdd(prm.partials + n_slots + (blockIdx.x >> 1), s1);
    }
}

}  // namespace kafka
