// Two lanes per pixel: the update of solve_bands_kernel for the largest blocks (n_params 14-16:
// 105-136 packed entries + the vectors do not fit one thread's registers, and spilling them or
// parking rows in shared memory costs more than sharing the block).  Measured on B200, 4 M px,
// supplied Jacobians, % of the HBM peak at 1 / 10 bands (bench.py --workload c5):
//     n_params        13        14        15        16
//     this kernel   57 / 61   60 / 67   60 / 67   61 / 68
//     one thread    75 / 81   63 / 55   49 / 44   33 / 29 (spills)
// so the pair kernel is the default from 14 and selectable at any size (KAFKA_SOLVE_PAIR, used by
// the tests).  ncu: 254 registers, 8 warps per SM, stalls on dependent-instruction latency (wait,
// scoreboards) -- two warps per scheduler cannot hide the factorisation's chain.
//
// Lane g (0/1) of a pair owns the rows r = 2k + g of the packed lower triangle, k = 0..K-1,
// K = ceil(P/2); row k of a lane sits at a[k(k+1) .. k(k+1)+2k+1] (lane 0's last slot of every
// row is unused, which keeps every register index a literal for both lanes).  An odd P is
// padded with a decoupled unit row.  Everything that crosses rows goes through shuffles
// inside the pair:
//   * right-looking Cholesky: column j is scaled by its owners and broadcast with
//     __shfl_sync(.., pair_base + row parity), then both lanes update their own rows --
//     per entry the same operations in the same order as the one-thread kernel;
//   * forward substitution: y_j broadcast from its owner; backward: each lane sums over its
//     own rows and one __shfl_xor adds the partner's share.
// Band values are staged per PIXEL by cp.async (the two lanes copy alternate values) and
// read by both lanes after __syncwarp.  128 threads = 64 pixels per CTA (see kPairPix); scratch and the
// deterministic reduction are those of solve_bands_kernel.
#pragma once

namespace kafka {

#ifndef KAFKA_PAIR_FROM_PARAMS
#define KAFKA_PAIR_FROM_PARAMS 14
#endif
constexpr int kPairFromParams = KAFKA_PAIR_FROM_PARAMS;   // state sizes from here on use the pair kernel
// 64 pixels (128 threads) per CTA: at 254 registers that is two CTAs per SM whose load phases (prior block,
// band rows) and compute phases (factorisation, substitutions) overlap each other; with one 128-pixel CTA
// per SM every warp waited on memory at the same time and the HBM pipe idled during the factorisation.
// CTAs 2k and 2k+1 add their partial sums into slot k of the scratch (zeroed by the host before the
// launch): the sum of two numbers does not depend on their order, so the norm stays bit-reproducible.
#ifndef KAFKA_PAIR_SPLIT
#define KAFKA_PAIR_SPLIT 2
#endif
constexpr int kPairSplit = KAFKA_PAIR_SPLIT;                  // CTAs per 128-pixel tile (experiment knob)
constexpr int kPairPix = kSolveThreads / kPairSplit;
constexpr int kPairThreads = 2 * kPairPix;

template <int P>
__device__ __forceinline__ double at_or_zero(const double (&v)[P], int idx) {
    return idx < P ? v[idx < P ? idx : 0] : 0.0;
}

__device__ __forceinline__ double pair_bcast(double v, unsigned lane) { return __shfl_sync(0xffffffffu, v, lane); }

// kAllIdentity: every band is a per-pixel array band with the identity mapper (see solve_bands_kernel).
template <int P, bool kAllIdentity = false>
__global__ void __launch_bounds__(kPairThreads, kPairSplit)
solve_bands_pair_kernel(const __grid_constant__ SolveParams prm, int n_dim_max) {
    constexpr int K = (P + 1) / 2;
    constexpr int PE = 2 * K;
    constexpr int LA = K * (K + 1);
    extern __shared__ __align__(16) double stages[];               // kAsyncStages x (3 + n_dim_max) x kPairPix
    __shared__ double s_part[2][kPairThreads / 32];
    __shared__ __align__(16) double s_uni[KAFKA_MAX_BANDS][P + 2];
    const int g = threadIdx.x & 1;
    const int slot = threadIdx.x >> 1;
    const unsigned pair_base = (threadIdx.x & 31u) & ~1u;
    const int64_t tile0 = (int64_t)blockIdx.x * kPairPix;
    const int64_t i = tile0 + slot;
    const int64_t ld = prm.ld;
    const int stage_doubles = (3 + n_dim_max) * kPairPix;
    const bool live = i < prm.n_pix;
    const bool in_ld = i < ld;
    double ssq = 0.0, bad = 0.0;

    // value v of a band sits at stage[v * kPairPix + slot]: 0 = w, 1 = y, 2 = H0, 3.. = Jacobian planes;
    // lane g copies the values with v % 2 == g: lane 0 takes w, H0 and the odd planes, lane 1 takes y and
    // the even planes.  Written with selects, not as one loop over v: that loop compiled to a jump table
    // on which the two lanes of every pair diverged (ncu: 40 % of the kernel's stall samples).
    auto issue = [&](int b) {
        if (b < prm.n_bands && in_ld) {
            const SolveBandDev& bd = prm.bands[b];
            double* st = stages + (b % kAsyncStages) * stage_doubles + slot;
            cp_async8(st + g * kPairPix, (g ? bd.y : bd.w) + i);
            if (kAllIdentity || bd.kind == KAFKA_BAND_ARRAYS) {
                cp_async8(st + (2 + g) * kPairPix, (g ? bd.jac : bd.h0) + i);
                // running pointers and a predicated, fully unrolled loop: no 64-bit multiply and no branch per copy
                const double* src = bd.jac + (int64_t)(1 + g) * ld + i;
                double* dst = st + (4 + g) * kPairPix;
                const int n_mine = ((kAllIdentity ? P : bd.n_planes) - g) / 2;    // planes 1+g, 3+g, ...
#pragma unroll
                for (int t = 0; t < KAFKA_MAX_DIM / 2; ++t, src += 2 * ld, dst += 2 * kPairPix)
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
        for (int k = threadIdx.x; k < prm.n_bands * (P + 1); k += kPairThreads) {
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
    const bool round32 = (prm.flags & KAFKA_ROUND_F32) != 0;
    bool domain_bad = false;
    double a[LA], rown[K], xp[P];
    {   // packed entry c = r(r+1)/2 + q of row r = 2k + g; row starts are 4k + 2g + 3 planes apart, so the
        // address is a running pointer (64-bit adds) instead of a 64-bit multiply per entry
        const int64_t step = bcast ? 1 : ld;
        const double* rowp = prm.pinv_f + (bcast ? 0 : il) + (int64_t)g * step;      // row g starts at c = g
        const int64_t g2 = 2 * (int64_t)g * step;
#pragma unroll
        for (int k = 0; k < K; ++k) {
            const double* ep = rowp;
#pragma unroll
            for (int q = 0; q < 2 * k + 2; ++q, ep += step) {
                const int r = 2 * k + g;
                const bool valid = q <= r && r < P;
                double v = (q == r) ? 1.0 : 0.0;                    // padding row: decoupled unit block
                if (valid) v = bcast ? *ep : ldg_stream(ep);
                a[k * (k + 1) + q] = v;
            }
            rowp += (4 * k + 3) * step + g2;
        }
    }
#pragma unroll
    for (int p = 0; p < P; ++p) xp[p] = ldg_stream(prm.x_prev + (int64_t)p * ld + il);
    {   // rhs = P_f^-1 x_f: every lane spreads its own entries over a full-length vector, the pair adds up
        double t[PE];
#pragma unroll
        for (int p = 0; p < PE; ++p) t[p] = 0.0;
        double xf[PE];
#pragma unroll
        for (int p = 0; p < PE; ++p) xf[p] = p < P ? __ldg(prm.x_f + (int64_t)(p < P ? p : 0) * ld + il) : 0.0;
#pragma unroll
        for (int k = 0; k < K; ++k) {
            const double xr = g ? xf[2 * k + 1] : xf[2 * k];
            double acc = 0.0;
#pragma unroll
            for (int q = 0; q < 2 * k + 2; ++q) {
                const bool lower = q < 2 * k + g;                   // strictly below the diagonal of row 2k+g
                const double v = a[k * (k + 1) + q];
                acc = fma(v, (q <= 2 * k + g) ? xf[q] : 0.0, acc);  // row r: sum_{q<=r} A(r,q) x_q
                t[q] = fma(lower ? v : 0.0, xr, t[q]);              // column part: A(r,q) x_r goes to row q
            }
            // acc belongs to row 2k+g
            t[2 * k] += g ? 0.0 : acc;
            t[2 * k + 1] += g ? acc : 0.0;
        }
#pragma unroll
        for (int k = 0; k < K; ++k) {
            const double e = t[2 * k] + __shfl_xor_sync(0xffffffffu, t[2 * k], 1);
            const double o = t[2 * k + 1] + __shfl_xor_sync(0xffffffffu, t[2 * k + 1], 1);
            rown[k] = g ? o : e;
        }
    }

    for (int b = 0; b < prm.n_bands; ++b) {
        issue(b + kAsyncStages - 1);
        cp_async_wait<kAsyncStages - 1>();
        __syncwarp();                                               // the partner's copies of band b are visible too
        const SolveBandDev& bd = prm.bands[b];
        const double* st = stages + (b % kAsyncStages) * stage_doubles;
        const double w = in_ld ? st[slot] : 0.0;
        const double y = in_ld ? st[kPairPix + slot] : 0.0;
        double J[P], h0;
        if (kAllIdentity || bd.kind == KAFKA_BAND_ARRAYS) {
            if (kAllIdentity || bd.identity) {
#pragma unroll
                for (int p = 0; p < P; ++p) J[p] = st[(3 + p) * kPairPix + slot];
            } else {
#pragma unroll
                for (int p = 0; p < P; ++p) J[p] = bd.src[p] >= 0 ? st[(3 + bd.src[p]) * kPairPix + slot] : 0.0;
            }
            h0 = st[2 * kPairPix + slot];
        } else if (bd.kind == KAFKA_BAND_UNIFORM) {
            const bool on = w != 0.0;
#pragma unroll
            for (int p = 0; p < P; ++p) J[p] = on ? s_uni[b][p] : 0.0;
            h0 = on ? s_uni[b][P] : 0.0;
        } else {
            band_row<P>(bd, 0, 0, xp, w, J, h0, round32, domain_bad);
        }
        double yp = y - h0;
#pragma unroll
        for (int p = 0; p < P; ++p) yp = fma(J[p], xp[p], yp);
        const double wy = w * yp;
#pragma unroll
        for (int k = 0; k < K; ++k) {
            const double jr = g ? at_or_zero<P>(J, 2 * k + 1) : J[2 * k];
            const double wj = w * jr;
            rown[k] = fma(wy, jr, rown[k]);
#pragma unroll
            for (int q = 0; q < 2 * k + 2; ++q)
                a[k * (k + 1) + q] = fma(wj, at_or_zero<P>(J, q), a[k * (k + 1) + q]);
        }
        if (prm.innovations && live && (b & 1) == g) stg_stream(prm.innovations + (int64_t)b * ld + i, y - h0);
        __syncwarp();                                               // both lanes are done with this stage
    }
    cp_async_wait<0>();

    if (live && prm.a_out) {
        double* rowp = prm.a_out + i + (int64_t)g * ld;                              // same walk as the prior load
        const int64_t g2 = 2 * (int64_t)g * ld;
#pragma unroll
        for (int k = 0; k < K; ++k) {
            double* ep = rowp;
#pragma unroll
            for (int q = 0; q < 2 * k + 2; ++q, ep += ld) {
                const int r = 2 * k + g;
                if (q <= r && r < P) stg_stream(ep, a[k * (k + 1) + q]);
            }
            rowp += (4 * k + 3) * ld + g2;
        }
    }

    // ---- Cholesky, right-looking; 1/L_jj is left on the diagonal ----
    bool ok = true;
#pragma unroll
    for (int j = 0; j < PE; ++j) {
        const int kj = j / 2, gj = j % 2;
        const double d = pair_bcast(a[kj * (kj + 1) + j], pair_base + gj);
        ok = ok && (d > 0.0);
        const double inv = rsqrt(d);
#pragma unroll
        for (int k = 0; k < K; ++k)
            if (k >= kj) a[k * (k + 1) + j] *= inv;                 // rows > j: L(r,j); row j itself: L(j,j), replaced next
        if (g == gj) a[kj * (kj + 1) + j] = inv;
        double col[PE];
#pragma unroll
        for (int r = 0; r < PE; ++r)
            if (r > j) col[r] = pair_bcast(a[(r / 2) * (r / 2 + 1) + j], pair_base + (r % 2));
#pragma unroll
        for (int k = 0; k < K; ++k)
            if (k > kj || (k == kj && gj == 0)) {
#pragma unroll
                for (int q = 0; q < 2 * k + 2; ++q)
                    if (q > j) a[k * (k + 1) + q] = fma(-a[k * (k + 1) + j], col[q], a[k * (k + 1) + q]);
            }
    }
    const bool indefinite = live && !ok && !domain_bad;             // solve_indefinite_kernel takes this pixel over
    ok = ok && !domain_bad;

    // ---- forward substitution: y_j broadcast from its owner ----
#pragma unroll
    for (int j = 0; j < PE; ++j) {
        const int kj = j / 2, gj = j % 2;
        const double yj = pair_bcast(rown[kj] * a[kj * (kj + 1) + j], pair_base + gj);
#pragma unroll
        for (int k = 0; k < K; ++k)
            if (k > kj || (k == kj && gj == 0)) rown[k] = fma(-a[k * (k + 1) + j], yj, rown[k]);
        if (g == gj) rown[kj] = yj;
    }
    // ---- backward substitution: each lane sums over its own rows, the pair adds up ----
#pragma unroll
    for (int jj = 0; jj < PE; ++jj) {
        const int j = PE - 1 - jj;
        const int kj = j / 2, gj = j % 2;
        double s = 0.0;
#pragma unroll
        for (int k = 0; k < K; ++k) {
            if (k > kj) s = fma(a[k * (k + 1) + j], rown[k], s);
            if (k == kj && gj == 0) s = fma(g ? a[k * (k + 1) + j] : 0.0, g ? rown[k] : 0.0, s);   // row j+1, lane 1 only
        }
        s += __shfl_xor_sync(0xffffffffu, s, 1);
        const double xj = (rown[kj] - s) * a[kj * (kj + 1) + j];
        if (g == gj) rown[kj] = xj;
    }

    if (prm.fwd || live) {
        // own rows out; a failed pixel stays at the forecast
#pragma unroll
        for (int k = 0; k < K; ++k) {
            const int r = 2 * k + g;
            if (r < P && live) {
                const double xn = ok ? rown[k] : __ldg(prm.x_f + (int64_t)r * ld + i);
                rown[k] = xn;
                stg_stream(prm.x_out + (int64_t)r * ld + i, xn);
                const double dlt = xn - (g ? at_or_zero<P>(xp, 2 * k + 1) : xp[2 * k]);
                ssq = fma(dlt, dlt, ssq);
            }
        }
        if (live && g == 0) bad = ok ? 0.0 : 1.0;
        if (prm.fwd) {                                              // all lanes of the warp take part in the shuffles
            double dx[P];
#pragma unroll
            for (int p = 0; p < P; ++p) {
                const double xv = pair_bcast(rown[p / 2], pair_base + (p % 2));
                dx[p] = live ? xv - __ldg(prm.x_f + (int64_t)p * ld + i) : 0.0;
            }
            if (live) {
                for (int b = g; b < prm.n_bands; b += 2) {
                    const SolveBandDev& bd = prm.bands[b];
                    double J[P], h0;
                    bool dummy = false;
                    band_row<P>(bd, ld, i, xp, bd.w[i], J, h0, round32, dummy);
                    double f = h0;
#pragma unroll
                    for (int p = 0; p < P; ++p) f = fma(J[p], dx[p], f);
                    stg_stream(prm.fwd + (int64_t)b * ld + i, f);
                }
            }
        }
    }

    ssq = warp_sum(ssq);
    bad = warp_sum(bad);
    {   // one bit per pixel: the even lanes' verdicts, packed into the half word of this warp's 16 pixels
        // (the words were zeroed by the host before the launch, so only a failure writes)
        unsigned m = __ballot_sync(0xffffffffu, indefinite && g == 0);
        if (m != 0 && (threadIdx.x & 31) == 0) {
            m = (m | (m >> 1)) & 0x33333333u;
            m = (m | (m >> 2)) & 0x0f0f0f0fu;
            m = (m | (m >> 4)) & 0x00ff00ffu;
            m = (m | (m >> 8)) & 0x0000ffffu;
            const int64_t first = tile0 + (threadIdx.x >> 5) * 16;  // first pixel of this warp
            atomicOr(prm.fail_bits + (first >> 5), m << (first & 31));
        }
    }
    if ((threadIdx.x & 31) == 0) {
        s_part[0][threadIdx.x >> 5] = ssq;
        s_part[1][threadIdx.x >> 5] = bad;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        double s0 = 0.0, s1 = 0.0;
#pragma unroll
        for (int k = 0; k < kPairThreads / 32; ++k) {
            s0 += s_part[0][k];
            s1 += s_part[1][k];
        }
        const unsigned n_slots = (gridDim.x + kPairSplit - 1) / kPairSplit;
        atomicAdd(prm.partials + blockIdx.x / kPairSplit, s0);
        atomicAdd(prm.partials + n_slots + blockIdx.x / kPairSplit, s1);
    }
}

}  // namespace kafka
