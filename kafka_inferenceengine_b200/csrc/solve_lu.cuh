// Pixels whose information matrix is not positive definite.
//
// The reference solves the block-diagonal system with a sparse LU (kafka/inference/solvers.py:133-134,
// scipy's spsolve): an indefinite block -- what the second-order correction of linear_kf.py:313-319 can
// leave behind, or a prior that is not a precision matrix -- still gets the solution of its equations.
// The solve kernels factorise by Cholesky and flag such a pixel instead (one bit per pixel in
// SolveParams::fail_bits, the pixel left at the forecast).  The second stage runs right behind them on the same
// stream, finds the set bits and solves those pixels by Gaussian elimination with partial pivoting -- the
// algorithm of a dense LU -- so that they get what the reference gives them: A^-1 b.  A pixel whose block is
// singular to working precision, or holds a NaN, stays at the forecast and stays counted.
//
// It is rare work, laid out for small footprint rather than speed: one warp takes one pixel at a time, the
// augmented matrix [A | b] (at most 16 x 17 doubles) lives in the warp's shared memory (lu.cuh).
// A and b are rebuilt from the inputs with the operations of the solve kernels (the kernels do not keep b,
// and store A only in iterations that can be the last).  The search for flagged pixels rides in the kernel
// that adds up the partial sums anyway (finish_solve_kernel), so a solve without such pixels pays nothing
// for it; every 128-pixel tile belongs to exactly one lane there and pixels are taken in index order, so the
// norm stays bit-reproducible.
#pragma once
#include "lu.cuh"

namespace kafka {

constexpr int kFinishWarps = 8;            // warps per CTA
constexpr int kFinishCtas = 296;           // at most two CTAs per SM; every warp takes 32 tiles per trip

// One pixel: rebuild [A | b] as the solve kernels build it, solve by elimination, store the state (and the
// forward-modelled values).  All 32 lanes call it; returns false when the block is singular or holds a NaN
// (the pixel then stays at the forecast).  dssq = change of the pixel's contribution to sum (x - x_prev)^2.
__device__ __forceinline__ bool solve_one_indefinite(const SolveParams& prm, int P, int64_t i, LuWork<1>& W, int lane,
                                                     double& dssq) {
    const int64_t ld = prm.ld;
    const bool bcast = (prm.flags & KAFKA_PRIOR_BROADCAST) != 0;
    const bool round32 = (prm.flags & KAFKA_ROUND_F32) != 0;
    const int r = lane;                                              // my row
    double xp[KAFKA_MAX_PARAMS], xf[KAFKA_MAX_PARAMS];
#pragma unroll
    for (int p = 0; p < KAFKA_MAX_PARAMS; ++p) {
        xp[p] = p < P ? prm.x_prev[(int64_t)p * ld + i] : 0.0;
        xf[p] = p < P ? prm.x_f[(int64_t)p * ld + i] : 0.0;
    }
    if (r < P)
        for (int q = 0; q <= r; ++q) {
            const double v = bcast ? prm.pinv_f[tri(r, q)] : prm.pinv_f[(int64_t)tri(r, q) * ld + i];
            W.m[r][q] = v;
            W.m[q][r] = v;
        }
    __syncwarp();
    double rhs = 0.0;
    if (r < P) {
#pragma unroll
        for (int q = 0; q < KAFKA_MAX_PARAMS; ++q)
            if (q < P) rhs = fma(W.m[r][q], xf[q], rhs);
    }
    __syncwarp();
    for (int b = 0; b < prm.n_bands; ++b) {
        const SolveBandDev& bd = prm.bands[b];
        const double w = bd.w[i], y = bd.y[i];
        double J[KAFKA_MAX_PARAMS], h0;
        bool dummy = false;
        band_row<KAFKA_MAX_PARAMS>(bd, ld, i, xp, w, J, h0, round32, dummy);
        double yp = y - h0;
#pragma unroll
        for (int p = 0; p < KAFKA_MAX_PARAMS; ++p)
            if (p < P) yp = fma(J[p], xp[p], yp);
        const double wy = w * yp;
        double jr = 0.0;
#pragma unroll
        for (int p = 0; p < KAFKA_MAX_PARAMS; ++p) jr = p == r ? J[p] : jr;
        const double wj = w * jr;
        if (r < P) {
            rhs = fma(wy, jr, rhs);
#pragma unroll
            for (int q = 0; q < KAFKA_MAX_PARAMS; ++q)
                if (q <= r) W.m[r][q] = fma(wj, J[q], W.m[r][q]);
        }
    }
    __syncwarp();
    if (r < P) {
        for (int q = 0; q < r; ++q) W.m[q][r] = W.m[r][q];
        W.m[r][P] = rhs;
    }
    __syncwarp();
    if (!warp_gepp(W, P, 1, lane)) return false;
    double xl[KAFKA_MAX_PARAMS];
#pragma unroll
    for (int p = 0; p < KAFKA_MAX_PARAMS; ++p) xl[p] = p < P ? W.m[p][P] : 0.0;
    double d = 0.0;
    if (r < P) {
        double mine_x = 0.0, mine_f = 0.0, mine_p = 0.0;
#pragma unroll
        for (int p = 0; p < KAFKA_MAX_PARAMS; ++p) {
            mine_x = p == r ? xl[p] : mine_x;
            mine_f = p == r ? xf[p] : mine_f;
            mine_p = p == r ? xp[p] : mine_p;
        }
        prm.x_out[(int64_t)r * ld + i] = mine_x;
        const double dn = mine_x - mine_p, dold = mine_f - mine_p;   // the solve kernel counted the forecast
        d = dn * dn - dold * dold;
    }
    dssq = warp_sum(d);
    if (prm.fwd) {
        for (int b = lane; b < prm.n_bands; b += 32) {
            const SolveBandDev& bd = prm.bands[b];
            double J[KAFKA_MAX_PARAMS], h0;
            bool dummy = false;
            band_row<KAFKA_MAX_PARAMS>(bd, ld, i, xp, bd.w[i], J, h0, round32, dummy);
            double f = h0;
#pragma unroll
            for (int p = 0; p < KAFKA_MAX_PARAMS; ++p)
                if (p < P) f = fma(J[p], xl[p] - xf[p], f);
            prm.fwd[(int64_t)b * ld + i] = f;
        }
    }
    __syncwarp();
    return true;
}

// Second stage of every solve: the flagged pixels, then the sums.  Lane l of a warp takes tile 32 t + l of the
// warp's trip t: its partial sum of squares and its count of pixels left at the forecast.  A tile with a
// non-zero count may hold flagged pixels: the warp reads the tile's four words of failure bits and solves
// those pixels one at a time, correcting the owning lane's two numbers in registers.  Sums: lane -> warp ->
// CTA in a fixed order, one pair per CTA into `cta_part`; the CTA that arrives last adds the pairs up in index
// order and writes the results.  Which tiles a lane meets depends only on n_tiles, so the norm has the same
// bits in every run.
__global__ void __launch_bounds__(kFinishWarps * 32)
finish_solve_kernel(const __grid_constant__ SolveParams prm, int P, unsigned n_tiles, double* __restrict__ cta_part,
                    unsigned* __restrict__ ticket, double* __restrict__ norm2, int32_t* __restrict__ n_bad,
                    bool report_lu) {
    __shared__ LuWork<1> s_work[kFinishWarps];
    __shared__ double s_sum[2][kFinishWarps];
    __shared__ bool s_last;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const unsigned n_warps = gridDim.x * kFinishWarps;
    double ssq = 0.0, bad = 0.0;
    int solved = 0;
    for (unsigned t0 = (blockIdx.x * kFinishWarps + warp) * 32u; t0 < n_tiles; t0 += n_warps * 32u) {
        const unsigned tile = t0 + lane;
        double my_ssq = 0.0, my_bad = 0.0;
        if (tile < n_tiles) {
            my_ssq = prm.partials[tile];
            my_bad = prm.partials[n_tiles + tile];
        }
        unsigned flagged = __ballot_sync(0xffffffffu, my_bad != 0.0);
        while (flagged) {                                            // rare
            const int src = __ffs(flagged) - 1;
            flagged &= flagged - 1;
            const unsigned ft = t0 + src;
            unsigned bits4 = lane < 4 ? prm.fail_bits[4 * ft + lane] : 0u;
            for (int wd = 0; wd < 4; ++wd) {
                unsigned bits = __shfl_sync(0xffffffffu, bits4, wd);
                while (bits) {
                    const int bit = __ffs(bits) - 1;
                    bits &= bits - 1;
                    const int64_t i = ((int64_t)ft * 4 + wd) * 32 + bit;
                    double dssq = 0.0;
                    if (solve_one_indefinite(prm, P, i, s_work[warp], lane, dssq)) {
                        if (lane == src) {
                            my_ssq += dssq;
                            my_bad -= 1.0;
                        }
                        ++solved;
                    }
                }
            }
        }
        ssq += my_ssq;
        bad += my_bad;
    }
    if (lane == 0 && solved) atomicAdd(prm.n_lu, solved);
    ssq = warp_sum(ssq);
    bad = warp_sum(bad);
    if (lane == 0) {
        s_sum[0][warp] = ssq;
        s_sum[1][warp] = bad;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        double a = 0.0, b = 0.0;
#pragma unroll
        for (int k = 0; k < kFinishWarps; ++k) {
            a += s_sum[0][k];
            b += s_sum[1][k];
        }
        cta_part[blockIdx.x] = a;
        cta_part[gridDim.x + blockIdx.x] = b;
        __threadfence();
        s_last = atomicAdd(ticket, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    if (warp == 0) {                                                 // at most kFinishCtas pairs: one warp, fixed order
        double a = 0.0, b = 0.0;
        for (unsigned k = lane; k < gridDim.x; k += 32) {
            a += __ldcg(cta_part + k);
            b += __ldcg(cta_part + gridDim.x + k);
        }
        a = warp_sum(a);
        b = warp_sum(b);
        if (lane == 0) {
            if (norm2) *norm2 = a;
            if (n_bad) {
                n_bad[0] = (int32_t)b;
                if (report_lu) n_bad[1] = __ldcg(prm.n_lu);
            }
            *ticket = 0u;
        }
    }
}

}  // namespace kafka
