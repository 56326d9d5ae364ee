// Fused normal-equation build + per-pixel fp64 Cholesky solve + convergence
// partial norm (sm_100a).  Replaces sort_band_data + variational_kalman_multiband
// (kafka/inference/solvers.py:81-145) and the norm of do_all_bands
// (kafka/linear_kf.py:293).
//
// One thread owns one pixel: the packed P x P block (P(P+1)/2 fp64) and the
// right-hand side live in registers, every global access is a coalesced planar
// access (consecutive threads = consecutive pixels), nothing is re-read.  The
// band rows are prefetched through a thread-private cp.async ring (see the
// kernel).  This kernel is HBM-bound; algorithmic bytes per pixel are in DESIGN.md.
#include <stdlib.h>

#include <type_traits>

#include "chol.cuh"

namespace kafka {

struct SolveBandDev {
    const double* h0;
    const double* jac;
    const double* y;
    const double* w;
    double row[KAFKA_MAX_PARAMS];
    double offset;
    int32_t kind;
    int32_t n_planes;               // ARRAYS: Jacobian planes of this band
    int8_t src[KAFKA_MAX_PARAMS];   // Jacobian plane feeding state slot p, -1 = none
    int8_t identity;                // ARRAYS: plane p feeds slot p for every p (n_planes == n_params)
};

struct SolveParams {
    SolveBandDev bands[KAFKA_MAX_BANDS];
    const double* x_prev;
    const double* x_f;
    const double* pinv_f;
    double* x_out;
    double* a_out;
    double* innovations;
    double* fwd;
    double* partials;     // [2][tiles]: sum of squares, pixels left at the forecast (tile = 128 pixels)
    uint32_t* fail_bits;  // one bit per pixel: Cholesky met a non-positive pivot (finish_solve_kernel takes over)
    int32_t* n_lu;        // [0] pixels solved by elimination, [1] arrival ticket of finish_solve_kernel
    int64_t n_pix;
    int64_t ld;
    int32_t n_bands;
    uint32_t flags;
};

#ifndef KAFKA_SOLVE_THREADS
#define KAFKA_SOLVE_THREADS 128
#endif
constexpr int kSolveThreads = KAFKA_SOLVE_THREADS;

// Water-Cloud model and its gradient (sar_forward_model.py:74-98); prm = A,B,C,D,E,cos(theta)
__device__ __forceinline__ void wcm_eval(const double* __restrict__ prm, double V, double SM, double& s0, double& g0,
                                         double& g1) {
    const double A = prm[0], Bc = prm[1], C = prm[2], Dc = prm[3], E = prm[4], mu = prm[5];
    const double tau = exp(-2.0 * Bc / mu * V);
    double z = pow(V, E), z1 = pow(V, E - 1.0);
    z = (z != z) ? 1.0 : z;                                 // the reference's NaN guards (:87-92)
    z1 = (z1 != z1) ? 1.0 : z1;
    const double surf = exp10((C + Dc * SM) / 10.0);
    s0 = A * z * mu * (1.0 - tau) + tau * surf;
    g0 = A * E * mu * z1 * (1.0 - tau) + 2.0 * A * Bc * z * tau - 2.0 * surf * Bc * tau / mu;
    g1 = Dc * 2.302585092994045684 * tau * surf * 0.1;
}

struct WcmParams {
    double prm[6];
};

__global__ void __launch_bounds__(256)
wcm_forward_kernel(const double* __restrict__ x, int64_t ld, int64_t n, const __grid_constant__ WcmParams wp,
                   double* __restrict__ sigma0, double* __restrict__ grad) {
    const int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x;
    if (i >= n) return;
    double s0, g0, g1;
    wcm_eval(wp.prm, x[i], x[ld + i], s0, g0, g1);
    sigma0[i] = s0;
    if (grad) {
        grad[i] = g0;
        grad[ld + i] = g1;
    }
}

template <int P>
__device__ __forceinline__ void band_row(const SolveBandDev& bd, int64_t ld, int64_t i, const double (&xp)[P],
                                         double w, double (&J)[P], double& h0, bool round32, bool& domain_bad) {
    if (bd.kind == KAFKA_BAND_LINEAR) {
        double s = bd.offset;
#pragma unroll
        for (int p = 0; p < P; ++p) {
            J[p] = w != 0.0 ? bd.row[p] : 0.0;
            s = fma(bd.row[p], xp[p], s);
        }
        h0 = w != 0.0 ? s : 0.0;
    } else if (bd.kind == KAFKA_BAND_WCM) {
        // Water-Cloud model (sar_forward_model.py:74-98): V = vegetation descriptor, SM = soil moisture
        double V = 0.0, SM = 0.0;
#pragma unroll
        for (int p = 0; p < P; ++p) {                       // static indexing keeps xp in registers
            V = bd.src[p] == 0 ? xp[p] : V;
            SM = bd.src[p] == 1 ? xp[p] : SM;
        }
        double s0, g0, g1;
        wcm_eval(bd.row, V, SM, s0, g0, g1);
        if (round32) {
            s0 = round_f32(s0);
            g0 = round_f32(g0);
            g1 = round_f32(g1);
        }
        const bool on = w != 0.0;
#pragma unroll
        for (int p = 0; p < P; ++p) J[p] = !on ? 0.0 : (bd.src[p] == 0 ? g0 : (bd.src[p] == 1 ? g1 : 0.0));
        h0 = on ? s0 : 0.0;
        if (on && !(V > 0.0 && SM > 0.0)) domain_bad = true;  // "Negative LAI!" / "Negative SM!" (:66-71)
    } else if (bd.kind == KAFKA_BAND_UNIFORM) {
        // one linearisation point for the whole image: every thread reads the same
        // D+1 values (L1 broadcast); unobserved pixels keep an empty row (utils.py:199-215)
#pragma unroll
        for (int p = 0; p < P; ++p) J[p] = (w != 0.0 && bd.src[p] >= 0) ? __ldg(bd.jac + bd.src[p]) : 0.0;
        h0 = w != 0.0 ? __ldg(bd.h0) : 0.0;
    } else {
#pragma unroll
        for (int p = 0; p < P; ++p) J[p] = bd.src[p] >= 0 ? ldg_stream(bd.jac + (int64_t)bd.src[p] * ld + i) : 0.0;
        h0 = ldg_stream(bd.h0 + i);
    }
}

// kMode: 0 = any mix of band kinds; 1 = every band a per-pixel array band with the identity mapper;
// 2 = every band shares one linearisation (KAFKA_BAND_UNIFORM).  1 and 2 are chosen by the host and drop the
// band-kind dispatch, the plane-count predicates and the mapper look-ups from the band loop.
template <int P, int kMode = 0>
__device__ __forceinline__ void band_row_staged(const SolveBandDev& bd, const double* __restrict__ st, int n_dim_max,
                                                const double* __restrict__ uni, const double (&xp)[P], double w,
                                                double (&J)[P], double& h0, bool round32, bool& domain_bad) {
    // stage layout: [w | y | h0 | jac plane 0 .. n_dim-1], kSolveThreads doubles each
    constexpr bool kAllIdentity = kMode == 1;
    if (kMode != 2 && (kAllIdentity || bd.kind == KAFKA_BAND_ARRAYS)) {
        if (kAllIdentity || bd.identity) {                                  // the usual case: literal offsets, no mapper look-ups
#pragma unroll
            for (int p = 0; p < P; ++p) J[p] = st[(3 + p) * kSolveThreads + threadIdx.x];
        } else {
#pragma unroll
            for (int p = 0; p < P; ++p)
                J[p] = bd.src[p] >= 0 ? st[(3 + bd.src[p]) * kSolveThreads + threadIdx.x] : 0.0;
        }
        h0 = st[2 * kSolveThreads + threadIdx.x];
    } else if (kMode == 2 || bd.kind == KAFKA_BAND_UNIFORM) {
        // one linearisation for the whole image, already laid out by state slot in shared memory
        // (uni = {J[0..P-1], h0}): warp-wide broadcast reads, no branch; unobserved pixels keep
        // an empty row (utils.py:199-215)
        const bool on = w != 0.0;
#pragma unroll
        for (int p = 0; p < P; ++p) J[p] = on ? uni[p] : 0.0;
        h0 = on ? uni[P] : 0.0;
    } else {
        band_row<P>(bd, 0, 0, xp, w, J, h0, round32, domain_bad);     // LINEAR / UNIFORM / WCM: nothing per pixel to load
    }
}

// ---------------------------------------------------------------------------
// Thread-private asynchronous pipeline.  Every thread copies ITS OWN pixel's band
// values (w, y, H0, J planes: 3+D doubles) into its own shared-memory slots with
// cp.async (LDGSTS, no registers held while in flight), kAsyncStages-1 bands ahead of the
// arithmetic, and reads them back after cp.async.wait_group: no thread ever reads
// what another thread copied, so the band loop needs no barrier of any kind.
// ---------------------------------------------------------------------------
#ifndef KAFKA_ASYNC_STAGES
#define KAFKA_ASYNC_STAGES 3
#endif
constexpr int kAsyncStages = KAFKA_ASYNC_STAGES;

// Packed entries kept in registers: all of them up to n_params 12 (78 entries).  For 13..15
// the first 124 - 4 n_params stay in registers and the remaining rows live in the thread's
// own shared-memory column (measured: 13 -> 71-76 %, 14 -> 55-63 %, 15 -> 44-49 % of the HBM peak,
// against 70 / 55 / 37 % with everything in registers and spilling).  From 14 the default is the
// two-lanes-per-pixel kernel of solve_pair.cuh (59-64 %); the split stays for 13 and for callers
// that pass sizes 14-15 without the pair kernel compiled in (KAFKA_PAIR_FROM_PARAMS).
template <int P>
struct SolveBlock {
    static constexpr int L = tri_count(P);
    static constexpr int R = (P >= 13 && P <= 15) ? 124 - 4 * P : L;
    using type = typename std::conditional<R == L, RegBlock<L>, SplitBlock<L, R, kSolveThreads>>::type;
};

// kMode 1: every band is a per-pixel array band whose P Jacobian planes map one-to-one onto the state (the
// emulator-in-the-loop and supplied-Jacobian cases); kMode 2: every band shares one linearisation (the first
// iteration after a prior reset); see band_row_staged.
template <int P, int kMode = 0>
__global__ void __launch_bounds__(kSolveThreads)
solve_bands_kernel(const __grid_constant__ SolveParams prm, int n_dim_max) {
    constexpr bool kAllIdentity = kMode == 1;
    constexpr int L = tri_count(P);
    constexpr int R = SolveBlock<P>::R;
    extern __shared__ __align__(16) double stages[];               // kAsyncStages x (3 + n_dim_max) x 128 [+ (L-R) x 128]
    __shared__ double s_part[2][kSolveThreads / 32];
    __shared__ __align__(16) double s_uni[KAFKA_MAX_BANDS][P + 2];   // UNIFORM bands: Jacobian by state slot, then H0
    const int64_t tile0 = (int64_t)blockIdx.x * kSolveThreads;
    const int64_t i = tile0 + threadIdx.x;
    const int64_t ld = prm.ld;
    const int stage_doubles = (3 + n_dim_max) * kSolveThreads;
    const bool live = i < prm.n_pix;
    const bool in_ld = i < ld;                                      // padding lanes inside the arrays still copy
    double ssq = 0.0, bad = 0.0;

    auto issue = [&](int b) {
        if (b < prm.n_bands && in_ld) {
            const SolveBandDev& bd = prm.bands[b];
            double* st = stages + (b % kAsyncStages) * stage_doubles + threadIdx.x;
            cp_async8(st, bd.w + i);
            cp_async8(st + kSolveThreads, bd.y + i);
            if (kMode != 2 && (kAllIdentity || bd.kind == KAFKA_BAND_ARRAYS)) {
                cp_async8(st + 2 * kSolveThreads, bd.h0 + i);
                // running pointers and a predicated, fully unrolled loop (a band has at most P planes): the
                // rolled loop cost a 64-bit multiply and a branch per copy (ncu: 17 % of the stall samples)
                const double* src = bd.jac + i;
                double* dst = st + 3 * kSolveThreads;
#pragma unroll
                for (int d = 0; d < P; ++d, src += ld, dst += kSolveThreads)
                    if (kAllIdentity || d < bd.n_planes) cp_async8(dst, src);
            }
        }
        cp_async_commit();                                          // one group per band, even when empty
    };
#pragma unroll
    for (int b = 0; b < kAsyncStages - 1; ++b) issue(b);

    bool any_uniform = kMode == 2;
    if (kMode == 0)
        for (int b = 0; b < prm.n_bands; ++b) any_uniform |= prm.bands[b].kind == KAFKA_BAND_UNIFORM;
    if (any_uniform) {                                              // same decision in every thread
        for (int k = threadIdx.x; k < prm.n_bands * (P + 1); k += kSolveThreads) {
            const int b = k / (P + 1), p = k - b * (P + 1);
            const SolveBandDev& bd = prm.bands[b];
            double v = 0.0;
            if (bd.kind == KAFKA_BAND_UNIFORM) v = p == P ? bd.h0[0] : (bd.src[p] >= 0 ? bd.jac[bd.src[p]] : 0.0);
            s_uni[b][p] = v;
        }
        __syncthreads();
    }

    const int64_t il = live ? i : tile0;
    typename SolveBlock<P>::type A;
    if constexpr (R < L) A.sm = stages + kAsyncStages * stage_doubles + threadIdx.x;
    double rhs[P], xp[P];
    const bool bcast = (prm.flags & KAFKA_PRIOR_BROADCAST) != 0;
    const bool round32 = (prm.flags & KAFKA_ROUND_F32) != 0;
    bool domain_bad = false, indefinite = false;
#pragma unroll
    for (int c = 0; c < L; ++c) A.set(c, bcast ? prm.pinv_f[c] : ldg_stream(prm.pinv_f + (int64_t)c * ld + il));
#pragma unroll
    for (int p = 0; p < P; ++p) {
        xp[p] = ldg_stream(prm.x_prev + (int64_t)p * ld + il);
        rhs[p] = 0.0;
    }
    {   // rhs = P_f^-1 x_f, one column at a time: x_f is not kept (it is re-read, from L2 at
        // best, only by the rare failed pixel and by the optional forward output)
#pragma unroll
        for (int q = 0; q < P; ++q) {
            const double xq = __ldg(prm.x_f + (int64_t)q * ld + il);
#pragma unroll
            for (int r = 0; r < P; ++r) rhs[r] = fma(A.get(r >= q ? tri(r, q) : tri(q, r)), xq, rhs[r]);
        }
    }

    for (int b = 0; b < prm.n_bands; ++b) {
        issue(b + kAsyncStages - 1);
        cp_async_wait<kAsyncStages - 1>();                          // this thread's copies of band b have landed
        const SolveBandDev& bd = prm.bands[b];
        const double* st = stages + (b % kAsyncStages) * stage_doubles;
        const double w = in_ld ? st[threadIdx.x] : 0.0;
        const double y = in_ld ? st[kSolveThreads + threadIdx.x] : 0.0;
        double J[P], h0;
        band_row_staged<P, kMode>(bd, st, n_dim_max, s_uni[b], xp, w, J, h0, round32, domain_bad);
        double yp = y - h0;
#pragma unroll
        for (int p = 0; p < P; ++p) yp = fma(J[p], xp[p], yp);
        const double wy = w * yp;
#pragma unroll
        for (int r = 0; r < P; ++r) {
            const double wj = w * J[r];
            rhs[r] = fma(wy, J[r], rhs[r]);
#pragma unroll
            for (int q = 0; q <= r; ++q) A.set(tri(r, q), fma(wj, J[q], A.get(tri(r, q))));
        }
        if (prm.innovations && live) stg_stream(prm.innovations + (int64_t)b * ld + i, y - h0);
    }
    cp_async_wait<0>();

    if (live) {
        if (prm.a_out) {
#pragma unroll
            for (int c = 0; c < L; ++c) stg_stream(prm.a_out + (int64_t)c * ld + i, A.get(c));
        }
        double invd[P];
        const bool spd = chol_factor_blk<P>(A, invd);
        const bool ok = spd && !domain_bad;
        indefinite = !spd && !domain_bad;
        chol_solve_blk<P>(A, invd, rhs);
#pragma unroll
        for (int p = 0; p < P; ++p) {
            const double xn = ok ? rhs[p] : __ldg(prm.x_f + (int64_t)p * ld + i);
            rhs[p] = xn;
            stg_stream(prm.x_out + (int64_t)p * ld + i, xn);
            const double d = xn - xp[p];
            ssq = fma(d, d, ssq);
        }
        bad = ok ? 0.0 : 1.0;
        if (prm.fwd) {
            for (int b = 0; b < prm.n_bands; ++b) {
                const SolveBandDev& bd = prm.bands[b];
                double J[P], h0;
                bool dummy = false;
                band_row<P>(bd, ld, i, xp, bd.w[i], J, h0, round32, dummy);
                double f = h0;
#pragma unroll
                for (int p = 0; p < P; ++p) f = fma(J[p], rhs[p] - __ldg(prm.x_f + (int64_t)p * ld + i), f);
                stg_stream(prm.fwd + (int64_t)b * ld + i, f);
            }
        }
    }

    ssq = warp_sum(ssq);
    bad = warp_sum(bad);
    const unsigned fail = __ballot_sync(0xffffffffu, indefinite);
    if ((threadIdx.x & 31) == 0) {
        s_part[0][threadIdx.x >> 5] = ssq;
        s_part[1][threadIdx.x >> 5] = bad;
        prm.fail_bits[blockIdx.x * (kSolveThreads / 32) + (threadIdx.x >> 5)] = fail;   // word w = pixels 32w .. 32w+31
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        double s0 = 0.0, s1 = 0.0;
#pragma unroll
        for (int k = 0; k < kSolveThreads / 32; ++k) {
            s0 += s_part[0][k];
            s1 += s_part[1][k];
        }
        prm.partials[blockIdx.x] = s0;
        prm.partials[gridDim.x + blockIdx.x] = s1;
        if (blockIdx.x == 0) prm.n_lu[0] = prm.n_lu[1] = 0;           // counter and ticket of the second stage
    }
}

}  // namespace kafka
#include "solve_pair.cuh"
#include "solve_lu.cuh"
namespace kafka {

// blend_prior (kf_tools.py:75-96): C = Pf + Cp, b = Pf mu_p + Cp x_f, x = C^-1 b.  The reference solves with a
// sparse LU (kf_tools.py:93-94); a pixel whose C is not positive definite -- a forecast block the second-order
// correction left indefinite -- is therefore solved by elimination with partial pivoting (warp_gepp): the warp
// takes its failed pixels one at a time through shared memory, lane r rebuilding row r of [C | b].
template <int P>
__global__ void __launch_bounds__(kSolveThreads)
blend_prior_kernel(const double* __restrict__ prior_mean, const double* __restrict__ prior_inv,
                   const double* __restrict__ x_f, const double* __restrict__ pinv_f, int64_t ld, int64_t n_pix,
                   double* __restrict__ x_out, double* __restrict__ c_out, int32_t* __restrict__ n_bad,
                   uint32_t flags) {
    constexpr int L = tri_count(P);
    __shared__ LuWork<1> s_work[kSolveThreads / 32];
    const int64_t i = (int64_t)blockIdx.x * kSolveThreads + threadIdx.x;
    const bool bcast = (flags & KAFKA_PRIOR_BROADCAST) != 0;
    bool failed = false;
    if (i < n_pix) {
        double Pf[L], Cp[L], mu[P], xf[P], b1[P], b2[P];
#pragma unroll
        for (int c = 0; c < L; ++c) {
            Pf[c] = ldg_stream(pinv_f + (int64_t)c * ld + i);
            Cp[c] = bcast ? prior_inv[c] : ldg_stream(prior_inv + (int64_t)c * ld + i);
        }
#pragma unroll
        for (int p = 0; p < P; ++p) {
            mu[p] = bcast ? prior_mean[p] : ldg_stream(prior_mean + (int64_t)p * ld + i);
            xf[p] = ldg_stream(x_f + (int64_t)p * ld + i);
        }
        symv<P>(Pf, mu, b1);       // the reference's cross weighting (kf_tools.py:90), kept as is
        symv<P>(Cp, xf, b2);
#pragma unroll
        for (int c = 0; c < L; ++c) {
            Pf[c] += Cp[c];
            stg_stream(c_out + (int64_t)c * ld + i, Pf[c]);
        }
#pragma unroll
        for (int p = 0; p < P; ++p) b1[p] += b2[p];
        double invd[P];
        const bool ok = chol_factor<P>(Pf, invd);
        chol_solve<P>(Pf, invd, b1);
#pragma unroll
        for (int p = 0; p < P; ++p) stg_stream(x_out + (int64_t)p * ld + i, ok ? b1[p] : xf[p]);
        failed = !ok;
    }
    unsigned todo = __ballot_sync(0xffffffffu, failed);
    int bad = 0;
    if (todo) {
        const int lane = threadIdx.x & 31;
        LuWork<1>& W = s_work[threadIdx.x >> 5];
        const int64_t i0 = (int64_t)blockIdx.x * kSolveThreads + (threadIdx.x & ~31);
        while (todo) {
            const int src = __ffs(todo) - 1;
            todo &= todo - 1;
            const int64_t j = i0 + src;
            if (lane < P) {
                const int r = lane;
                double b = 0.0;
                for (int q = 0; q < P; ++q) {                       // row r of Pf and Cp, b_r as symv adds it up
                    const int c = r >= q ? tri(r, q) : tri(q, r);
                    const double pf = pinv_f[(int64_t)c * ld + j];
                    b = fma(pf, bcast ? prior_mean[q] : prior_mean[(int64_t)q * ld + j], b);
                    W.m[r][q] = pf;
                }
                double b2 = 0.0;
                for (int q = 0; q < P; ++q) {
                    const int c = r >= q ? tri(r, q) : tri(q, r);
                    const double cp = bcast ? prior_inv[c] : prior_inv[(int64_t)c * ld + j];
                    b2 = fma(cp, x_f[(int64_t)q * ld + j], b2);
                    W.m[r][q] += cp;
                }
                W.m[r][P] = b + b2;
            }
            __syncwarp();
            const bool ok = warp_gepp(W, P, 1, lane);
            if (ok && lane < P) x_out[(int64_t)lane * ld + j] = W.m[lane][P];
            bad += ok ? 0 : 1;
            __syncwarp();
        }
    }
    if ((threadIdx.x & 31) == 0 && bad != 0 && n_bad) atomicAdd(n_bad, bad);   // integer count: order-independent
}

}  // namespace kafka

using namespace kafka;

extern "C" {

// scratch, T = tiles of 128 pixels: [0, T) partial sums of squares | [T, 2T) pixels left at the forecast |
// [2T, 4T) one bit per pixel "Cholesky failed" (4 words per tile) | [4T] pixels solved by the pivoted
// elimination (int32) and the second stage's arrival ticket (uint32) | 3 spare | 2 kFinishCtas per-CTA sums
int64_t kafka_solve_scratch_doubles(int64_t n_pix) {
    if (n_pix < 0) return -1;
    return 4 * ((n_pix + kSolveThreads - 1) / kSolveThreads) + 4 + 2 * kFinishCtas;
}

int kafka_solve_bands(int64_t n_pix, int64_t ld, int n_params, const kafka_solve_band* bands, int n_bands,
                      const double* x_prev, const double* x_f, const double* pinv_f, double* x_out, double* a_out,
                      double* innovations, double* fwd, double* norm2, int32_t* n_bad, double* scratch,
                      uint32_t flags, void* stream) {
    KAFKA_CHECK_ARG(n_pix >= 0 && ld >= n_pix, "kafka_solve_bands: need 0 <= n_pix <= ld (n_pix=%lld ld=%lld)",
                    (long long)n_pix, (long long)ld);
    KAFKA_CHECK_ARG(n_params >= 1 && n_params <= KAFKA_MAX_PARAMS, "kafka_solve_bands: n_params=%d outside 1..%d",
                    n_params, KAFKA_MAX_PARAMS);
    KAFKA_CHECK_ARG(n_bands >= 0 && n_bands <= KAFKA_MAX_BANDS, "kafka_solve_bands: n_bands=%d outside 0..%d", n_bands,
                    KAFKA_MAX_BANDS);
    KAFKA_CHECK_ARG(n_bands == 0 || bands, "kafka_solve_bands: bands is null");
    cudaStream_t s = (cudaStream_t)stream;
    const bool report_lu = (flags & KAFKA_SOLVE_COUNT_LU) != 0;
    if (n_pix == 0) {
        if (norm2) KAFKA_CUDA(cudaMemsetAsync(norm2, 0, sizeof(double), s));
        if (n_bad) KAFKA_CUDA(cudaMemsetAsync(n_bad, 0, (report_lu ? 2 : 1) * sizeof(int32_t), s));
        return 0;
    }
    KAFKA_CHECK_ARG(x_prev && x_f && pinv_f && x_out && scratch, "kafka_solve_bands: null pointer");
    SolveParams prm;
    memset(&prm, 0, sizeof(prm));
    for (int b = 0; b < n_bands; ++b) {
        SolveBandDev& d = prm.bands[b];
        const kafka_solve_band& h = bands[b];
        KAFKA_CHECK_ARG(h.y && h.w, "kafka_solve_bands: band %d lacks y/w", b);
        d.y = h.y;
        d.w = h.w;
        d.kind = h.kind;
        for (int p = 0; p < KAFKA_MAX_PARAMS; ++p) d.src[p] = -1;
        if (h.kind == KAFKA_BAND_LINEAR) {
            for (int p = 0; p < n_params; ++p) d.row[p] = h.row[p];
            d.offset = h.offset;
        } else if (h.kind == KAFKA_BAND_WCM) {
            for (int p = 0; p < 6; ++p) d.row[p] = h.row[p];
            for (int k = 0; k < 2; ++k) {
                int slot = h.mapper[k];
                KAFKA_CHECK_ARG(slot >= 0 && slot < n_params && d.src[slot] < 0,
                                "kafka_solve_bands: band %d maps WCM input %d to slot %d", b, k, slot);
                d.src[slot] = (int8_t)k;
            }
            KAFKA_CHECK_ARG(h.row[5] > 0.0, "kafka_solve_bands: band %d has cos(incidence angle) <= 0", b);
        } else if (h.kind == KAFKA_BAND_ARRAYS || h.kind == KAFKA_BAND_UNIFORM) {
            KAFKA_CHECK_ARG(h.h0 && h.jac, "kafka_solve_bands: band %d lacks h0/jac", b);
            KAFKA_CHECK_ARG(h.n_dim >= 1 && h.n_dim <= KAFKA_MAX_DIM, "kafka_solve_bands: band %d n_dim=%d", b, h.n_dim);
            d.h0 = h.h0;
            d.jac = h.jac;
            d.n_planes = h.n_dim;
            for (int k = 0; k < h.n_dim; ++k) {
                int slot = h.mapper[k];
                KAFKA_CHECK_ARG(slot >= 0 && slot < n_params, "kafka_solve_bands: band %d maps plane %d to slot %d", b, k,
                                slot);
                KAFKA_CHECK_ARG(d.src[slot] < 0, "kafka_solve_bands: band %d maps two planes to slot %d", b, slot);
                d.src[slot] = (int8_t)k;
            }
            d.identity = h.kind == KAFKA_BAND_ARRAYS && h.n_dim == n_params;
            for (int k = 0; k < h.n_dim; ++k) d.identity = d.identity && h.mapper[k] == k;
        } else {
            set_error("kafka_solve_bands: band %d has unknown kind %d", b, h.kind);
            return KAFKA_E_BADARG;
        }
    }
    prm.x_prev = x_prev;
    prm.x_f = x_f;
    prm.pinv_f = pinv_f;
    prm.x_out = x_out;
    prm.a_out = a_out;
    prm.innovations = innovations;
    prm.fwd = fwd;
    prm.partials = scratch;
    const unsigned grid = (unsigned)((n_pix + kSolveThreads - 1) / kSolveThreads);
    prm.fail_bits = reinterpret_cast<uint32_t*>(scratch + 2 * (size_t)grid);
    prm.n_lu = reinterpret_cast<int32_t*>(scratch + 4 * (size_t)grid);
    prm.n_pix = n_pix;
    prm.ld = ld;
    prm.n_bands = n_bands;
    prm.flags = flags;
    int n_dim_max = 0;
    for (int b = 0; b < n_bands; ++b)
        if (prm.bands[b].kind == KAFKA_BAND_ARRAYS && prm.bands[b].n_planes > n_dim_max) n_dim_max = prm.bands[b].n_planes;
    const size_t ring = (size_t)kAsyncStages * (3 + n_dim_max) * kSolveThreads * sizeof(double);
    // second stage (solve_lu.cuh): pixels whose Cholesky failed are solved by pivoted elimination, then the
    // partial sums are added up in a fixed order
    auto finish = [&]() -> int {
        unsigned ctas = (grid + kFinishWarps * 32 - 1) / (kFinishWarps * 32);
        ctas = ctas < (unsigned)kFinishCtas ? ctas : (unsigned)kFinishCtas;
        finish_solve_kernel<<<ctas, kFinishWarps * 32, 0, s>>>(prm, n_params, grid, scratch + 4 * (size_t)grid + 4,
                                                               reinterpret_cast<unsigned*>(prm.n_lu + 1), norm2, n_bad,
                                                               report_lu);
        return launch_status("finish_solve_kernel");
    };
    int rc = KAFKA_E_UNSUPPORTED;
    if (n_params >= kPairFromParams || (flags & KAFKA_SOLVE_PAIR)) {
        bool all_identity_pair = n_bands > 0;
        for (int b = 0; b < n_bands; ++b) all_identity_pair = all_identity_pair && prm.bands[b].identity;
        KAFKA_DISPATCH_P(n_params, {
            auto kern = all_identity_pair ? solve_bands_pair_kernel<kP, true> : solve_bands_pair_kernel<kP, false>;
            cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(ring / 2));
            // the two CTAs of a tile add into one scratch slot, failure bits are only ever set: all start at zero
            if (e == cudaSuccess) e = cudaMemsetAsync(scratch, 0, (4 * (size_t)grid + 1) * sizeof(double), s);
            if (e != cudaSuccess) {
                set_error("launch set-up failed: %s", cudaGetErrorString(e));
                rc = (int)e;
            } else {
                kern<<<(unsigned)((n_pix + kPairPix - 1) / kPairPix), kPairThreads, ring / kPairSplit, s>>>(prm, n_dim_max);
                rc = launch_status("solve_bands_pair_kernel");
            }
        });
        if (rc != 0) return rc;
        return finish();
    }
    bool all_identity = n_bands > 0, all_uniform = n_bands > 0;
    for (int b = 0; b < n_bands; ++b) {
        all_identity = all_identity && prm.bands[b].identity;
        all_uniform = all_uniform && prm.bands[b].kind == KAFKA_BAND_UNIFORM;
    }
    KAFKA_DISPATCH_P(n_params, {
        auto kern = all_identity ? solve_bands_kernel<kP, 1> : all_uniform ? solve_bands_kernel<kP, 2> : solve_bands_kernel<kP, 0>;
        const size_t smem = ring + (size_t)(SolveBlock<kP>::L - SolveBlock<kP>::R) * kSolveThreads * sizeof(double);
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) {
            set_error("cudaFuncSetAttribute failed: %s", cudaGetErrorString(e));
            rc = (int)e;
        } else {
            kern<<<grid, kSolveThreads, smem, s>>>(prm, n_dim_max);
            rc = launch_status("solve_bands_kernel");
        }
    });
    if (rc != 0) return rc;
    return finish();
}

int kafka_wcm_forward(const double* x, int64_t ld, int64_t n, const double* params_host, double* sigma0, double* grad,
                      void* stream) {
    KAFKA_CHECK_ARG(n >= 0 && ld >= n, "kafka_wcm_forward: need 0 <= n <= ld");
    KAFKA_CHECK_ARG(params_host, "kafka_wcm_forward: params is null");
    KAFKA_CHECK_ARG(params_host[5] > 0.0, "kafka_wcm_forward: cos(incidence angle) must be > 0");
    if (n == 0) return 0;
    KAFKA_CHECK_ARG(x && sigma0, "kafka_wcm_forward: null pointer");
    WcmParams wp;
    for (int k = 0; k < 6; ++k) wp.prm[k] = params_host[k];
    wcm_forward_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(x, ld, n, wp, sigma0, grad);
    return launch_status("wcm_forward_kernel");
}

int kafka_blend_prior(const double* prior_mean, const double* prior_inv, const double* x_f, const double* pinv_f,
                      int64_t ld, int64_t n_pix, int n_params, double* x_out, double* c_out, int32_t* n_bad,
                      uint32_t flags, void* stream) {
    KAFKA_CHECK_ARG(n_pix >= 0 && ld >= n_pix, "kafka_blend_prior: need 0 <= n_pix <= ld");
    KAFKA_CHECK_ARG(n_params >= 1 && n_params <= KAFKA_MAX_PARAMS, "kafka_blend_prior: n_params=%d", n_params);
    cudaStream_t s = (cudaStream_t)stream;
    if (n_bad) KAFKA_CUDA(cudaMemsetAsync(n_bad, 0, sizeof(int32_t), s));
    if (n_pix == 0) return 0;
    KAFKA_CHECK_ARG(prior_mean && prior_inv && x_f && pinv_f && x_out && c_out, "kafka_blend_prior: null pointer");
    const unsigned grid = (unsigned)((n_pix + kSolveThreads - 1) / kSolveThreads);
    int rc = KAFKA_E_UNSUPPORTED;
    KAFKA_DISPATCH_P(n_params, {
        blend_prior_kernel<kP><<<grid, kSolveThreads, 0, s>>>(prior_mean, prior_inv, x_f, pinv_f, ld, n_pix, x_out,
                                                             c_out, n_bad, flags);
        rc = launch_status("blend_prior_kernel");
    });
    return rc;
}

}  // extern "C"
