// Gaussian-process (ARD squared-exponential) emulator kernels for sm_100a.
//
// Replaces run_emulator + gp.predict / gp.hessian of the reference
// (kafka/inference/utils.py:68-106,206; kafka/inference/kf_tools.py:28).
//
// Maths (see include/kafka_b200.h and oracle/gp.py).  With a per-emulator centre c,
// X' = X - c, x' = x - c, zh_jd = theta_d X'_jd:
//     k_j   = exp( e_j - 0.5 sum_d theta_d x'_d^2 + sum_d zh_jd x'_d ),
//             e_j = ln s_f - 0.5 sum_d theta_d X'_jd^2
//     w_j   = k_j alpha_j,   mean = sum_j w_j,   G_d = sum_j w_j zh_jd
//     dmean/dx_d = G_d - theta_d x'_d mean
// The table stores zh and e pre-multiplied by L = 1024/ln2, so the dot product IS the
// base-2^(1/1024) exponent: k = rint(a), r = a - k is exact, 2^(k/1024) comes from a
// 1024-entry table and exp(r ln2/1024) from a degree-3 polynomial (|r| <= 1/2, truncation
// 5.4e-16).  The per-pixel factor exp(-0.5 sum theta x'^2) is pulled out of the loop and
// applied once at the end (a warp keeps it inside the exponent when one of its pixels is so
// far from the centre that the un-shifted exponent could overflow).  Per (pixel, band,
// training point): 2D FMAs + 8 fp64 ops, i.e. 28 fp64-pipe instructions at D=10 against
// the 5D+28 = 78 flops of the textbook form.  The training table (T rows of {zh[D], e,
// alpha}) is staged once per CTA in shared memory by a TMA bulk copy and then read as
// warp-wide broadcasts.  The kernel is bound by the fp64 pipe, not by HBM.
#include <math.h>
#include <math_constants.h>

#include "common.cuh"

namespace kafka {

// exponent granularity: 2^(1/1024) (1024-entry table + degree-3 polynomial).  benchmarks/gp_variants.py
// can still build the 256-entry / degree-4 variant (KAFKA_EXP_LOG2N=8, table under benchmarks/exp_tables/):
// on B200 it is 6 % slower, and the 64-entry / degree-5 one measured in round 1 another 4.5 %.
#ifndef KAFKA_EXP_LOG2N
#define KAFKA_EXP_LOG2N 10
#endif
#if KAFKA_EXP_LOG2N != 10 && KAFKA_EXP_LOG2N != 8
#error "KAFKA_EXP_LOG2N must be 10 (product) or 8 (benchmark variant)"
#endif
// unchecked exponent insert in the hot loop (see exp_scaled); with LOG2N 10: 5.3 % faster than 8 + checked
#ifndef KAFKA_EXP_CLAMP
#define KAFKA_EXP_CLAMP 1
#endif
// both measured on B200 (profiles/r02_gp_variants.txt): 0.45 % and 0.85 %, 1.35 % together
#ifndef KAFKA_EXP_INSERT_FIRST
#define KAFKA_EXP_INSERT_FIRST 1
#endif
#ifndef KAFKA_GP_PREFETCH_HEAD
#define KAFKA_GP_PREFETCH_HEAD 1
#endif
constexpr int kExpLog2N = KAFKA_EXP_LOG2N;
constexpr int kExpN = 1 << kExpLog2N;
// in global memory (not __constant__): every CTA copies it to shared memory, and constant
// loads at 32 different addresses per warp would serialise
__device__ __align__(16) const double g_exp2_tab[kExpN] = {
#if KAFKA_EXP_LOG2N == 8
#include "../../benchmarks/exp_tables/exp_table256.inc"
#else
#include "exp_table1024.inc"
#endif
};

// ---- table layout -----------------------------------------------------------
__host__ __device__ constexpr int dpad(int D) { return (D + 1) & ~1; }
__host__ __device__ constexpr int tab_hdr(int D) { return 2 * dpad(D) + 4; }   // theta_e, centre, s_f, nugget, wide, L*cbar
// row: zh[DP], e, alpha, then the pair the linearise loop reads instead of {e, alpha}:
//   e2  = e + L (ln|alpha| - cbar)   -- the weight's magnitude folded into the exponent (cbar = mean ln|alpha|,
//                                       applied once per pixel after the loop): one DMUL fewer per evaluation
//   sgn = 0 or 0x80000000 as an integer in the slot's low word -- XOR-ed into the result's high word: the
//                                       weight's sign (a zero weight gets e2 = -2^29, i.e. an exact or near zero)
__host__ __device__ constexpr int tab_row(int D) { return dpad(D) + 4; }

// ---- exp in base 2^(1/N): argument already scaled by L = N/ln2 ----------------------
#if KAFKA_EXP_LOG2N == 8
constexpr double kL = 3.69329930467574627073e+02;        // 256/ln2
constexpr double kInvL = 2.70760617406228627432e-03;     // ln2/256
constexpr uint32_t kHiUnder = 0x410feb6cu;               // high word of 708 L
__constant__ double c_exp_poly[6] = {0.0, 2.23939519087515698716e-12, 3.30830268054137129301e-09,
                                      3.66556559691010622906e-06, 2.70760617406228627432e-03,
                                      6755399441055744.0 /* 1.5 * 2^52 */};
#else
constexpr double kL = 1.47731972187029850829e+03;        // 1024/ln2
constexpr double kInvL = 6.76901543515571568581e-04;     // ln2/1024
constexpr uint32_t kHiUnder = 0x412feb6cu;               // high word of 708 L
__constant__ double c_exp_poly[6] = {0.0, 0.0, 5.16922293834589264532e-11, 2.29097849806881639316e-07,
                                      6.76901543515571568581e-04, 6755399441055744.0 /* 1.5 * 2^52 */};
#endif
constexpr double kFarShift = 600.0 * kL;                 // beyond this the per-pixel factor stays in the exponent

// exp(a / L) for finite a < 709 L.  fp64-pipe cost: 3 adds + 5 / 4 / 3 FMAs + 1 mul.
// kChecked: the only range test is one unsigned compare on the high word (exactly 0 below
// -708 L), so the function is branch-free and several evaluations interleave.
// !kChecked: for callers that guarantee |a| < 2^30 (then k below is exact): the binary exponent
// is clamped with one integer max instead, and what would underflow comes out as a number of
// the order of 2^-1021 instead of 0 -- three integer instructions fewer per evaluation.
// |rel err| < 6e-16.  NaN / +inf / overflow are NOT handled here: callers check their inputs
// once per pixel, outside the hot loop.
// The evaluation comes in two halves so that a caller with several arguments can issue all the
// table loads before the first polynomial (source order is what the scheduler follows when
// registers are tight: with the load written after the polynomial it was issued right before its
// use and every evaluation waited out the shared-memory latency).
struct ExpPart {
    double r;       // a - rint(a), exact, |r| <= 1/2
    double tv;      // 2^((k mod N)/N) from the table
    int k;          // rint(a)
};

__device__ __forceinline__ ExpPart exp_lookup(double a, const double* __restrict__ tab) {
    ExpPart e;
    const double t = a + c_exp_poly[5];              // magic-number add: rint(a) in the low word (F2I/I2F measured slower)
    e.k = __double2loint(t);
    e.r = a - (t - c_exp_poly[5]);
    e.tv = tab[e.k & (kExpN - 1)];
    return e;
}

// sgn (0 or 0x80000000) is XOR-ed into the result's high word: the sign of the result.
template <bool kChecked>
__device__ __forceinline__ double exp_finish(const ExpPart& e, double a, int sgn = 0) {
    const double r = e.r;
    const int k = e.k;
#if KAFKA_EXP_LOG2N == 8
    double p = fma(r, c_exp_poly[1], c_exp_poly[2]);
    p = fma(p, r, c_exp_poly[3]);
#else
    double p = fma(r, c_exp_poly[2], c_exp_poly[3]);
#endif
    p = fma(p, r, c_exp_poly[4]);
    p = fma(p, r, 1.0);
    if constexpr (!kChecked && KAFKA_EXP_INSERT_FIRST) {
        // binary exponent and sign go into the TABLE value (available long before the polynomial is), so
        // that nothing but the multiply stands between the polynomial and the first use of the result:
        // tv 2^ex stays a normal number for ex >= -1021, and p is within 1 % of 1
        const int ex = max(k >> kExpLog2N, -1021);
        const double tvs = __hiloint2double((__double2hiint(e.tv) + (ex << 20)) ^ sgn, __double2loint(e.tv));
        return tvs * p;
    }
    const double v = e.tv * p;                           // in (0.99, 2.01)
    if constexpr (!kChecked) {
        const int ex = max(k >> kExpLog2N, -1021);
        return __hiloint2double((__double2hiint(v) + (ex << 20)) ^ sgn, __double2loint(v));
    } else {
        const int hi_n = __double2hiint(v) + ((k >> kExpLog2N) << 20);
        const int lo_n = __double2loint(v);
        int hi, lo;
        // selp keeps this branch-free (the compiler otherwise jumps around the exponent insert)
        asm("{\n\t.reg .pred p;\n\t"
            "setp.ge.u32 p, %2, %3;\n\t"
            "selp.b32 %0, 0, %4, p;\n\t"
            "selp.b32 %1, 0, %5, p;\n\t}"
            : "=r"(hi), "=r"(lo)
            : "r"(__double2hiint(a)), "r"(0x80000000u + kHiUnder), "r"(hi_n), "r"(lo_n));
        return __hiloint2double(hi ^ sgn, lo);
    }
}

template <bool kChecked = true>
__device__ __forceinline__ double exp_scaled(double a, const double* __restrict__ tab) {
    return exp_finish<kChecked>(exp_lookup(a, tab), a);
}

// ---- linearise: mean + Jacobian for every (pixel, band) -------------------------
struct GpBandDev {
    const double* table;
    int n_train;
    int mapper[KAFKA_MAX_DIM];
};
struct GpLinParams {
    GpBandDev bands[KAFKA_MAX_BANDS];
    const double* x;
    const uint8_t* obs_mask;
    double* h0;
    double* jac;
    int64_t ld;
    int64_t n_pix;
    uint32_t flags;
    int32_t n_bands;
};

// tuning knobs (defaults are the measured best; benchmarks/gp_variants.py sweeps them)
#ifndef KAFKA_GP_THREADS
#define KAFKA_GP_THREADS 128
#endif
#ifndef KAFKA_GP_UNROLL
#define KAFKA_GP_UNROLL 1     // 1: 21.15 ms, 2: 21.78, 4: 21.19, 8: 21.20 (4 M px x 10 bands x 250 points)
#endif
#ifndef KAFKA_GP_MINBLOCKS
#define KAFKA_GP_MINBLOCKS 1
#endif
#ifndef KAFKA_GP_FORCE_PPT
#define KAFKA_GP_FORCE_PPT 0
#endif
#ifndef KAFKA_GP_PPT_BIG
#define KAFKA_GP_PPT_BIG 2    // pixels per thread on large images
#endif
constexpr int kGpPptBig = KAFKA_GP_PPT_BIG;
constexpr int kGpThreads = KAFKA_GP_THREADS;
constexpr int kGpUnroll = KAFKA_GP_UNROLL;

// Inner loop over the training points of one band for PPT pixels per thread.
// rows[j] = {L zh[DP], L e, alpha} in shared memory, read as warp-wide broadcasts.
template <int D, int PPT, bool kShiftInside>
__device__ __forceinline__ void gp_accumulate(const double* __restrict__ rows, int T, const double (&xr)[PPT][D],
                                              const double (&sh)[PPT], double (&mu)[PPT], double (&G)[PPT][D],
                                              const double* __restrict__ tab64) {
    constexpr int DP = dpad(D), RS = tab_row(D);
    const double* const rows_end = rows + (size_t)T * RS;       // the row pointer is the loop variable: no separate counter
    // KAFKA_GP_PREFETCH_HEAD: the two operands the first FMA of a row needs ({e2, sgn} and zh[0..1]) are
    // fetched at the end of the previous trip, behind the G sums, instead of right before their use
    double2 ea_n = *reinterpret_cast<const double2*>(rows + DP + 2);
    double2 z0_n = *reinterpret_cast<const double2*>(rows);
#pragma unroll kGpUnroll
    for (const double* row = rows; row != rows_end; row += RS) {
        double zh[DP];
#pragma unroll
        for (int d = (KAFKA_GP_PREFETCH_HEAD ? 2 : 0); d < DP; d += 2) {
            double2 v = *reinterpret_cast<const double2*>(row + d);
            zh[d] = v.x;
            zh[d + 1] = v.y;
        }
        double2 ea;
        if constexpr (KAFKA_GP_PREFETCH_HEAD) {
            ea = ea_n;
            zh[0] = z0_n.x;
            zh[1] = z0_n.y;
        } else {
            ea = *reinterpret_cast<const double2*>(row + DP + 2);                 // {e2, sgn}
        }
        const int sgn = __double2loint(ea.y);                                    // 0 or 0x80000000: the weight's sign
        double a[PPT];
        ExpPart ep[PPT];
#pragma unroll
        for (int pp = 0; pp < PPT; ++pp) {
            a[pp] = kShiftInside ? ea.x + sh[pp] : ea.x;
#pragma unroll
            for (int d = 0; d < D; ++d) a[pp] = fma(zh[d], xr[pp][d], a[pp]);
        }
#pragma unroll
        for (int pp = 0; pp < PPT; ++pp) ep[pp] = exp_lookup(a[pp], tab64);
#pragma unroll
        for (int pp = 0; pp < PPT; ++pp) {
            const double w = exp_finish<kShiftInside || !KAFKA_EXP_CLAMP>(ep[pp], a[pp], sgn);  // = k_j alpha_j / exp(cbar)
            mu[pp] += w;
#pragma unroll
            for (int d = 0; d < D; ++d) G[pp][d] = fma(w, zh[d], G[pp][d]);
        }
        if constexpr (KAFKA_GP_PREFETCH_HEAD) {           // one row past the table on the last trip: allocated, unused
            ea_n = *reinterpret_cast<const double2*>(row + RS + DP + 2);
            z0_n = *reinterpret_cast<const double2*>(row + RS);
        }
    }
}

template <int D, int PPT>
__global__ void __launch_bounds__(kGpThreads, (KAFKA_GP_MINBLOCKS == 1 && PPT == 2 && D <= 10 && KAFKA_GP_THREADS == 128) ? 3 : KAFKA_GP_MINBLOCKS)
gp_linearize_kernel(const __grid_constant__ GpLinParams prm) {
    constexpr int DP = dpad(D), HDR = tab_hdr(D), RS = tab_row(D);
    extern __shared__ __align__(16) double smem[];
    __shared__ uint64_t bar;
    __shared__ __align__(16) double s_exp[kExpN];

    // band is the FAST grid index: the CTAs of all bands of one pixel tile run together, so
    // the tile's state is fetched from HBM once and served to the other bands by L2
    const int band_idx = (int)(blockIdx.x % prm.n_bands);
    const int64_t tile_idx = blockIdx.x / prm.n_bands;
    const GpBandDev& band = prm.bands[band_idx];
    const int T = band.n_train;
    const uint32_t bytes = (uint32_t)((HDR + T * RS) * sizeof(double));

    if (threadIdx.x == 0) mbar_init(&bar, 1);
    __syncthreads();
    if (threadIdx.x == 0) {
        mbar_expect_tx(&bar, bytes + (uint32_t)sizeof(s_exp));
        tma_bulk_g2s(s_exp, g_exp2_tab, (uint32_t)sizeof(s_exp), &bar);
        constexpr uint32_t kChunk = 16384;
        for (uint32_t off = 0; off < bytes; off += kChunk)
            tma_bulk_g2s(reinterpret_cast<char*>(smem) + off, reinterpret_cast<const char*>(band.table) + off,
                         min(kChunk, bytes - off), &bar);
    }

    // pixel states while the table is in flight
    const int64_t tile0 = tile_idx * (kGpThreads * PPT);
    int64_t pix[PPT];
    bool live[PPT];
    double xr[PPT][D];
    bool any_live = false;
#pragma unroll
    for (int pp = 0; pp < PPT; ++pp) {
        pix[pp] = tile0 + pp * kGpThreads + threadIdx.x;
        bool in = pix[pp] < prm.n_pix;
        live[pp] = in && (prm.obs_mask == nullptr || prm.obs_mask[(int64_t)band_idx * prm.ld + pix[pp]] != 0);
        any_live |= live[pp];
#pragma unroll
        for (int d = 0; d < D; ++d)
            xr[pp][d] = in ? prm.x[(int64_t)band.mapper[d] * prm.ld + pix[pp]] : 0.0;
    }

    mbar_wait(&bar, 0);

    const double* theta_e = smem;
    const double* centre = smem + DP;
    const double* rows = smem + HDR;

    double sh[PPT], mu[PPT], G[PPT][D];
    bool far = smem[2 * DP + 2] != 0.0, bad[PPT];             // a "wide" table takes the checked path
#pragma unroll
    for (int pp = 0; pp < PPT; ++pp) {
        double s = 0.0;
#pragma unroll
        for (int d = 0; d < D; ++d) {
            xr[pp][d] -= centre[d];
            s = fma(theta_e[d] * xr[pp][d], xr[pp][d], s);
            G[pp][d] = 0.0;
        }
        sh[pp] = (-0.5 * kL) * s;                          // scaled like the table
        mu[pp] = 0.0;
        far |= live[pp] && !(sh[pp] > -kFarShift);         // also true for NaN; padding lanes do not count
        // a non-finite state (NaN / inf in any input) is caught here, once per pixel, and
        // turned into NaN outputs below: exp_scaled itself assumes finite arguments
        bad[pp] = !(fabs(s) <= 1.79769313486231570815e308);
    }

    if (__any_sync(0xffffffffu, any_live)) {
        // common case: exp(sh) is applied after the loop; a warp with a far-away pixel keeps
        // sh inside the exponent so that nothing can overflow
        if (!__any_sync(0xffffffffu, far)) {
            gp_accumulate<D, PPT, false>(rows, T, xr, sh, mu, G, s_exp);
            // L * mean ln|alpha|, taken out of every row's exponent (read here, not before the loop: more
            // registers live across it push the polynomial's constants back into the loop)
            const double lcbar = smem[2 * DP + 3];
#pragma unroll
            for (int pp = 0; pp < PPT; ++pp) {
                // the pixel's own factor is recomputed (same operations, same bits as sh[pp]) rather than
                // kept in registers across the loop
                double s2 = 0.0;
#pragma unroll
                for (int d = 0; d < D; ++d) s2 = fma(theta_e[d] * xr[pp][d], xr[pp][d], s2);
                const double scale = exp_scaled((-0.5 * kL) * s2 + lcbar, s_exp);
                mu[pp] *= scale;
#pragma unroll
                for (int d = 0; d < D; ++d) G[pp][d] *= scale;
            }
        } else {
            gp_accumulate<D, PPT, true>(rows, T, xr, sh, mu, G, s_exp);
            const double scale = exp_scaled(smem[2 * DP + 3], s_exp);
#pragma unroll
            for (int pp = 0; pp < PPT; ++pp) {
                mu[pp] *= scale;
#pragma unroll
                for (int d = 0; d < D; ++d) G[pp][d] *= scale;
            }
        }
#pragma unroll
        for (int pp = 0; pp < PPT; ++pp)
#pragma unroll
            for (int d = 0; d < D; ++d) G[pp][d] *= kInvL;  // table rows are L * zh
    }

    const bool r32 = (prm.flags & KAFKA_ROUND_F32) != 0;
#pragma unroll
    for (int pp = 0; pp < PPT; ++pp) {
        if (pix[pp] >= prm.n_pix) continue;
        if (bad[pp]) {
            mu[pp] = CUDART_NAN;
#pragma unroll
            for (int d = 0; d < D; ++d) G[pp][d] = CUDART_NAN;
        }
        double m = live[pp] ? mu[pp] : 0.0;
        prm.h0[(int64_t)band_idx * prm.ld + pix[pp]] = r32 ? round_f32(m) : m;
        if (prm.jac) {
#pragma unroll
            for (int d = 0; d < D; ++d) {
                double g = live[pp] ? G[pp][d] - theta_e[d] * xr[pp][d] * mu[pp] : 0.0;
                prm.jac[((int64_t)band_idx * D + d) * prm.ld + pix[pp]] = r32 ? round_f32(g) : g;
            }
        }
    }
}

// ---- Hessian: hess_de = sum_j w_j u_jd u_je - delta_de theta_d mean, u_jd = zh_jd - theta_d x'_d
template <int D>
__global__ void __launch_bounds__(kGpThreads)
gp_hessian_kernel(const double* __restrict__ xin, int64_t ld, int64_t n_pts, const double* __restrict__ table, int T,
                  double* __restrict__ hess) {
    constexpr int DP = dpad(D), HDR = tab_hdr(D), RS = tab_row(D), L = tri_count(D);
    extern __shared__ __align__(16) double smem[];
    __shared__ double s_exp[kExpN];
    const int total = HDR + T * RS;
    for (int k = threadIdx.x; k < total; k += kGpThreads) smem[k] = table[k];
    for (int k = threadIdx.x; k < kExpN; k += blockDim.x) s_exp[k] = g_exp2_tab[k];
    __syncthreads();
    int64_t i = (int64_t)blockIdx.x * kGpThreads + threadIdx.x;
    if (i >= n_pts) return;
    const double* theta_e = smem;
    const double* centre = smem + DP;
    const double* rows = smem + HDR;
    double xr[D], tx[D], acc[L], mu = 0.0, s = 0.0;
#pragma unroll
    for (int d = 0; d < D; ++d) {
        xr[d] = xin[(int64_t)d * ld + i] - centre[d];
        tx[d] = theta_e[d] * xr[d];
        s = fma(tx[d], xr[d], s);
    }
#pragma unroll
    for (int c = 0; c < L; ++c) acc[c] = 0.0;
    const double sh = (-0.5 * kL) * s;
    for (int j = 0; j < T; ++j) {
        const double* row = rows + j * RS;
        double a = row[DP] + sh, u[D];
#pragma unroll
        for (int d = 0; d < D; ++d) {
            a = fma(row[d], xr[d], a);
            u[d] = fma(row[d], kInvL, -tx[d]);
        }
        double w = exp_scaled(a, s_exp) * row[DP + 1];
        mu += w;
#pragma unroll
        for (int r = 0; r < D; ++r) {
            double wu = w * u[r];
#pragma unroll
            for (int q = 0; q <= r; ++q) acc[tri(r, q)] = fma(wu, u[q], acc[tri(r, q)]);
        }
    }
#pragma unroll
    for (int r = 0; r < D; ++r)
#pragma unroll
        for (int q = 0; q <= r; ++q) {
            double v = acc[tri(r, q)] - (r == q ? theta_e[r] * mu : 0.0);
            hess[((int64_t)r * D + q) * ld + i] = v;
            hess[((int64_t)q * D + r) * ld + i] = v;
        }
}

// ---- second-order correction of the posterior precision (kf_tools.py:26-72) ----------
// One thread per pixel, bands in sequence: the D(D+1)/2 Hessian sums of a band are
// accumulated in registers, scaled by w * innovation and added into the packed P x P
// correction block in global memory (coalesced read-modify-write, one plane at a time).
struct GpHessCorrParams {
    GpBandDev bands[KAFKA_MAX_BANDS];
    const double* x;
    const double* innovations;
    const double* w;
    double* corr;
    double* a_inout;
    int64_t ld;
    int64_t n_pix;
    int32_t n_bands;
    int32_t n_params;
};

template <int D>
__global__ void __launch_bounds__(128)
gp_hessian_correction_kernel(const __grid_constant__ GpHessCorrParams prm) {
    constexpr int DP = dpad(D), HDR = tab_hdr(D), RS = tab_row(D), L = tri_count(D);
    extern __shared__ __align__(16) double smem[];
    __shared__ double s_exp[kExpN];
    for (int k = threadIdx.x; k < kExpN; k += blockDim.x) s_exp[k] = g_exp2_tab[k];
    const int64_t i = (int64_t)blockIdx.x * 128 + threadIdx.x;
    const bool in = i < prm.n_pix;
    const int LP = tri_count(prm.n_params);
    // the output block starts from zero (or from a_inout when correcting in place)
    double* out = prm.corr ? prm.corr : prm.a_inout;
    const double sign = prm.corr ? 1.0 : -1.0;
    if (in && prm.corr)
        for (int c = 0; c < LP; ++c) prm.corr[(int64_t)c * prm.ld + i] = 0.0;
    for (int b = 0; b < prm.n_bands; ++b) {
        const GpBandDev& band = prm.bands[b];
        const int T = band.n_train;
        __syncthreads();                                   // previous band's table no longer in use
        for (int k = threadIdx.x; k < HDR + T * RS; k += 128) smem[k] = band.table[k];
        __syncthreads();
        if (!in) continue;
        const double c = prm.w[(int64_t)b * prm.ld + i] * prm.innovations[(int64_t)b * prm.ld + i];
        if (c == 0.0) continue;                            // masked pixel: zero block (kf_tools.py:49-51)
        const double* theta_e = smem;
        const double* centre = smem + DP;
        const double* rows = smem + HDR;
        double xr[D], tx[D], acc[L], mu = 0.0, s = 0.0;
#pragma unroll
        for (int d = 0; d < D; ++d) {
            xr[d] = prm.x[(int64_t)band.mapper[d] * prm.ld + i] - centre[d];
            tx[d] = theta_e[d] * xr[d];
            s = fma(tx[d], xr[d], s);
        }
#pragma unroll
        for (int k = 0; k < L; ++k) acc[k] = 0.0;
        const double sh = (-0.5 * kL) * s;
        for (int j = 0; j < T; ++j) {
            const double* row = rows + j * RS;
            double a = row[DP] + sh, u[D];
#pragma unroll
            for (int d = 0; d < D; ++d) {
                a = fma(row[d], xr[d], a);
                u[d] = fma(row[d], kInvL, -tx[d]);
            }
            const double wj = exp_scaled(a, s_exp) * row[DP + 1];
            mu += wj;
#pragma unroll
            for (int r = 0; r < D; ++r) {
                const double wu = wj * u[r];
#pragma unroll
                for (int q = 0; q <= r; ++q) acc[tri(r, q)] = fma(wu, u[q], acc[tri(r, q)]);
            }
        }
        const double cs = c * sign;
#pragma unroll
        for (int r = 0; r < D; ++r)
#pragma unroll
            for (int q = 0; q <= r; ++q) {
                const double h = acc[tri(r, q)] - (r == q ? theta_e[r] * mu : 0.0);
                const int sr = band.mapper[r], sq = band.mapper[q];
                const int cidx = sr >= sq ? tri(sr, sq) : tri(sq, sr);
                // two emulator inputs mapped to one state slot would hit the same entry twice
                // per thread; the host rejects such mappers, so this is a plain RMW
                double* dst = out + (int64_t)cidx * prm.ld + i;
                *dst = fma(cs, h, *dst);
            }
    }
    // producing corr AND correcting a_inout: second pass a -= corr
    if (in && prm.corr && prm.a_inout)
        for (int c = 0; c < LP; ++c)
            prm.a_inout[(int64_t)c * prm.ld + i] -= prm.corr[(int64_t)c * prm.ld + i];
}

// ---- predictive variance: s_f - k^T Q^-1 k, one warp per point ---------------------
__global__ void __launch_bounds__(128)
gp_variance_kernel(const double* __restrict__ xin, int64_t ld, int64_t n_pts, const double* __restrict__ table, int T,
                   int D, const double* __restrict__ inv_q, double* __restrict__ var) {
    extern __shared__ double kbuf[];             // [4 warps][T]
    __shared__ double s_exp[kExpN];
    for (int k = threadIdx.x; k < kExpN; k += blockDim.x) s_exp[k] = g_exp2_tab[k];
    __syncthreads();
    const int DP = dpad(D), HDR = tab_hdr(D), RS = tab_row(D);
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    int64_t i = (int64_t)blockIdx.x * 4 + wid;
    if (i >= n_pts) return;
    double* kv = kbuf + (size_t)wid * T;
    double s = 0.0;
    for (int d = 0; d < D; ++d) {
        double xc = xin[(int64_t)d * ld + i] - table[DP + d];
        s = fma(table[d] * xc, xc, s);
    }
    for (int j = lane; j < T; j += 32) {
        const double* row = table + HDR + (size_t)j * RS;
        double a = row[DP] - (0.5 * kL) * s;
        for (int d = 0; d < D; ++d) a = fma(row[d], xin[(int64_t)d * ld + i] - table[DP + d], a);
        kv[j] = exp_scaled(a, s_exp);
    }
    __syncwarp();
    double acc = 0.0;
    for (int r = 0; r < T; ++r) {
        const double* qrow = inv_q + (size_t)r * T;
        double part = 0.0;
        for (int j = lane; j < T; j += 32) part = fma(qrow[j], kv[j], part);
        part = warp_sum(part);
        acc = fma(kv[r], part, acc);
    }
    if (lane == 0) var[i] = table[2 * DP] - acc;
}

// ---- predictive variance with fp64 tensor-core tiles --------------------------------------
// var_i = s_f - k_i^T Q^-1 k_i for a tile of MT points is the one GEMM-shaped piece of the path
// ((N x T).(T x T), 2 T^2 flops per point): Y = K Q^-1 with K the MT x T kernel matrix of
// the tile, then var_i = s_f - sum_n K_in Y_in.  An FMA formulation of it reads three registers per
// instruction and is held to 2/3 of the fp64 pipe by the register file (benchmarks/probe.cu); the
// tensor path (mma.sync.m8n8k4.f64, DMMA) takes whole fragments per instruction and is not.
// CTA = 8 warps.  Phase 1: K (exp of the scaled dot products, as in the linearise kernel) into shared
// memory, zero padded to a multiple of 8 columns, row stride = 8 mod 16 doubles so that the fragment
// loads of two rows fall on different bank halves.  Phase 2, per 256 columns of Y: warp w owns 32
// columns and all MT rows (MT/8 x 4 accumulator tiles); the k index of the fragments is permuted so that
// a lane fetches its operands of two consecutive mma steps as one pair: k = 8 s + 2 (lane % 4) + {0, 1},
// the A pair from shared memory, the B pair from Q^-1 read through L1/L2 as Q^-1[n][k] (it is
// symmetric, so the pair is contiguous).  Q^-1 is streamed once per tile: 8 T^2 / MT bytes per point
// instead of the 8 T^2 of the one-warp-per-point kernel below.
__device__ __forceinline__ void dmma_m8n8k4(double (&c)[2], double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(c[0]), "+d"(c[1])
                 : "d"(a), "d"(b));
}

constexpr int kVarThreads = 256;
__host__ __device__ constexpr int var_cols(int T) { return (T + 7) / 8 * 8; }
__host__ __device__ constexpr int var_stride(int T) { return (var_cols(T) + 15) / 16 * 16 + 8; }

template <int MT>
__global__ void __launch_bounds__(kVarThreads, 1)
gp_variance_mma_kernel(const double* __restrict__ xin, int64_t ld, int64_t n_pts, const double* __restrict__ table, int T,
                       int D, const double* __restrict__ inv_q, double* __restrict__ var) {
    constexpr int A = MT / 8;                                   // row tiles per warp
    extern __shared__ __align__(16) double vs[];                // Ks[MT][TS] | xs[MT][16] | sh[MT] | red[8][MT]
    __shared__ __align__(16) double s_exp[kExpN];
    const int Tp = var_cols(T), TS = var_stride(T);
    const int DP = dpad(D), HDR = tab_hdr(D), RS = tab_row(D);
    double* Ks = vs;
    double* xs = Ks + (size_t)MT * TS;
    double* shv = xs + MT * 16;
    double* red = shv + MT;
    const int64_t p0 = (int64_t)blockIdx.x * MT;
    for (int k = threadIdx.x; k < kExpN; k += kVarThreads) s_exp[k] = g_exp2_tab[k];
    for (int k = threadIdx.x; k < MT * 16; k += kVarThreads) {
        const int i = k >> 4, d = k & 15;
        const int64_t p = p0 + i;
        xs[k] = (d < D && p < n_pts) ? xin[(int64_t)d * ld + p] - table[DP + d] : 0.0;   // padding points sit at the centre
    }
    __syncthreads();
    if (threadIdx.x < MT) {
        double s = 0.0;
        for (int d = 0; d < D; ++d) s = fma(table[d] * xs[threadIdx.x * 16 + d], xs[threadIdx.x * 16 + d], s);
        shv[threadIdx.x] = (-0.5 * kL) * s;
    }
    __syncthreads();
    // phase 1: K_ij = s_f exp(-1/2 |x_i - X_j|^2_theta); consecutive threads take consecutive training points
    for (int idx = threadIdx.x; idx < MT * Tp; idx += kVarThreads) {
        const int i = idx / Tp, j = idx - i * Tp;
        double v = 0.0;
        if (j < T) {
            const double* row = table + HDR + (size_t)j * RS;
            double a = row[DP] + shv[i];
            for (int d = 0; d < D; ++d) a = fma(row[d], xs[i * 16 + d], a);
            v = exp_scaled(a, s_exp);
        }
        Ks[(size_t)i * TS + j] = v;
    }
    __syncthreads();

    // phase 2
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int g = lane >> 2, t4 = lane & 3;
    double acc[A];
#pragma unroll
    for (int i = 0; i < A; ++i) acc[i] = 0.0;
    for (int nc0 = 0; nc0 < Tp; nc0 += 256) {
        const int n0 = nc0 + 32 * wid;
        if (n0 >= Tp) break;                                    // warp-uniform
        double c[A][4][2];
#pragma unroll
        for (int i = 0; i < A; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) c[i][j][0] = c[i][j][1] = 0.0;
        const double* qrow[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int n = n0 + 8 * j + g;
            qrow[j] = inv_q + (size_t)(n < T ? n : T - 1) * T;  // columns past T: any finite row, their K is zero
        }
        for (int k8 = 0; k8 < Tp; k8 += 8) {
            const int k = k8 + 2 * t4;
            double b0[4], b1[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                b0[j] = k < T ? __ldg(qrow[j] + k) : 0.0;
                b1[j] = k + 1 < T ? __ldg(qrow[j] + k + 1) : 0.0;
            }
#pragma unroll
            for (int i = 0; i < A; ++i) {
                const double2 av = *reinterpret_cast<const double2*>(Ks + (size_t)(8 * i + g) * TS + k);
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    dmma_m8n8k4(c[i][j], av.x, b0[j]);
                    dmma_m8n8k4(c[i][j], av.y, b1[j]);
                }
            }
        }
        // var partial: Y_in K_in over this warp's columns (K is zero past T)
#pragma unroll
        for (int i = 0; i < A; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int col = n0 + 8 * j + 2 * t4;
                if (col < Tp) {
                    const double2 kv = *reinterpret_cast<const double2*>(Ks + (size_t)(8 * i + g) * TS + col);
                    acc[i] = fma(c[i][j][0], kv.x, acc[i]);
                    acc[i] = fma(c[i][j][1], kv.y, acc[i]);
                }
            }
    }
#pragma unroll
    for (int i = 0; i < A; ++i) {
        double v = acc[i];
        v += __shfl_xor_sync(0xffffffffu, v, 1);
        v += __shfl_xor_sync(0xffffffffu, v, 2);
        if (t4 == 0) red[wid * MT + 8 * i + g] = v;
    }
    __syncthreads();
    if (threadIdx.x < MT && p0 + threadIdx.x < n_pts) {
        double s = 0.0;
#pragma unroll
        for (int w = 0; w < kVarThreads / 32; ++w) s += red[w * MT + threadIdx.x];      // fixed order
        var[p0 + threadIdx.x] = table[2 * DP] - s;
    }
}

template <int MT>
static size_t variance_mma_smem(int T) {
    return ((size_t)MT * var_stride(T) + MT * 16 + MT + (kVarThreads / 32) * MT) * sizeof(double);
}

template <int D, int PPT>
static int launch_linearize(const GpLinParams& prm, int n_bands, int max_T, cudaStream_t s) {
    // one row more than the table: the rotated loop fetches the next row's operands on its last trip too
    size_t smem = (size_t)(tab_hdr(D) + (max_T + 1) * tab_row(D)) * sizeof(double);
    if (smem > 200 * 1024) {
        set_error("kafka_gp_linearize: training table of %zu bytes does not fit shared memory", smem);
        return KAFKA_E_TOOLARGE;
    }
    auto kern = gp_linearize_kernel<D, PPT>;
    KAFKA_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int64_t tiles = (prm.n_pix + kGpThreads * PPT - 1) / (kGpThreads * PPT);
    if (tiles * n_bands > 0x7fffffffLL) {
        set_error("kafka_gp_linearize: %lld pixel tiles x %d bands exceed the grid limit", (long long)tiles, n_bands);
        return KAFKA_E_TOOLARGE;
    }
    kern<<<(unsigned)(tiles * n_bands), kGpThreads, smem, s>>>(prm);
    return launch_status("gp_linearize_kernel");
}

template <int D>
static int dispatch_ppt(const GpLinParams& prm, int n_bands, int max_T, cudaStream_t s) {
    // two pixels per thread halve the shared-memory operand traffic per FMA; only
    // worth it when the grid still fills the machine several times over
    int64_t ctas2 = (prm.n_pix + kGpPptBig * kGpThreads - 1) / (kGpPptBig * kGpThreads) * n_bands;
    bool two = ctas2 >= 8 * (int64_t)num_sms();
    if (KAFKA_GP_FORCE_PPT == 1) two = false;
    if (KAFKA_GP_FORCE_PPT == 2) two = true;
    if (two) return launch_linearize<D, kGpPptBig>(prm, n_bands, max_T, s);
    return launch_linearize<D, 1>(prm, n_bands, max_T, s);
}

#define KAFKA_DISPATCH_D(D_, ...) \
    switch (D_) { \
        case 1: { constexpr int kD = 1; __VA_ARGS__; } break; \
        case 2: { constexpr int kD = 2; __VA_ARGS__; } break; \
        case 3: { constexpr int kD = 3; __VA_ARGS__; } break; \
        case 4: { constexpr int kD = 4; __VA_ARGS__; } break; \
        case 5: { constexpr int kD = 5; __VA_ARGS__; } break; \
        case 6: { constexpr int kD = 6; __VA_ARGS__; } break; \
        case 7: { constexpr int kD = 7; __VA_ARGS__; } break; \
        case 8: { constexpr int kD = 8; __VA_ARGS__; } break; \
        case 9: { constexpr int kD = 9; __VA_ARGS__; } break; \
        case 10: { constexpr int kD = 10; __VA_ARGS__; } break; \
        case 11: { constexpr int kD = 11; __VA_ARGS__; } break; \
        case 12: { constexpr int kD = 12; __VA_ARGS__; } break; \
        case 13: { constexpr int kD = 13; __VA_ARGS__; } break; \
        case 14: { constexpr int kD = 14; __VA_ARGS__; } break; \
        case 15: { constexpr int kD = 15; __VA_ARGS__; } break; \
        case 16: { constexpr int kD = 16; __VA_ARGS__; } break; \
        default: break; \
    }

}  // namespace kafka

using namespace kafka;

extern "C" {

int64_t kafka_gp_table_doubles(int n_train, int n_dim) {
    if (n_train < 1 || n_dim < 1 || n_dim > KAFKA_MAX_DIM) return -1;
    return (int64_t)tab_hdr(n_dim) + (int64_t)n_train * tab_row(n_dim);
}

int kafka_gp_pack_host(const double* inputs, const double* theta, const double* invqt, int T, int D,
                       double* table) {
    KAFKA_CHECK_ARG(inputs && theta && invqt && table, "kafka_gp_pack_host: null pointer");
    KAFKA_CHECK_ARG(T >= 1 && D >= 1 && D <= KAFKA_MAX_DIM, "kafka_gp_pack_host: T=%d D=%d unsupported", T, D);
    for (int j = 0; j < T; ++j)
        KAFKA_CHECK_ARG(fabs(invqt[j]) <= 1.79769313486231570815e308, "kafka_gp_pack_host: weight %d is not finite", j);
    const int DP = dpad(D), HDR = tab_hdr(D), RS = tab_row(D);
    for (int k = 0; k < HDR + T * RS; ++k) table[k] = 0.0;
    double* theta_e = table;
    double* centre = table + DP;
    for (int d = 0; d < D; ++d) {
        theta_e[d] = exp(theta[d]);
        long double s = 0;
        for (int j = 0; j < T; ++j) s += inputs[(size_t)j * D + d];
        centre[d] = (double)(s / T);
    }
    table[2 * DP] = exp(theta[D]);
    table[2 * DP + 1] = exp(theta[D + 1]);
    // cbar = mean ln|alpha| over the non-zero weights: keeps the folded exponents e2 as small as the plain ones
    long double lsum = 0;
    int n_nz = 0;
    for (int j = 0; j < T; ++j)
        if (invqt[j] != 0.0) {
            lsum += log(fabs(invqt[j]));
            ++n_nz;
        }
    double cbar = n_nz ? (double)(lsum / n_nz) : 0.0;
    if (!(fabs(cbar) < 300.0)) cbar = 0.0;
    table[2 * DP + 3] = kL * cbar;
    double wide = 0.0;
    for (int j = 0; j < T; ++j) {
        double* row = table + HDR + (size_t)j * RS;
        double c = 0.0;
        long double cl = 0;                           // the same sum in extended precision, for e2
        for (int d = 0; d < D; ++d) {
            double xc = inputs[(size_t)j * D + d] - centre[d];
            double zh = theta_e[d] * xc;
            c = fma(zh, xc, c);
            cl += (long double)zh * xc;
            row[d] = kL * zh;                         // rows are stored in units of 1/L
        }
        row[DP] = kL * (theta[D] - 0.5 * c);
        row[DP + 1] = invqt[j];
        const double al = invqt[j];
        double lshift = 0.0;                          // ln|alpha| - cbar
        uint64_t sgn = 0;
        if (al != 0.0) {
            const long double ll = logl(fabsl((long double)al)) - (long double)cbar;
            lshift = (double)ll;
            // one rounding in all: an outlying training point puts |e2| at 1e10 and more, where every
            // intermediate rounding in double would show in the kernel value
            row[DP + 2] = (double)((long double)kL * (((long double)theta[D] - 0.5L * cl) + ll));
            sgn = al < 0.0 ? 0x80000000ull : 0;
        } else {
            row[DP + 2] = -536870912.0;               // a zero weight: an exponent far below any underflow
        }
        memcpy(row + DP + 3, &sgn, sizeof(sgn));
        // exponent of this row at any pixel the kernel does not treat as far (sum theta x'^2 < 1200):
        // |a| <= L (|ln s_f| + c/2 + |ln|alpha| - cbar| + sqrt(1200 c)).  The unchecked exp needs |a| < 2^30;
        // 2^29 leaves room for the zero-weight rows' offset of -2^29.
        const double bound = kL * (fabs(theta[D]) + 0.5 * c + fabs(lshift) + sqrt(1200.0 * c));
        if (!(bound < 536870912.0) || fabs(lshift) > 100.0) wide = 1.0;
        // ... and its unshifted value must stay a finite double: with the per-pixel factor exp(-x'.x'/2)
        // hoisted out of the loop the exponent reaches ln s_f + (ln|alpha| - cbar) + max(z.x' - c/2), where
        // z.x' <= sqrt(1200 c) on the near path, i.e. at most +600 at c = 1200.  Past ~ln(DBL_MAX) = 709.8
        // the unchecked exponent insert would wrap; such a table keeps the factor inside the exponent.
        const double zmax = c >= 1200.0 ? 600.0 : sqrt(1200.0 * c) - 0.5 * c;
        if (!(theta[D] + lshift + zmax < 700.0)) wide = 1.0;
    }
    table[2 * DP + 2] = wide;
    return 0;
}

int kafka_gp_linearize(const double* x, int64_t ld, int64_t n_pix, const kafka_gp_band* bands, int n_bands,
                       const uint8_t* obs_mask, double* h0, double* jac, uint32_t flags, void* stream) {
    KAFKA_CHECK_ARG(n_pix >= 0 && ld >= n_pix, "kafka_gp_linearize: need 0 <= n_pix <= ld");
    KAFKA_CHECK_ARG(n_bands >= 1 && n_bands <= KAFKA_MAX_BANDS, "kafka_gp_linearize: n_bands=%d outside 1..%d", n_bands,
                    KAFKA_MAX_BANDS);
    KAFKA_CHECK_ARG(bands, "kafka_gp_linearize: bands is null");
    if (n_pix == 0) return 0;
    KAFKA_CHECK_ARG(x && h0, "kafka_gp_linearize: null pointer");
    GpLinParams prm;
    memset(&prm, 0, sizeof(prm));
    const int D = bands[0].n_dim;
    KAFKA_CHECK_ARG(D >= 1 && D <= KAFKA_MAX_DIM, "kafka_gp_linearize: n_dim=%d outside 1..%d", D, KAFKA_MAX_DIM);
    int max_T = 0;
    for (int b = 0; b < n_bands; ++b) {
        KAFKA_CHECK_ARG(bands[b].n_dim == D, "kafka_gp_linearize: all bands of one call must share n_dim");
        KAFKA_CHECK_ARG(bands[b].table && bands[b].n_train >= 1, "kafka_gp_linearize: band %d has no table", b);
        KAFKA_CHECK_ARG((reinterpret_cast<uintptr_t>(bands[b].table) & 15u) == 0,
                        "kafka_gp_linearize: band %d table must be 16-byte aligned", b);
        prm.bands[b].table = bands[b].table;
        prm.bands[b].n_train = bands[b].n_train;
        for (int d = 0; d < D; ++d) {
            KAFKA_CHECK_ARG(bands[b].mapper[d] >= 0, "kafka_gp_linearize: negative mapper entry");
            prm.bands[b].mapper[d] = bands[b].mapper[d];
        }
        max_T = bands[b].n_train > max_T ? bands[b].n_train : max_T;
    }
    prm.x = x;
    prm.obs_mask = obs_mask;
    prm.h0 = h0;
    prm.jac = jac;
    prm.ld = ld;
    prm.n_pix = n_pix;
    prm.flags = flags;
    prm.n_bands = n_bands;
    int rc = KAFKA_E_UNSUPPORTED;
    KAFKA_DISPATCH_D(D, rc = dispatch_ppt<kD>(prm, n_bands, max_T, (cudaStream_t)stream));
    return rc;
}

int kafka_gp_predict(const double* xin, int64_t ld, int64_t n_pts, const double* table, int n_train, int n_dim,
                     const double* inv_q, double* mean, double* jac, double* var, void* stream) {
    KAFKA_CHECK_ARG(n_dim >= 1 && n_dim <= KAFKA_MAX_DIM, "kafka_gp_predict: n_dim=%d", n_dim);
    KAFKA_CHECK_ARG(!var || inv_q, "kafka_gp_predict: variance requested without inv_q");
    if (n_pts == 0) return 0;
    KAFKA_CHECK_ARG(mean, "kafka_gp_predict: mean is null");
    kafka_gp_band band;
    band.table = table;
    band.n_train = n_train;
    band.n_dim = n_dim;
    for (int d = 0; d < KAFKA_MAX_DIM; ++d) band.mapper[d] = d;
    int rc = kafka_gp_linearize(xin, ld, n_pts, &band, 1, nullptr, mean, jac, 0u, stream);
    if (rc != 0 || !var) return rc;
    // tensor-core tiles when the tile's kernel matrix fits shared memory (64 points up to T ~ 340, 32 up to ~ 750)
    if (!getenv("KAFKA_GP_VARIANCE_NAIVE")) {
        const size_t s64 = variance_mma_smem<64>(n_train), s32 = variance_mma_smem<32>(n_train);
        if (s64 <= 180 * 1024) {
            KAFKA_CUDA(cudaFuncSetAttribute(gp_variance_mma_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)s64));
            gp_variance_mma_kernel<64><<<(unsigned)((n_pts + 63) / 64), kVarThreads, s64, (cudaStream_t)stream>>>(
                xin, ld, n_pts, table, n_train, n_dim, inv_q, var);
            return launch_status("gp_variance_mma_kernel");
        }
        if (s32 <= 200 * 1024) {
            KAFKA_CUDA(cudaFuncSetAttribute(gp_variance_mma_kernel<32>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)s32));
            gp_variance_mma_kernel<32><<<(unsigned)((n_pts + 31) / 32), kVarThreads, s32, (cudaStream_t)stream>>>(
                xin, ld, n_pts, table, n_train, n_dim, inv_q, var);
            return launch_status("gp_variance_mma_kernel");
        }
    }
    size_t smem = (size_t)4 * n_train * sizeof(double);
    KAFKA_CUDA(cudaFuncSetAttribute(gp_variance_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    gp_variance_kernel<<<(unsigned)((n_pts + 3) / 4), 128, smem, (cudaStream_t)stream>>>(xin, ld, n_pts, table, n_train,
                                                                                       n_dim, inv_q, var);
    return launch_status("gp_variance_kernel");
}

int kafka_gp_hessian_correction(const double* x, int64_t ld, int64_t n_pix, int n_params,
                                const kafka_gp_band* bands, int n_bands, const double* innovations,
                                const double* w, double* corr, double* a_inout, void* stream) {
    KAFKA_CHECK_ARG(n_pix >= 0 && ld >= n_pix, "kafka_gp_hessian_correction: need 0 <= n_pix <= ld");
    KAFKA_CHECK_ARG(n_params >= 1 && n_params <= KAFKA_MAX_PARAMS, "kafka_gp_hessian_correction: n_params=%d", n_params);
    KAFKA_CHECK_ARG(n_bands >= 1 && n_bands <= KAFKA_MAX_BANDS && bands, "kafka_gp_hessian_correction: n_bands=%d", n_bands);
    KAFKA_CHECK_ARG(corr || a_inout, "kafka_gp_hessian_correction: no output given");
    if (n_pix == 0) return 0;
    KAFKA_CHECK_ARG(x && innovations && w, "kafka_gp_hessian_correction: null pointer");
    GpHessCorrParams prm;
    memset(&prm, 0, sizeof(prm));
    const int D = bands[0].n_dim;
    KAFKA_CHECK_ARG(D >= 1 && D <= KAFKA_MAX_DIM, "kafka_gp_hessian_correction: n_dim=%d", D);
    int max_T = 0;
    for (int b = 0; b < n_bands; ++b) {
        KAFKA_CHECK_ARG(bands[b].n_dim == D, "kafka_gp_hessian_correction: all bands of one call must share n_dim");
        KAFKA_CHECK_ARG(bands[b].table && bands[b].n_train >= 1, "kafka_gp_hessian_correction: band %d has no table", b);
        prm.bands[b].table = bands[b].table;
        prm.bands[b].n_train = bands[b].n_train;
        uint32_t seen = 0;
        for (int d = 0; d < D; ++d) {
            int slot = bands[b].mapper[d];
            KAFKA_CHECK_ARG(slot >= 0 && slot < n_params && !(seen & (1u << slot)),
                            "kafka_gp_hessian_correction: band %d maps input %d to slot %d", b, d, slot);
            seen |= 1u << slot;
            prm.bands[b].mapper[d] = slot;
        }
        max_T = bands[b].n_train > max_T ? bands[b].n_train : max_T;
    }
    prm.x = x;
    prm.innovations = innovations;
    prm.w = w;
    prm.corr = corr;
    prm.a_inout = a_inout;
    prm.ld = ld;
    prm.n_pix = n_pix;
    prm.n_bands = n_bands;
    prm.n_params = n_params;
    // one row more than the table: the rotated loop fetches the next row's operands on its last trip too
    size_t smem = (size_t)(tab_hdr(D) + (max_T + 1) * tab_row(D)) * sizeof(double);
    if (smem > 200 * 1024) {
        set_error("kafka_gp_hessian_correction: training table of %zu bytes does not fit shared memory", smem);
        return KAFKA_E_TOOLARGE;
    }
    int rc = KAFKA_E_UNSUPPORTED;
    unsigned grid = (unsigned)((n_pix + 127) / 128);
    KAFKA_DISPATCH_D(D, {
        auto kern = gp_hessian_correction_kernel<kD>;
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) {
            set_error("cudaFuncSetAttribute failed: %s", cudaGetErrorString(e));
            rc = (int)e;
        } else {
            kern<<<grid, 128, smem, (cudaStream_t)stream>>>(prm);
            rc = launch_status("gp_hessian_correction_kernel");
        }
    });
    return rc;
}

int kafka_gp_hessian(const double* xin, int64_t ld, int64_t n_pts, const double* table, int n_train, int n_dim,
                     double* hess, void* stream) {
    KAFKA_CHECK_ARG(n_dim >= 1 && n_dim <= KAFKA_MAX_DIM, "kafka_gp_hessian: n_dim=%d", n_dim);
    KAFKA_CHECK_ARG(n_pts >= 0 && ld >= n_pts, "kafka_gp_hessian: need 0 <= n_pts <= ld");
    if (n_pts == 0) return 0;
    KAFKA_CHECK_ARG(xin && table && hess && n_train >= 1, "kafka_gp_hessian: null pointer");
    size_t smem = (size_t)(tab_hdr(n_dim) + n_train * tab_row(n_dim)) * sizeof(double);
    if (smem > 200 * 1024) {
        set_error("kafka_gp_hessian: training table of %zu bytes does not fit shared memory", smem);
        return KAFKA_E_TOOLARGE;
    }
    int rc = KAFKA_E_UNSUPPORTED;
    unsigned grid = (unsigned)((n_pts + kGpThreads - 1) / kGpThreads);
    KAFKA_DISPATCH_D(n_dim, {
        auto kern = gp_hessian_kernel<kD>;
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) {
            set_error("cudaFuncSetAttribute failed: %s", cudaGetErrorString(e));
            rc = (int)e;
        } else {
            kern<<<grid, kGpThreads, smem, (cudaStream_t)stream>>>(xin, ld, n_pts, table, n_train, hess);
            rc = launch_status("gp_hessian_kernel");
        }
    });
    return rc;
}

}  // extern "C"
