// Warp-cooperative Gaussian elimination with partial pivoting on a small dense system held in shared
// memory: the fall-back for per-pixel blocks that are not positive definite, where the reference's sparse
// LU (scipy spsolve / splu: kafka/inference/solvers.py:133-134, kf_tools.py:93-94, kf_tools.py:236-243)
// still returns the solution of the equations and a Cholesky factorisation has none to give.  Rare work,
// laid out for a small footprint: one warp, one system at a time, lane r owns row r during the
// elimination, lane c owns right-hand side c during the back substitution.
#pragma once
#include "common.cuh"

namespace kafka {

template <int NRHS>
struct LuWork {
    static constexpr int kStride = KAFKA_MAX_PARAMS + NRHS;
    double m[KAFKA_MAX_PARAMS][kStride];     // [A | B]: P x P matrix, then n_rhs columns
};

// Solves A X = B in place (X replaces B).  Returns false -- to every lane -- when a pivot is zero or not a
// number, or a component of the solution is not finite.  All 32 lanes must call it.
template <int NRHS>
__device__ __forceinline__ bool warp_gepp(LuWork<NRHS>& W, int P, int n_rhs, int lane) {
    bool ok = true;
    const int width = P + n_rhs;
    for (int k = 0; k < P; ++k) {
        // pivot: the first row of maximal |entry| in column k, rows k..P-1 (what idamax picks)
        double best = (lane >= k && lane < P) ? fabs(W.m[lane][k]) : -1.0;
        if (best != best) best = INFINITY;                           // a NaN wins and fails the solve below
        int who = lane;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const double ob = __shfl_xor_sync(0xffffffffu, best, o);
            const int ow = __shfl_xor_sync(0xffffffffu, who, o);
            if (ob > best || (ob == best && ow < who)) {
                best = ob;
                who = ow;
            }
        }
        if (!(best > 0.0) || best == INFINITY) ok = false;
        if (who != k && lane < width) {                              // lane q swaps column q of the two rows
            const double t = W.m[k][lane];
            W.m[k][lane] = W.m[who][lane];
            W.m[who][lane] = t;
        }
        __syncwarp();
        if (lane > k && lane < P) {
            const double f = W.m[lane][k] / W.m[k][k];
            for (int q = k + 1; q < width; ++q) W.m[lane][q] = fma(-f, W.m[k][q], W.m[lane][q]);
        }
        __syncwarp();
    }
    bool finite = true;
    if (lane < n_rhs) {                                              // back substitution, one lane per right-hand side
        const int c = P + lane;
        for (int k = P - 1; k >= 0; --k) {
            double v = W.m[k][c];
            for (int q = k + 1; q < P; ++q) v = fma(-W.m[k][q], W.m[q][c], v);
            v /= W.m[k][k];
            W.m[k][c] = v;
            finite = finite && (v - v == 0.0);
        }
    }
    __syncwarp();
    return ok && __all_sync(0xffffffffu, finite);
}

}  // namespace kafka
