// Per-thread fp64 Cholesky on a packed lower-triangular P x P block held in
// registers (P is a compile-time constant so every index is static).
#pragma once
#include "common.cuh"

namespace kafka {

// In place A -> L (A = L L^T), diagonal stored as 1/L_jj in invd.  Returns false
// when the block is not positive definite.
// Storage of one packed block.  RegBlock keeps every entry in registers; SplitBlock keeps the
// first R entries in registers and the rest in the thread's own shared-memory column (stride
// `STRIDE` doubles), for state sizes whose block does not fit the register file (n_params
// 13..16).  All indices are literals after unrolling, so `c < R` folds at compile time.
template <int L>
struct RegBlock {
    double v[L];
    __device__ __forceinline__ double get(int c) const { return v[c]; }
    __device__ __forceinline__ void set(int c, double x) { v[c] = x; }
};

template <int L, int R, int STRIDE>
struct SplitBlock {
    double v[R];
    // this thread's column: entry c >= R at sm[(c - R) * STRIDE].  volatile on purpose: it keeps
    // the compiler from hoisting dozens of these loads ahead of their use, which is what made
    // the unrolled factorisation spill.
    volatile double* sm;
    __device__ __forceinline__ double get(int c) const { return c < R ? v[c < R ? c : 0] : sm[(c - R) * STRIDE]; }
    __device__ __forceinline__ void set(int c, double x) {
        if (c < R)
            v[c < R ? c : 0] = x;
        else
            sm[(c - R) * STRIDE] = x;
    }
};

// Right-looking (outer-product) Cholesky: once column j is known, the whole trailing block is
// updated with INDEPENDENT FMAs, so the dependent chain per column is rsqrt -> scale -> one
// FMA instead of a j-term dot product.  With 8 warps per SM the solve kernels are bound by
// instruction latency, and this halves the critical path of the factorisation.
// All loops run over the full compile-time range 0..P-1 with a guard instead of a
// data-dependent bound: after unrolling every guard folds away and every index is a literal
// (with triangular bounds the compiler refuses to unroll some sizes and drops the block into
// local memory).  A -> L in place (A = L L^T), diagonal stored as 1/L_jj in invd; returns
// false when the block is not positive definite.
template <int P, typename Blk>
__device__ __forceinline__ bool chol_factor_blk(Blk& A, double (&invd)[P]) {
    bool ok = true;
#pragma unroll
    for (int j = 0; j < P; ++j) {
        const double d = A.get(tri(j, j));
        ok = ok && (d > 0.0);
        const double inv = rsqrt(d);   // <= 1 ulp in fp64
        invd[j] = inv;
        A.set(tri(j, j), d * inv);
#pragma unroll
        for (int i = 0; i < P; ++i)
            if (i > j) A.set(tri(i, j), A.get(tri(i, j)) * inv);
#pragma unroll
        for (int i = 0; i < P; ++i)
#pragma unroll
            for (int k = 0; k < P; ++k)
                if (i > j && k > j && k <= i)
                    A.set(tri(i, k), fma(-A.get(tri(i, j)), A.get(tri(k, j)), A.get(tri(i, k))));
    }
    return ok;
}

// Solve L L^T x = b in place (column-oriented substitutions, same reasoning).
template <int P, typename Blk>
__device__ __forceinline__ void chol_solve_blk(const Blk& Lm, const double (&invd)[P], double (&b)[P]) {
#pragma unroll
    for (int k = 0; k < P; ++k) {
        b[k] *= invd[k];
#pragma unroll
        for (int i = 0; i < P; ++i)
            if (i > k) b[i] = fma(-Lm.get(tri(i, k)), b[k], b[i]);
    }
#pragma unroll
    for (int kk = 0; kk < P; ++kk) {
        const int k = P - 1 - kk;
        b[k] *= invd[k];
#pragma unroll
        for (int i = 0; i < P; ++i)
            if (i < k) b[i] = fma(-Lm.get(tri(k, i)), b[k], b[i]);
    }
}

// y = S x for a packed symmetric block.
template <int P, typename Blk>
__device__ __forceinline__ void symv_blk(const Blk& S, const double (&x)[P], double (&y)[P]) {
#pragma unroll
    for (int r = 0; r < P; ++r) {
        double s = 0.0;
#pragma unroll
        for (int q = 0; q < P; ++q) s = fma(S.get(r >= q ? tri(r, q) : tri(q, r)), x[q], s);
        y[r] = s;
    }
}

// plain-array front ends (blocks that live entirely in registers)
template <int P>
__device__ __forceinline__ bool chol_factor(double (&A)[tri_count(P)], double (&invd)[P]) {
    return chol_factor_blk<P>(reinterpret_cast<RegBlock<tri_count(P)>&>(A), invd);
}
template <int P>
__device__ __forceinline__ void chol_solve(const double (&Lm)[tri_count(P)], const double (&invd)[P], double (&b)[P]) {
    chol_solve_blk<P>(reinterpret_cast<const RegBlock<tri_count(P)>&>(Lm), invd, b);
}
template <int P>
__device__ __forceinline__ void symv(const double (&S)[tri_count(P)], const double (&x)[P], double (&y)[P]) {
    symv_blk<P>(reinterpret_cast<const RegBlock<tri_count(P)>&>(S), x, y);
}

#ifdef KAFKA_DEV_ONLY_P      /* development builds: instantiate one state size only (fast compile, SASS inspection) */
#define KAFKA_DISPATCH_P(P_, ...) \
    switch (P_) { \
        case KAFKA_DEV_ONLY_P: { constexpr int kP = KAFKA_DEV_ONLY_P; __VA_ARGS__; } break; \
        default: break; \
    }
#else
#define KAFKA_DISPATCH_P(P_, ...) \
    switch (P_) { \
        case 1: { constexpr int kP = 1; __VA_ARGS__; } break; \
        case 2: { constexpr int kP = 2; __VA_ARGS__; } break; \
        case 3: { constexpr int kP = 3; __VA_ARGS__; } break; \
        case 4: { constexpr int kP = 4; __VA_ARGS__; } break; \
        case 5: { constexpr int kP = 5; __VA_ARGS__; } break; \
        case 6: { constexpr int kP = 6; __VA_ARGS__; } break; \
        case 7: { constexpr int kP = 7; __VA_ARGS__; } break; \
        case 8: { constexpr int kP = 8; __VA_ARGS__; } break; \
        case 9: { constexpr int kP = 9; __VA_ARGS__; } break; \
        case 10: { constexpr int kP = 10; __VA_ARGS__; } break; \
        case 11: { constexpr int kP = 11; __VA_ARGS__; } break; \
        case 12: { constexpr int kP = 12; __VA_ARGS__; } break; \
        case 13: { constexpr int kP = 13; __VA_ARGS__; } break; \
        case 14: { constexpr int kP = 14; __VA_ARGS__; } break; \
        case 15: { constexpr int kP = 15; __VA_ARGS__; } break; \
        case 16: { constexpr int kP = 16; __VA_ARGS__; } break; \
        default: break; \
    }
#endif

}  // namespace kafka
