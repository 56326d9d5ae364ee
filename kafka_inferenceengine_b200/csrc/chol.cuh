// Per-thread fp64 Cholesky on a packed lower-triangular P x P block held in
// registers (P is a compile-time constant so every index is static).
#pragma once
#include "common.cuh"

namespace kafka {

// In place A -> L (A = L L^T), diagonal stored as 1/L_jj in invd.  Returns false
// when the block is not positive definite.
// All loops run over the full compile-time range 0..P-1 with a guard instead of a
// data-dependent bound: after unrolling every guard folds away and every index is a
// literal (with triangular bounds the compiler refuses to unroll some sizes and drops the
// block into local memory).
template <int P>
__device__ __forceinline__ bool chol_factor(double (&A)[tri_count(P)], double (&invd)[P]) {
    bool ok = true;
#pragma unroll
    for (int j = 0; j < P; ++j) {
        double d = A[tri(j, j)];
#pragma unroll
        for (int k = 0; k < P; ++k)
            if (k < j) d = fma(-A[tri(j, k)], A[tri(j, k)], d);
        ok = ok && (d > 0.0);
        double inv = rsqrt(d);   // <= 1 ulp in fp64
        invd[j] = inv;
        A[tri(j, j)] = d * inv;
#pragma unroll
        for (int i = 0; i < P; ++i) {
            if (i > j) {
                double s = A[tri(i, j)];
#pragma unroll
                for (int k = 0; k < P; ++k)
                    if (k < j) s = fma(-A[tri(i, k)], A[tri(j, k)], s);
                A[tri(i, j)] = s * inv;
            }
        }
    }
    return ok;
}

// Solve L L^T x = b in place.
template <int P>
__device__ __forceinline__ void chol_solve(const double (&Lm)[tri_count(P)], const double (&invd)[P], double (&b)[P]) {
#pragma unroll
    for (int i = 0; i < P; ++i) {
        double s = b[i];
#pragma unroll
        for (int k = 0; k < P; ++k)
            if (k < i) s = fma(-Lm[tri(i, k)], b[k], s);
        b[i] = s * invd[i];
    }
#pragma unroll
    for (int ii = 0; ii < P; ++ii) {
        const int i = P - 1 - ii;
        double s = b[i];
#pragma unroll
        for (int k = 0; k < P; ++k)
            if (k > i) s = fma(-Lm[tri(k, i)], b[k], s);
        b[i] = s * invd[i];
    }
}

// y = S x for a packed symmetric block.
template <int P>
__device__ __forceinline__ void symv(const double (&S)[tri_count(P)], const double (&x)[P], double (&y)[P]) {
#pragma unroll
    for (int r = 0; r < P; ++r) {
        double s = 0.0;
#pragma unroll
        for (int q = 0; q < P; ++q) s = fma(S[r >= q ? tri(r, q) : tri(q, r)], x[q], s);
        y[r] = s;
    }
}

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

}  // namespace kafka
