// State propagation ("advance") kernels: batched elementwise / per-block work on
// planar state and packed precision blocks (sm_100a, HBM-bound).  Replace the
// propagators of kafka/inference/kf_tools.py:174-353.
#include "chol.cuh"
#include "lu.cuh"

namespace kafka {

constexpr int kAdvThreads = 256;

struct CarryParams {
    double mean[KAFKA_MAX_PARAMS];
    double block[KAFKA_MAX_PARAMS * (KAFKA_MAX_PARAMS + 1) / 2];   // packed prior precision, float32-rounded
};

// propagate_information_filter_LAI (kf_tools.py:292-314), any n_params / slot
__global__ void __launch_bounds__(kAdvThreads)
advance_carry_kernel(const double* __restrict__ x_a, const double* __restrict__ a, const double* __restrict__ q,
                     int64_t ld, int64_t n_pix, int P, int slot, const __grid_constant__ CarryParams cp,
                     double* __restrict__ x_f, double* __restrict__ pinv_f) {
    const int64_t i = (int64_t)blockIdx.x * kAdvThreads + threadIdx.x;
    if (i >= n_pix) return;
    const int cs = tri(slot, slot);
    const double xs = ldg_stream(x_a + (int64_t)slot * ld + i);
    const double d = ldg_stream(a + (int64_t)cs * ld + i);
    const double qq = ldg_stream(q + (int64_t)slot * ld + i);
    const double inflated = 1.0 / ((1.0 / d) + qq);                 // kf_tools.py:308
    for (int p = 0; p < P; ++p) stg_stream(x_f + (int64_t)p * ld + i, p == slot ? xs : cp.mean[p]);
    const int L = tri_count(P);
    for (int c = 0; c < L; ++c)
        stg_stream(pinv_f + (int64_t)c * ld + i, c == cs ? round_f32(inflated) : cp.block[c]);   // float32 block_diag (:312)
}

// propagate_information_filter_approx_SLOW (kf_tools.py:247-289)
__global__ void __launch_bounds__(kAdvThreads)
advance_approx_kernel(const double* __restrict__ a, const double* __restrict__ q, int64_t ld, int64_t n_pix, int P,
                      double* __restrict__ pinv_f) {
    const int64_t i = (int64_t)blockIdx.x * kAdvThreads + threadIdx.x;
    if (i >= n_pix) return;
    for (int r = 0; r < P; ++r) {
        for (int c = 0; c < r; ++c) stg_stream(pinv_f + (int64_t)tri(r, c) * ld + i, 0.0);
        const double m = ldg_stream(a + (int64_t)tri(r, r) * ld + i);
        const double qq = ldg_stream(q + (int64_t)r * ld + i);
        stg_stream(pinv_f + (int64_t)tri(r, r) * ld + i, m * (1.0 / (1.0 + m * qq)));   // M.dot(D), D = 1/(1+m q) (:285-288)
    }
}

// propagate_standard_kalman (kf_tools.py:174-205), diagonal M
__global__ void __launch_bounds__(kAdvThreads)
advance_standard_kernel(const double* __restrict__ x_a, const double* __restrict__ p_a, const double* __restrict__ m,
                        const double* __restrict__ q, int64_t ld, int64_t n_pix, int P, double* __restrict__ x_f,
                        double* __restrict__ p_f) {
    const int64_t i = (int64_t)blockIdx.x * kAdvThreads + threadIdx.x;
    if (i >= n_pix) return;
    for (int p = 0; p < P; ++p) {
        double x = ldg_stream(x_a + (int64_t)p * ld + i);
        stg_stream(x_f + (int64_t)p * ld + i, m ? ldg_stream(m + (int64_t)p * ld + i) * x : x);
    }
    for (int r = 0; r < P; ++r)
        for (int c = 0; c <= r; ++c) {
            double v = ldg_stream(p_a + (int64_t)tri(r, c) * ld + i);
            if (r == c) v += ldg_stream(q + (int64_t)r * ld + i);
            stg_stream(p_f + (int64_t)tri(r, c) * ld + i, v);
        }
}

// propagate_information_filter_SLOW (kf_tools.py:208-245):
//   (I + Pinv Q)^-1 Pinv  =  Pinv - (S Pinv)^T (I + S Pinv S)^-1 (S Pinv),  S = sqrt(Q)
// The inner matrix is symmetric positive definite whenever Pinv is, so one
// Cholesky per pixel replaces the reference's global sparse LU.  When it is not
// (a block the second-order correction left indefinite), the warp solves the
// reference's own system (I + Pinv Q) X = Pinv for that pixel by elimination with
// partial pivoting (lu.cuh); only a singular block is handed on unchanged and counted.
template <int P>
__global__ void __launch_bounds__(128)
advance_exact_kernel(const double* __restrict__ a, const double* __restrict__ q, int64_t ld, int64_t n_pix,
                     double* __restrict__ pinv_f, int32_t* __restrict__ n_bad) {
    constexpr int L = tri_count(P);
    __shared__ LuWork<KAFKA_MAX_PARAMS> s_work[4];
    const int64_t i = (int64_t)blockIdx.x * 128 + threadIdx.x;
    int bad = 0;
    bool failed = false;
    if (i < n_pix) {
        double A[L], M[L], s[P], invd[P];
#pragma unroll
        for (int c = 0; c < L; ++c) A[c] = ldg_stream(a + (int64_t)c * ld + i);
#pragma unroll
        for (int p = 0; p < P; ++p) s[p] = sqrt(ldg_stream(q + (int64_t)p * ld + i));
#pragma unroll
        for (int r = 0; r < P; ++r)
#pragma unroll
            for (int c = 0; c <= r; ++c) M[tri(r, c)] = s[r] * A[tri(r, c)] * s[c] + (r == c ? 1.0 : 0.0);
        const bool ok = chol_factor<P>(M, invd);
        failed = !ok;
        // column by column: w = L^-1 (S Pinv e_col); out -= w w^T accumulated over columns
        double out[L];
#pragma unroll
        for (int c = 0; c < L; ++c) out[c] = A[c];
        // W = L^-1 B with B = S Pinv (P x P); out_rc = Pinv_rc - sum_k W_kr W_kc
        double W[P][P];
#pragma unroll
        for (int col = 0; col < P; ++col) {
#pragma unroll
            for (int k = 0; k < P; ++k) {
                double v = s[k] * A[k >= col ? tri(k, col) : tri(col, k)];
#pragma unroll
                for (int m2 = 0; m2 < k; ++m2) v = fma(-M[tri(k, m2)], W[m2][col], v);
                W[k][col] = v * invd[k];
            }
        }
#pragma unroll
        for (int r = 0; r < P; ++r)
#pragma unroll
            for (int c = 0; c <= r; ++c) {
                double v = out[tri(r, c)];
#pragma unroll
                for (int k = 0; k < P; ++k) v = fma(-W[k][r], W[k][c], v);
                stg_stream(pinv_f + (int64_t)tri(r, c) * ld + i, ok ? v : A[tri(r, c)]);
            }
    }
    unsigned todo = __ballot_sync(0xffffffffu, failed);
    if (todo) {
        const int lane = threadIdx.x & 31;
        LuWork<KAFKA_MAX_PARAMS>& W = s_work[threadIdx.x >> 5];
        const int64_t i0 = (int64_t)blockIdx.x * 128 + (threadIdx.x & ~31);
        while (todo) {
            const int src = __ffs(todo) - 1;
            todo &= todo - 1;
            const int64_t j = i0 + src;
            if (lane < P) {                                         // row r of [I + Pinv Q | Pinv]
                const int r = lane;
                for (int c = 0; c < P; ++c) {
                    const double v = a[(int64_t)(r >= c ? tri(r, c) : tri(c, r)) * ld + j];
                    W.m[r][c] = fma(v, q[(int64_t)c * ld + j], r == c ? 1.0 : 0.0);
                    W.m[r][P + c] = v;
                }
            }
            __syncwarp();
            const bool ok = warp_gepp(W, P, P, lane);
            if (ok && lane < P)
                for (int r = lane; r < P; ++r) pinv_f[(int64_t)tri(r, lane) * ld + j] = W.m[r][P + lane];   // lower triangle of X
            bad += ok ? 0 : 1;
            __syncwarp();
        }
    }
    if ((threadIdx.x & 31) == 0 && bad != 0 && n_bad) atomicAdd(n_bad, bad);
}

}  // namespace kafka

using namespace kafka;

static int check_adv(const char* who, int64_t n_pix, int n_params, int64_t ld) {
    KAFKA_CHECK_ARG(n_pix >= 0 && ld >= n_pix, "%s: need 0 <= n_pix <= ld", who);
    KAFKA_CHECK_ARG(n_params >= 1 && n_params <= KAFKA_MAX_PARAMS, "%s: n_params=%d outside 1..%d", who, n_params,
                    KAFKA_MAX_PARAMS);
    return 0;
}

extern "C" {

int kafka_advance_carry_slot(const double* x_a, const double* a_packed, const double* q, int64_t ld, int64_t n_pix,
                             int n_params, int carry_slot, const double* prior_mean_host,
                             const double* prior_inv_host, double* x_f, double* pinv_f, void* stream) {
    if (int e = check_adv("kafka_advance_carry_slot", n_pix, n_params, ld)) return e;
    KAFKA_CHECK_ARG(carry_slot >= 0 && carry_slot < n_params, "kafka_advance_carry_slot: slot %d outside state", carry_slot);
    KAFKA_CHECK_ARG(prior_mean_host && prior_inv_host, "kafka_advance_carry_slot: null prior");
    if (n_pix == 0) return 0;
    KAFKA_CHECK_ARG(x_a && a_packed && q && x_f && pinv_f, "kafka_advance_carry_slot: null pointer");
    CarryParams cp;
    memset(&cp, 0, sizeof(cp));
    for (int p = 0; p < n_params; ++p) cp.mean[p] = prior_mean_host[p];
    for (int r = 0; r < n_params; ++r)
        for (int c = 0; c <= r; ++c) cp.block[tri(r, c)] = (double)(float)prior_inv_host[r * n_params + c];
    advance_carry_kernel<<<(unsigned)((n_pix + kAdvThreads - 1) / kAdvThreads), kAdvThreads, 0, (cudaStream_t)stream>>>(
        x_a, a_packed, q, ld, n_pix, n_params, carry_slot, cp, x_f, pinv_f);
    return launch_status("advance_carry_kernel");
}

int kafka_advance_approx(const double* a_packed, const double* q, int64_t ld, int64_t n_pix, int n_params,
                         double* pinv_f, void* stream) {
    if (int e = check_adv("kafka_advance_approx", n_pix, n_params, ld)) return e;
    if (n_pix == 0) return 0;
    KAFKA_CHECK_ARG(a_packed && q && pinv_f, "kafka_advance_approx: null pointer");
    advance_approx_kernel<<<(unsigned)((n_pix + kAdvThreads - 1) / kAdvThreads), kAdvThreads, 0, (cudaStream_t)stream>>>(
        a_packed, q, ld, n_pix, n_params, pinv_f);
    return launch_status("advance_approx_kernel");
}

int kafka_advance_standard(const double* x_a, const double* p_packed, const double* m, const double* q, int64_t ld,
                           int64_t n_pix, int n_params, double* x_f, double* p_f, void* stream) {
    if (int e = check_adv("kafka_advance_standard", n_pix, n_params, ld)) return e;
    if (n_pix == 0) return 0;
    KAFKA_CHECK_ARG(x_a && p_packed && q && x_f && p_f, "kafka_advance_standard: null pointer");
    advance_standard_kernel<<<(unsigned)((n_pix + kAdvThreads - 1) / kAdvThreads), kAdvThreads, 0,
                              (cudaStream_t)stream>>>(x_a, p_packed, m, q, ld, n_pix, n_params, x_f, p_f);
    return launch_status("advance_standard_kernel");
}

int kafka_advance_exact(const double* a_packed, const double* q, int64_t ld, int64_t n_pix, int n_params,
                        double* pinv_f, int32_t* n_bad, void* stream) {
    if (int e = check_adv("kafka_advance_exact", n_pix, n_params, ld)) return e;
    cudaStream_t s = (cudaStream_t)stream;
    if (n_bad) KAFKA_CUDA(cudaMemsetAsync(n_bad, 0, sizeof(int32_t), s));
    if (n_pix == 0) return 0;
    KAFKA_CHECK_ARG(a_packed && q && pinv_f, "kafka_advance_exact: null pointer");
    int rc = KAFKA_E_UNSUPPORTED;
    KAFKA_DISPATCH_P(n_params, {
        advance_exact_kernel<kP><<<(unsigned)((n_pix + 127) / 128), 128, 0, s>>>(a_packed, q, ld, n_pix, pinv_f, n_bad);
        rc = launch_status("advance_exact_kernel");
    });
    return rc;
}

}  // extern "C"
