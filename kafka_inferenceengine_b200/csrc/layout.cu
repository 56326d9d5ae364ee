// State-mask compaction, pixel indexing and boundary layout conversion (sm_100a).
// All of it is HBM-bound byte/index work: coalesced loads, shared-memory
// transposes, no tensor cores.
#include <stdarg.h>
#include <string.h>

#include "common.cuh"

namespace kafka {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

// ---------------------------------------------------------------------------
// compaction: count -> scan -> write
// ---------------------------------------------------------------------------
constexpr int kCompactThreads = 256;
constexpr int kCompactPerThread = 16;
constexpr int kCompactTile = kCompactThreads * kCompactPerThread;  // 4096 mask bytes / CTA

__device__ __forceinline__ uint32_t load_mask_bits(const uint8_t* __restrict__ mask, int64_t base,
                                                   int64_t n, bool aligned) {
    // bit k of the result = mask[base + k] != 0, k < 16
    uint32_t bits = 0;
    if (aligned && base + kCompactPerThread <= n) {
        uint4 v = __ldcs(reinterpret_cast<const uint4*>(mask + base));
        uint32_t wds[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int wi = 0; wi < 4; ++wi)
#pragma unroll
            for (int b = 0; b < 4; ++b)
                bits |= (((wds[wi] >> (8 * b)) & 0xffu) != 0u ? 1u : 0u) << (wi * 4 + b);
    } else {
        for (int k = 0; k < kCompactPerThread; ++k)
            if (base + k < n && mask[base + k] != 0) bits |= 1u << k;
    }
    return bits;
}

__global__ void __launch_bounds__(kCompactThreads)
compact_count_kernel(const uint8_t* __restrict__ mask, int64_t n, int32_t* __restrict__ block_counts,
                     bool aligned) {
    __shared__ int warp_counts[kCompactThreads / 32];
    int64_t base = (int64_t)blockIdx.x * kCompactTile + (int64_t)threadIdx.x * kCompactPerThread;
    int c = base < n ? __popc(load_mask_bits(mask, base, n, aligned)) : 0;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
    if ((threadIdx.x & 31) == 0) warp_counts[threadIdx.x >> 5] = c;
    __syncthreads();
    if (threadIdx.x == 0) {
        int t = 0;
        for (int w2 = 0; w2 < kCompactThreads / 32; ++w2) t += warp_counts[w2];
        block_counts[blockIdx.x] = t;
    }
}

// exclusive scan of block_counts[0..n_blocks) in place; total at [n_blocks]
__global__ void __launch_bounds__(1024) compact_scan_kernel(int32_t* counts, int n_blocks) {
    __shared__ int64_t part[1024];
    int per = (n_blocks + 1023) / 1024;
    int lo = threadIdx.x * per, hi = min(lo + per, n_blocks);
    int64_t s = 0;
    for (int k = lo; k < hi; ++k) s += counts[k];
    part[threadIdx.x] = s;
    __syncthreads();
    for (int o = 1; o < 1024; o <<= 1) {   // Hillis-Steele inclusive scan
        int64_t v = threadIdx.x >= o ? part[threadIdx.x - o] : 0;
        __syncthreads();
        part[threadIdx.x] += v;
        __syncthreads();
    }
    int64_t run = threadIdx.x == 0 ? 0 : part[threadIdx.x - 1];
    for (int k = lo; k < hi; ++k) {
        int c = counts[k];
        counts[k] = (int32_t)run;
        run += c;
    }
    if (threadIdx.x == 1023) counts[n_blocks] = (int32_t)part[1023];
}

__global__ void __launch_bounds__(kCompactThreads)
compact_write_kernel(const uint8_t* __restrict__ mask, int64_t n, const int32_t* __restrict__ block_offsets,
                     int32_t* __restrict__ idx, int64_t capacity, bool aligned) {
    __shared__ int warp_counts[kCompactThreads / 32];
    int64_t base = (int64_t)blockIdx.x * kCompactTile + (int64_t)threadIdx.x * kCompactPerThread;
    uint32_t bits = base < n ? load_mask_bits(mask, base, n, aligned) : 0u;
    int c = __popc(bits);
    int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    int incl = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        int v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += v;
    }
    if (lane == 31) warp_counts[wid] = incl;
    __syncthreads();
    int woff = 0;
    for (int w2 = 0; w2 < wid; ++w2) woff += warp_counts[w2];
    int64_t out = (int64_t)block_offsets[blockIdx.x] + woff + incl - c;
    while (bits) {
        int k = __ffs(bits) - 1;
        bits &= bits - 1;
        if (out < capacity) idx[out] = (int32_t)(base + k);
        ++out;
    }
}

__global__ void gather_band_kernel(const double* __restrict__ obs, const double* __restrict__ inv_var,
                                   const uint8_t* __restrict__ obs_mask, const int32_t* __restrict__ idx,
                                   int64_t n_pix, double* __restrict__ y, double* __restrict__ w,
                                   uint8_t* __restrict__ m_out) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_pix) return;
    int32_t j = idx[i];
    bool m = obs_mask[j] != 0;
    y[i] = m ? obs[j] : 0.0;
    w[i] = m ? inv_var[j] : 0.0;   // gate, never multiply: masked pixels hold inf
    if (m_out) m_out[i] = m ? 1 : 0;
}

// all bands of one date in one launch (blockIdx.y = band): for small images the launch is the cost
__global__ void gather_bands_kernel(const double* __restrict__ obs, const double* __restrict__ inv_var,
                                    const uint8_t* __restrict__ obs_mask, int64_t raster_stride,
                                    const int32_t* __restrict__ idx, int64_t n_pix, int64_t ld,
                                    double* __restrict__ y, double* __restrict__ w, uint8_t* __restrict__ m_out) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_pix) return;
    const int64_t src = (int64_t)blockIdx.y * raster_stride + idx[i];
    const int64_t dst = (int64_t)blockIdx.y * ld + i;
    const bool m = obs_mask[src] != 0;
    y[dst] = m ? obs[src] : 0.0;
    w[dst] = m ? inv_var[src] : 0.0;
    if (m_out) m_out[dst] = m ? 1 : 0;
}

// MODIS broadband-albedo ingest (observations.py:288-310).  Products and sums are rounded one by
// one (__dmul_rn / __dadd_rn: no FMA contraction), in numpy's order, so the result is bit-identical
// to the reader's.
template <typename KT>
__global__ void gather_bhr_band_kernel(const KT* __restrict__ kernels, int64_t plane_stride,
                                       const uint8_t* __restrict__ mask, const uint8_t* __restrict__ qa,
                                       const int32_t* __restrict__ idx, int64_t n_pix, double* __restrict__ y,
                                       double* __restrict__ w, uint8_t* __restrict__ m_out) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_pix) return;
    const int32_t j = idx[i];
    const bool m = mask[j] != 0;
    const double k0 = (double)kernels[j], k1 = (double)kernels[plane_stride + j], k2 = (double)kernels[2 * plane_stride + j];
    const double bhr = __dadd_rn(__dadd_rn(__dmul_rn(k0, 1.0), __dmul_rn(k1, 0.189184)), __dmul_rn(k2, -1.377622));
    const uint8_t q = qa[j];
    double sigma = 0.0;                                            // R_mat = zeros_like(bhr)        (:299)
    // np.maximum(2.5e-3, v): unlike fmax it hands a NaN through
    auto np_maximum = [](double v) { return (v > 2.5e-3 || v != v) ? v : 2.5e-3; };
    if (q == 0) sigma = np_maximum(__dmul_rn(bhr, 0.05));          // (:301)
    if (q == 1) sigma = np_maximum(__dmul_rn(bhr, 0.07));          // (:302)
    y[i] = m ? bhr : 0.0;
    w[i] = m ? 1.0 / __dmul_rn(sigma, sigma) : 0.0;                // 1/R^2 (:306); mask set but qa > 1 -> inf, as the reader
    if (m_out) m_out[i] = m ? 1 : 0;
}

template <typename RawT>
__global__ void gather_scaled_band_kernel(const RawT* __restrict__ raw, const uint8_t* __restrict__ ext_mask,
                                          const int32_t* __restrict__ idx, int64_t n_pix, double divisor,
                                          double rel_sigma, double* __restrict__ y, double* __restrict__ w,
                                          uint8_t* __restrict__ m_out) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_pix) return;
    int32_t j = idx[i];
    const RawT r = raw[j];
    bool m = r > (RawT)0;                                        // mask = rho_surface > 0   (:168)
    if (ext_mask) m = m && ext_mask[j] != 0;
    if constexpr (sizeof(RawT) == 4) {
        // a float32 raster stays float32 through the reader's numpy expressions (scalars do not
        // widen an array): every step is rounded in float32, the result widened at the end
        const float rho = m ? __fdiv_rn((float)r, (float)divisor) : 0.0f;
        const float sig = __fmul_rn(rho, (float)rel_sigma);
        y[i] = (double)rho;
        w[i] = m ? (double)__fdiv_rn(1.0f, __fmul_rn(sig, sig)) : 0.0;
    } else {
        const double rho = m ? __ddiv_rn((double)r, divisor) : 0.0;  // rho_surface / 10000.     (:169)
        const double sig = __dmul_rn(rho, rel_sigma);                // R_mat = rho_surface*0.05 (:174)
        y[i] = rho;
        w[i] = m ? __ddiv_rn(1.0, __dmul_rn(sig, sig)) : 0.0;        // 1./(R_mat)**2, gated     (:175-178)
    }
    if (m_out) m_out[i] = m ? 1 : 0;
}

// Sentinel-1 backscatter ingest (Sentinel1_Observations.py:107-130,147-149,176-187): the reader works
// in the raster's own type (float32 from the NetCDF, or float64), so the products are rounded in that
// type, one by one, before the inverse variance is widened to fp64 by setdiag.
template <typename RawT>
__device__ __forceinline__ RawT s1_inverse_variance(RawT unc);
template <>
__device__ __forceinline__ float s1_inverse_variance<float>(float unc) {
    return __fdiv_rn(1.0f, __fmul_rn(unc, unc));
}
template <>
__device__ __forceinline__ double s1_inverse_variance<double>(double unc) {
    return __ddiv_rn(1.0, __dmul_rn(unc, unc));
}
__device__ __forceinline__ float s1_scale(float v, float f) { return __fmul_rn(v, f); }
__device__ __forceinline__ double s1_scale(double v, double f) { return __dmul_rn(v, f); }

template <typename RawT>
__global__ void gather_backscatter_band_kernel(const RawT* __restrict__ raw, const int32_t* __restrict__ idx,
                                               int64_t n_pix, RawT missing, RawT rel_sigma, double* __restrict__ y,
                                               double* __restrict__ w, uint8_t* __restrict__ m_out) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_pix) return;
    const RawT r = raw[idx[i]];
    const bool m = !(r == missing);                               // mask[backscatter == WRONG_VALUE] = False (:148)
    const RawT unc = s1_scale(r, rel_sigma);                      // unc = backscatter*0.05            (:124)
    y[i] = m ? (double)r : 0.0;
    w[i] = m ? (double)s1_inverse_variance<RawT>(unc) : 0.0;     // 1./(R_mat)**2, gated              (:183-187)
    if (m_out) m_out[i] = m ? 1 : 0;
}

__global__ void scatter_output_kernel(const double* __restrict__ x, const double* __restrict__ a,
                                      int64_t ld, int64_t n_pix, int n_params,
                                      const int32_t* __restrict__ idx, int64_t n_raster,
                                      float* __restrict__ mean_r, float* __restrict__ unc_r) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_pix) return;
    int64_t j = idx[i];
    for (int p = 0; p < n_params; ++p) {
        if (mean_r) mean_r[(int64_t)p * n_raster + j] = (float)x[(int64_t)p * ld + i];
        if (unc_r) unc_r[(int64_t)p * n_raster + j] = (float)(1.0 / sqrt(a[(int64_t)tri(p, p) * ld + i]));
    }
}

__global__ void __launch_bounds__(256)
state_uniform_kernel(const double* __restrict__ x, int64_t ld, int64_t n_pix, int P, int32_t* __restrict__ n_diff) {
    int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x;
    int diff = 0;
    if (i < n_pix)
        for (int p = 0; p < P; ++p)
            diff |= __double_as_longlong(ldg_stream(x + (int64_t)p * ld + i)) !=
                    __double_as_longlong(__ldg(x + (int64_t)p * ld));
    diff = __reduce_add_sync(0xffffffffu, diff);
    if ((threadIdx.x & 31) == 0 && diff) atomicAdd(n_diff, diff);     // integer count: order-independent
}

// ---------------------------------------------------------------------------
// interleaved <-> planar through a padded shared-memory tile
// ---------------------------------------------------------------------------
constexpr int kTilePix = 128;

template <bool kToPlanar, bool kDiag>
__global__ void __launch_bounds__(kTilePix)
interleave_kernel(const double* __restrict__ src, double* __restrict__ dst, int64_t n_pix, int P, int64_t ld) {
    extern __shared__ double tile[];          // [kTilePix][P|1]  (odd stride: no bank conflicts)
    const int stride = P | 1;
    int64_t i0 = (int64_t)blockIdx.x * kTilePix;
    int np = (int)min((int64_t)kTilePix, n_pix - i0);
    if (kToPlanar) {
        for (int k = threadIdx.x; k < np * P; k += kTilePix) tile[(k / P) * stride + (k % P)] = src[i0 * P + k];
        __syncthreads();
        if ((int)threadIdx.x < np)
            for (int p = 0; p < P; ++p) dst[(int64_t)p * ld + i0 + threadIdx.x] = tile[threadIdx.x * stride + p];
    } else {
        if ((int)threadIdx.x < np)
            for (int p = 0; p < P; ++p) {
                int64_t plane = kDiag ? tri(p, p) : p;
                tile[threadIdx.x * stride + p] = src[plane * ld + i0 + threadIdx.x];
            }
        __syncthreads();
        for (int k = threadIdx.x; k < np * P; k += kTilePix) dst[i0 * P + k] = tile[(k / P) * stride + (k % P)];
    }
}

constexpr int kBlkPix = 32;

template <bool kToPacked>
__global__ void __launch_bounds__(128)
blocks_kernel(const double* __restrict__ src, double* __restrict__ dst, int64_t n_pix, int P, int64_t ld) {
    extern __shared__ double tile[];          // [kBlkPix][P*P + 1]
    const int PP = P * P, stride = PP + 1, L = tri_count(P);
    int64_t i0 = (int64_t)blockIdx.x * kBlkPix;
    int np = (int)min((int64_t)kBlkPix, n_pix - i0);
    if (kToPacked) {
        for (int k = threadIdx.x; k < np * PP; k += blockDim.x) tile[(k / PP) * stride + (k % PP)] = src[i0 * PP + k];
        __syncthreads();
        for (int k = threadIdx.x; k < L * kBlkPix; k += blockDim.x) {
            int c = k / kBlkPix, px = k % kBlkPix;
            if (px >= np) continue;
            int r = 0;
            while (tri_count(r + 1) <= c) ++r;
            int q = c - tri_count(r);
            dst[(int64_t)c * ld + i0 + px] = tile[px * stride + r * P + q];
        }
    } else {
        for (int k = threadIdx.x; k < L * kBlkPix; k += blockDim.x) {
            int c = k / kBlkPix, px = k % kBlkPix;
            if (px >= np) continue;
            int r = 0;
            while (tri_count(r + 1) <= c) ++r;
            int q = c - tri_count(r);
            double v = src[(int64_t)c * ld + i0 + px];
            tile[px * stride + r * P + q] = v;
            tile[px * stride + q * P + r] = v;
        }
        __syncthreads();
        for (int k = threadIdx.x; k < np * PP; k += blockDim.x) dst[i0 * PP + k] = tile[(k / PP) * stride + (k % PP)];
    }
}

}  // namespace kafka

using namespace kafka;

extern "C" {

const char* kafka_last_error(void) { return g_err; }
int kafka_version(void) { return 100; }

int64_t kafka_compact_scratch_elems(int64_t n_raster) {
    return (n_raster + kCompactTile - 1) / kCompactTile + 2;
}

int kafka_compact_mask(const uint8_t* mask, int64_t n_raster, int32_t* idx, int64_t idx_capacity,
                       int32_t* scratch, int64_t* n_true_host, void* stream) {
    KAFKA_CHECK_ARG(mask && scratch && n_true_host, "kafka_compact_mask: null pointer");
    KAFKA_CHECK_ARG(n_raster >= 0 && n_raster < ((int64_t)1 << 31), "kafka_compact_mask: n_raster=%lld out of range",
                    (long long)n_raster);
    cudaStream_t s = (cudaStream_t)stream;
    *n_true_host = 0;
    if (n_raster == 0) return 0;
    int n_blocks = (int)((n_raster + kCompactTile - 1) / kCompactTile);
    bool aligned = (reinterpret_cast<uintptr_t>(mask) & 15u) == 0;
    compact_count_kernel<<<n_blocks, kCompactThreads, 0, s>>>(mask, n_raster, scratch, aligned);
    compact_scan_kernel<<<1, 1024, 0, s>>>(scratch, n_blocks);
    if (int e = launch_status("compact_count/scan")) return e;
    int32_t total = 0;
    KAFKA_CUDA(cudaMemcpyAsync(&total, scratch + n_blocks, sizeof(int32_t), cudaMemcpyDeviceToHost, s));
    if (idx) {
        compact_write_kernel<<<n_blocks, kCompactThreads, 0, s>>>(mask, n_raster, scratch, idx, idx_capacity, aligned);
        if (int e = launch_status("compact_write")) return e;
    }
    KAFKA_CUDA(cudaStreamSynchronize(s));
    *n_true_host = total;
    if (idx && total > idx_capacity) {
        set_error("kafka_compact_mask: %d True pixels but idx holds %lld", total, (long long)idx_capacity);
        return KAFKA_E_TOOLARGE;
    }
    return 0;
}

int kafka_gather_band(const double* obs, const double* inv_var, const uint8_t* obs_mask,
                      const int32_t* idx, int64_t n_pix, double* y, double* w, uint8_t* m_out,
                      void* stream) {
    KAFKA_CHECK_ARG(n_pix >= 0, "kafka_gather_band: n_pix < 0");
    if (n_pix == 0) return 0;
    KAFKA_CHECK_ARG(obs && inv_var && obs_mask && idx && y && w, "kafka_gather_band: null pointer");
    int threads = 256;
    gather_band_kernel<<<(unsigned)((n_pix + threads - 1) / threads), threads, 0, (cudaStream_t)stream>>>(
        obs, inv_var, obs_mask, idx, n_pix, y, w, m_out);
    return launch_status("gather_band");
}

int kafka_gather_bands(const double* obs, const double* inv_var, const uint8_t* obs_mask, int64_t raster_stride,
                       int n_bands, const int32_t* idx, int64_t n_pix, int64_t ld, double* y, double* w,
                       uint8_t* m_out, void* stream) {
    KAFKA_CHECK_ARG(n_pix >= 0 && ld >= n_pix, "kafka_gather_bands: need 0 <= n_pix <= ld");
    KAFKA_CHECK_ARG(n_bands >= 0 && n_bands <= KAFKA_MAX_BANDS, "kafka_gather_bands: n_bands=%d outside 0..%d", n_bands,
                    KAFKA_MAX_BANDS);
    if (n_pix == 0 || n_bands == 0) return 0;
    KAFKA_CHECK_ARG(obs && inv_var && obs_mask && idx && y && w, "kafka_gather_bands: null pointer");
    const int threads = 256;
    dim3 grid((unsigned)((n_pix + threads - 1) / threads), (unsigned)n_bands);
    gather_bands_kernel<<<grid, threads, 0, (cudaStream_t)stream>>>(obs, inv_var, obs_mask, raster_stride, idx, n_pix, ld,
                                                                    y, w, m_out);
    return launch_status("gather_bands");
}

int kafka_gather_bhr_band(const void* kernels, int kernel_type, int64_t plane_stride, const uint8_t* mask,
                          const uint8_t* qa, const int32_t* idx, int64_t n_pix, double* y, double* w, uint8_t* m_out,
                          void* stream) {
    KAFKA_CHECK_ARG(n_pix >= 0, "kafka_gather_bhr_band: n_pix < 0");
    if (n_pix == 0) return 0;
    KAFKA_CHECK_ARG(kernels && mask && qa && idx && y && w, "kafka_gather_bhr_band: null pointer");
    KAFKA_CHECK_ARG(plane_stride > 0, "kafka_gather_bhr_band: plane_stride must be positive");
    const int threads = 256;
    const unsigned grid = (unsigned)((n_pix + threads - 1) / threads);
    cudaStream_t s = (cudaStream_t)stream;
    if (kernel_type == KAFKA_RAW_F32)
        gather_bhr_band_kernel<float><<<grid, threads, 0, s>>>((const float*)kernels, plane_stride, mask, qa, idx, n_pix, y, w, m_out);
    else if (kernel_type == KAFKA_RAW_F64)
        gather_bhr_band_kernel<double><<<grid, threads, 0, s>>>((const double*)kernels, plane_stride, mask, qa, idx, n_pix, y, w, m_out);
    else {
        set_error("kafka_gather_bhr_band: kernel_type %d is neither KAFKA_RAW_F32 nor KAFKA_RAW_F64", kernel_type);
        return KAFKA_E_BADARG;
    }
    return launch_status("gather_bhr_band");
}

int kafka_gather_scaled_band(const void* raw, int raw_type, const uint8_t* ext_mask, const int32_t* idx,
                             int64_t n_pix, double divisor, double rel_sigma, double* y, double* w,
                             uint8_t* m_out, void* stream) {
    KAFKA_CHECK_ARG(n_pix >= 0, "kafka_gather_scaled_band: n_pix < 0");
    KAFKA_CHECK_ARG(divisor > 0.0 && rel_sigma > 0.0, "kafka_gather_scaled_band: divisor and rel_sigma must be > 0");
    if (n_pix == 0) return 0;
    KAFKA_CHECK_ARG(raw && idx && y && w, "kafka_gather_scaled_band: null pointer");
    const int threads = 256;
    const unsigned grid = (unsigned)((n_pix + threads - 1) / threads);
    cudaStream_t s = (cudaStream_t)stream;
    switch (raw_type) {
        case KAFKA_RAW_I16:
            gather_scaled_band_kernel<int16_t><<<grid, threads, 0, s>>>((const int16_t*)raw, ext_mask, idx, n_pix, divisor, rel_sigma, y, w, m_out);
            break;
        case KAFKA_RAW_U16:
            gather_scaled_band_kernel<uint16_t><<<grid, threads, 0, s>>>((const uint16_t*)raw, ext_mask, idx, n_pix, divisor, rel_sigma, y, w, m_out);
            break;
        case KAFKA_RAW_F32:
            gather_scaled_band_kernel<float><<<grid, threads, 0, s>>>((const float*)raw, ext_mask, idx, n_pix, divisor, rel_sigma, y, w, m_out);
            break;
        default:
            set_error("kafka_gather_scaled_band: unknown raw_type %d", raw_type);
            return KAFKA_E_BADARG;
    }
    return launch_status("gather_scaled_band_kernel");
}

int kafka_gather_backscatter_band(const void* raw, int raw_type, const int32_t* idx, int64_t n_pix,
                                  double missing_value, double rel_sigma, double* y, double* w, uint8_t* m_out,
                                  void* stream) {
    KAFKA_CHECK_ARG(n_pix >= 0, "kafka_gather_backscatter_band: n_pix < 0");
    KAFKA_CHECK_ARG(rel_sigma > 0.0, "kafka_gather_backscatter_band: rel_sigma must be > 0");
    if (n_pix == 0) return 0;
    KAFKA_CHECK_ARG(raw && idx && y && w, "kafka_gather_backscatter_band: null pointer");
    const int threads = 256;
    const unsigned grid = (unsigned)((n_pix + threads - 1) / threads);
    cudaStream_t s = (cudaStream_t)stream;
    if (raw_type == KAFKA_RAW_F32)
        gather_backscatter_band_kernel<float><<<grid, threads, 0, s>>>((const float*)raw, idx, n_pix, (float)missing_value,
                                                                       (float)rel_sigma, y, w, m_out);
    else if (raw_type == KAFKA_RAW_F64)
        gather_backscatter_band_kernel<double><<<grid, threads, 0, s>>>((const double*)raw, idx, n_pix, missing_value,
                                                                        rel_sigma, y, w, m_out);
    else {
        set_error("kafka_gather_backscatter_band: raw_type %d is neither KAFKA_RAW_F32 nor KAFKA_RAW_F64", raw_type);
        return KAFKA_E_BADARG;
    }
    return launch_status("gather_backscatter_band");
}

int kafka_scatter_output(const double* x, const double* a_packed, int64_t ld, int64_t n_pix,
                         int n_params, const int32_t* idx, int64_t n_raster, float* mean_raster,
                         float* unc_raster, void* stream) {
    KAFKA_CHECK_ARG(n_pix >= 0 && ld >= n_pix, "kafka_scatter_output: bad n_pix/ld");
    KAFKA_CHECK_ARG(n_params >= 1 && n_params <= KAFKA_MAX_PARAMS, "kafka_scatter_output: n_params=%d", n_params);
    if (n_pix == 0) return 0;
    KAFKA_CHECK_ARG(idx && (!mean_raster || x) && (!unc_raster || a_packed), "kafka_scatter_output: null pointer");
    int threads = 256;
    scatter_output_kernel<<<(unsigned)((n_pix + threads - 1) / threads), threads, 0, (cudaStream_t)stream>>>(
        x, a_packed, ld, n_pix, n_params, idx, n_raster, mean_raster, unc_raster);
    return launch_status("scatter_output");
}

int kafka_output_rasters(const double* x, const double* a_packed, int64_t ld, int64_t n_pix, int n_params,
                         const int32_t* idx, int64_t n_raster, float* mean_raster, float* unc_raster, int zero_first,
                         void* host_dst, void* stream) {
    KAFKA_CHECK_ARG(mean_raster && n_raster >= 0, "kafka_output_rasters: null raster");
    KAFKA_CHECK_ARG(n_params >= 1 && n_params <= KAFKA_MAX_PARAMS, "kafka_output_rasters: n_params=%d", n_params);
    KAFKA_CHECK_ARG(!unc_raster || unc_raster == mean_raster + (int64_t)n_params * n_raster,
                    "kafka_output_rasters: the two rasters must be one allocation, mean first");
    const size_t bytes = (size_t)(unc_raster ? 2 : 1) * n_params * n_raster * sizeof(float);
    if (zero_first && bytes) KAFKA_CUDA(cudaMemsetAsync(mean_raster, 0, bytes, (cudaStream_t)stream));
    int rc = kafka_scatter_output(x, a_packed, ld, n_pix, n_params, idx, n_raster, mean_raster, unc_raster, stream);
    if (rc != 0) return rc;
    if (host_dst && bytes)
        KAFKA_CUDA(cudaMemcpyAsync(host_dst, mean_raster, bytes, cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    return 0;
}

int kafka_state_uniform(const double* x, int64_t ld, int64_t n_pix, int n_params, int32_t* n_diff, void* stream) {
    KAFKA_CHECK_ARG(n_pix >= 0 && ld >= n_pix, "kafka_state_uniform: need 0 <= n_pix <= ld");
    KAFKA_CHECK_ARG(n_params >= 1 && n_params <= KAFKA_MAX_PARAMS, "kafka_state_uniform: n_params=%d", n_params);
    KAFKA_CHECK_ARG(n_diff, "kafka_state_uniform: n_diff is null");
    KAFKA_CUDA(cudaMemsetAsync(n_diff, 0, sizeof(int32_t), (cudaStream_t)stream));
    if (n_pix == 0) return 0;
    KAFKA_CHECK_ARG(x, "kafka_state_uniform: x is null");
    state_uniform_kernel<<<(unsigned)((n_pix + 255) / 256), 256, 0, (cudaStream_t)stream>>>(x, ld, n_pix, n_params, n_diff);
    return launch_status("state_uniform_kernel");
}

// A few words from device memory straight into pinned host memory, written by the SMs over
// PCIe instead of by a copy engine: a scalar read-back queued as a memcpy would wait behind
// whatever large device->host transfer (the previous date's output rasters) occupies the engine.
__global__ void store_host_words_kernel(const uint64_t* __restrict__ src, uint64_t* __restrict__ dst, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) dst[i] = src[i];
    __threadfence_system();
}

int kafka_read_small(const void* src_dev, void* dst_pinned_host, int64_t n_bytes, void* stream) {
    KAFKA_CHECK_ARG(src_dev && dst_pinned_host, "kafka_read_small: null pointer");
    KAFKA_CHECK_ARG(n_bytes > 0 && n_bytes <= 65536 && n_bytes % 8 == 0,
                    "kafka_read_small: n_bytes=%lld must be a multiple of 8 in 8..65536", (long long)n_bytes);
    KAFKA_CHECK_ARG(((uintptr_t)src_dev | (uintptr_t)dst_pinned_host) % 8 == 0, "kafka_read_small: pointers must be 8-byte aligned");
    void* dst_dev = nullptr;
    KAFKA_CUDA(cudaHostGetDevicePointer(&dst_dev, dst_pinned_host, 0));     // fails for pageable memory
    const int n = (int)(n_bytes / 8);
    store_host_words_kernel<<<(n + 255) / 256, 256, 0, (cudaStream_t)stream>>>((const uint64_t*)src_dev, (uint64_t*)dst_dev, n);
    return launch_status("store_host_words_kernel");
}

static int check_layout(const char* who, int64_t n_pix, int n_params, int64_t ld) {
    KAFKA_CHECK_ARG(n_pix >= 0 && ld >= n_pix, "%s: need 0 <= n_pix <= ld (n_pix=%lld ld=%lld)", who,
                    (long long)n_pix, (long long)ld);
    KAFKA_CHECK_ARG(n_params >= 1 && n_params <= KAFKA_MAX_PARAMS, "%s: n_params=%d outside 1..%d", who, n_params,
                    KAFKA_MAX_PARAMS);
    return 0;
}

int kafka_interleaved_to_planar(const double* x_il, int64_t n_pix, int n_params, double* x_pl, int64_t ld,
                                void* stream) {
    if (int e = check_layout("kafka_interleaved_to_planar", n_pix, n_params, ld)) return e;
    if (n_pix == 0) return 0;
    KAFKA_CHECK_ARG(x_il && x_pl, "kafka_interleaved_to_planar: null pointer");
    size_t smem = (size_t)kTilePix * (n_params | 1) * sizeof(double);
    interleave_kernel<true, false><<<(unsigned)((n_pix + kTilePix - 1) / kTilePix), kTilePix, smem, (cudaStream_t)stream>>>(
        x_il, x_pl, n_pix, n_params, ld);
    return launch_status("interleaved_to_planar");
}

int kafka_planar_to_interleaved(const double* x_pl, int64_t ld, int64_t n_pix, int n_params, double* x_il,
                                void* stream) {
    if (int e = check_layout("kafka_planar_to_interleaved", n_pix, n_params, ld)) return e;
    if (n_pix == 0) return 0;
    KAFKA_CHECK_ARG(x_il && x_pl, "kafka_planar_to_interleaved: null pointer");
    size_t smem = (size_t)kTilePix * (n_params | 1) * sizeof(double);
    interleave_kernel<false, false><<<(unsigned)((n_pix + kTilePix - 1) / kTilePix), kTilePix, smem, (cudaStream_t)stream>>>(
        x_pl, x_il, n_pix, n_params, ld);
    return launch_status("planar_to_interleaved");
}

int kafka_packed_diagonal(const double* a_packed, int64_t ld, int64_t n_pix, int n_params, double* diag_il,
                          void* stream) {
    if (int e = check_layout("kafka_packed_diagonal", n_pix, n_params, ld)) return e;
    if (n_pix == 0) return 0;
    KAFKA_CHECK_ARG(a_packed && diag_il, "kafka_packed_diagonal: null pointer");
    size_t smem = (size_t)kTilePix * (n_params | 1) * sizeof(double);
    interleave_kernel<false, true><<<(unsigned)((n_pix + kTilePix - 1) / kTilePix), kTilePix, smem, (cudaStream_t)stream>>>(
        a_packed, diag_il, n_pix, n_params, ld);
    return launch_status("packed_diagonal");
}

static int blocks_launch(bool to_packed, const double* src, double* dst, int64_t n_pix, int P, int64_t ld,
                         cudaStream_t s) {
    size_t smem = (size_t)kBlkPix * (P * P + 1) * sizeof(double);
    unsigned grid = (unsigned)((n_pix + kBlkPix - 1) / kBlkPix);
    if (to_packed) {
        KAFKA_CUDA(cudaFuncSetAttribute(blocks_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        blocks_kernel<true><<<grid, 128, smem, s>>>(src, dst, n_pix, P, ld);
    } else {
        KAFKA_CUDA(cudaFuncSetAttribute(blocks_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        blocks_kernel<false><<<grid, 128, smem, s>>>(src, dst, n_pix, P, ld);
    }
    return launch_status("blocks<->packed");
}

int kafka_blocks_to_packed(const double* blocks, int64_t n_pix, int n_params, double* a_packed, int64_t ld,
                           void* stream) {
    if (int e = check_layout("kafka_blocks_to_packed", n_pix, n_params, ld)) return e;
    if (n_pix == 0) return 0;
    KAFKA_CHECK_ARG(blocks && a_packed, "kafka_blocks_to_packed: null pointer");
    return blocks_launch(true, blocks, a_packed, n_pix, n_params, ld, (cudaStream_t)stream);
}

int kafka_packed_to_blocks(const double* a_packed, int64_t ld, int64_t n_pix, int n_params, double* blocks,
                           void* stream) {
    if (int e = check_layout("kafka_packed_to_blocks", n_pix, n_params, ld)) return e;
    if (n_pix == 0) return 0;
    KAFKA_CHECK_ARG(blocks && a_packed, "kafka_packed_to_blocks: null pointer");
    return blocks_launch(false, a_packed, blocks, n_pix, n_params, ld, (cudaStream_t)stream);
}

}  // extern "C"
