/*
 * kafka_b200.h -- C ABI of libkafka_b200.so: the B200 (sm_100a) implementation of
 * KaFKA's per-pixel assimilation path.
 *
 * The reference (QCDIS/KaFKA-InferenceEngine) is pure Python: there is no FFI
 * to mirror, so every entry point below names the reference FUNCTION it
 * replaces (file:line relative to the reference root).  INTEGRATION.md shows the
 * ctypes binding a maintainer would add to the reference.
 *
 * Conventions
 *  - plain pointers and sizes only; every data pointer is a DEVICE pointer into
 *    caller-owned memory unless the name ends in _host;
 *  - `stream` is a cudaStream_t passed as void* (NULL = default stream);
 *  - return value: 0 = OK, <0 = bad argument (KAFKA_E_*), >0 = cudaError_t;
 *    kafka_last_error() gives the text for the calling thread;
 *  - nothing here allocates or frees device memory; scratch is passed in.
 *
 * Device layouts ("planar", pixel index fastest; `ld` = elements between planes,
 * ld >= n_pix):
 *  - state vector      x[p*ld + i]            p < n_params           (fp64)
 *  - precision blocks  a[c*ld + i]            c = r(r+1)/2 + q, q<=r (fp64,
 *                      packed lower triangle of the symmetric P x P block)
 *  - per-band vectors  y[b*ld + i], w[b*ld + i], h0[b*ld + i]
 *  - Jacobians         jac[(b*n_dim + d)*ld + i]   d = emulator input dimension
 * The reference's interleaved vector x[n_params*i + p]
 * (kafka/inference/utils.py:201) and its block-diagonal sparse matrices are
 * converted at the boundary by the layout functions below.
 */
#ifndef KAFKA_B200_H
#define KAFKA_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define KAFKA_MAX_PARAMS 16
#define KAFKA_MAX_BANDS 16
#define KAFKA_MAX_DIM 16

#define KAFKA_E_BADARG (-1)
#define KAFKA_E_UNSUPPORTED (-2)
#define KAFKA_E_TOOLARGE (-3)

/* band operator kinds for kafka_solve_bands */
#define KAFKA_BAND_ARRAYS 0 /* h0/jac arrays (emulator output or supplied Jacobians) */
#define KAFKA_BAND_LINEAR 1 /* constant row: H0 = row.x_prev + offset, J = row        */
#define KAFKA_BAND_UNIFORM 2 /* ONE linearisation shared by all pixels: h0[0], jac[d];   \
                                the reference's "one unique state vector" shortcut      \
                                (kafka/inference/utils.py:72-74)                        */

#define KAFKA_BAND_WCM 3     /* Water-Cloud SAR backscatter model, closed form:          \
                                row[0..4] = A,B,C,D,E, row[5] = cos(incidence angle),    \
                                mapper[0] = state slot of the vegetation descriptor      \
                                (LAI), mapper[1] = slot of soil moisture                 \
                                (kafka/observation_operators/sar_forward_model.py:13-113) */

/* raw raster element types for kafka_gather_scaled_band */
#define KAFKA_RAW_I16 0
#define KAFKA_RAW_U16 1
#define KAFKA_RAW_F32 2
#define KAFKA_RAW_F64 3 /* kafka_gather_bhr_band and kafka_gather_backscatter_band only */

/* flags */
#define KAFKA_ROUND_F32 1u      /* round H0/J to float32 as the reference stores them \
                                   (kafka/inference/utils.py:193-195)                  */
#define KAFKA_PRIOR_BROADCAST 2u /* pinv_f is ONE packed block shared by all pixels    */
#define KAFKA_SOLVE_PAIR 4u       /* kafka_solve_bands: use the two-lanes-per-pixel kernel at any size
                                    (the default from 14 parameters; for tests) */
#define KAFKA_SOLVE_COUNT_LU 8u  /* kafka_solve_bands: n_bad points at TWO int32: [0] pixels left at the
                                    forecast, [1] pixels solved by pivoted elimination */

const char* kafka_last_error(void);
int kafka_version(void);

/* ---- a1: state-mask compaction and pixel indexing ------------------------
 * Replaces `state_mask.sum()` (kafka/linear_kf.py:79-80) and the boolean-mask
 * gathers `arr[state_mask]` (kafka/inference/solvers.py:90-92).  idx[k] = linear
 * (row-major) raster index of the k-th True pixel.  `scratch` holds at least
 * kafka_compact_scratch_elems(n_raster) int32; *n_true_host receives the count
 * (the call synchronises the stream).  idx may be NULL to only count. */
int64_t kafka_compact_scratch_elems(int64_t n_raster);
int kafka_compact_mask(const uint8_t* mask, int64_t n_raster, int32_t* idx, int64_t idx_capacity,
                       int32_t* scratch, int64_t* n_true_host, void* stream);

/* Gather one band's rasters into compacted planar vectors, replacing
 * sort_band_data's R/y/mask extraction (kafka/inference/solvers.py:90-93):
 *   m = obs_mask[idx]; y = m ? obs[idx] : 0; w = m ? inv_var[idx] : 0
 * (w is gated instead of multiplied: masked pixels carry inv_var = inf,
 * kafka/input_output/Sentinel2_Observations.py:175-178).  m_out may be NULL. */
int kafka_gather_band(const double* obs, const double* inv_var, const uint8_t* obs_mask,
                      const int32_t* idx, int64_t n_pix, double* y, double* w, uint8_t* m_out,
                      void* stream);

/* The same gather for all bands of one date in one launch: band b's rasters start
 * at obs + b*raster_stride (likewise inv_var, obs_mask), its compacted rows at
 * y + b*ld (likewise w, m_out).  Same values as n_bands calls of kafka_gather_band. */
int kafka_gather_bands(const double* obs, const double* inv_var, const uint8_t* obs_mask,
                       int64_t raster_stride, int n_bands, const int32_t* idx, int64_t n_pix,
                       int64_t ld, double* y, double* w, uint8_t* m_out, void* stream);

/* Observation ingest fused with the gather, replacing the arithmetic of
 * Sentinel2Observations.get_band_data
 * (kafka/input_output/Sentinel2_Observations.py:166-179): from the RAW surface
 * reflectance raster (scaled integers) to compacted rows
 *   m = raw[idx] > 0 (and ext_mask[idx] != 0 when ext_mask is given)
 *   y = m ? raw[idx] / divisor      : 0        (divisor = 10000)
 *   w = m ? 1 / (rel_sigma * y)^2   : 0        (rel_sigma = 0.05)
 * Integer rasters are converted to fp64 first, as numpy does; a KAFKA_RAW_F32 raster
 * stays float32 through the reader's expressions, so there every step is rounded in
 * float32 and the results widened at the end (bit-identical to the reader in both
 * cases).  Only the raw raster (2 or 4 bytes per pixel) has to cross PCIe instead of
 * the 17 bytes of (float64 reflectance, float64 inverse variance, mask). */
int kafka_gather_scaled_band(const void* raw, int raw_type, const uint8_t* ext_mask, const int32_t* idx,
                             int64_t n_pix, double divisor, double rel_sigma, double* y, double* w,
                             uint8_t* m_out, void* stream);

/* Output stage, replaces KafkaOutput.dump_data's scatter
 * (kafka/input_output/observations.py:374-375,392-393): for parameter p writes
 * mean_raster[p][idx[i]] = (float)x[p][i] and, if diag is given,
 * unc_raster[p][idx[i]] = (float)(1/sqrt(diag[p][i])).  Rasters (n_params x
 * n_raster float32) must be zero-filled by the caller. */
int kafka_scatter_output(const double* x, const double* a_packed, int64_t ld, int64_t n_pix,
                         int n_params, const int32_t* idx, int64_t n_raster, float* mean_raster,
                         float* unc_raster, void* stream);

/* The whole output stage of one date in one call (KafkaOutput.dump_data,
 * kafka/input_output/observations.py:360-399, up to the point where it hands the rasters to GDAL):
 * optionally zero the device rasters, scatter as kafka_scatter_output, then read the rasters back
 * into host_dst (page-locked: the copy is then asynchronous) -- all queued on `stream`.  The two
 * rasters must be one allocation: unc_raster == mean_raster + n_params * n_raster (or NULL), and
 * host_dst receives them in that order.  zero_first may be 0 when the buffer was zero-filled once
 * and is only ever written through the same idx. */
int kafka_output_rasters(const double* x, const double* a_packed, int64_t ld, int64_t n_pix,
                         int n_params, const int32_t* idx, int64_t n_raster, float* mean_raster,
                         float* unc_raster, int zero_first, void* host_dst, void* stream);

/* Observation ingest of the MODIS broadband-albedo reader fused with the gather,
 * replacing the arithmetic of BHRObservations.get_band_data
 * (kafka/input_output/observations.py:288-310).  kernels: the band's three BRDF
 * kernel weights (isotropic, volumetric, geometric), plane k at
 * kernels + k*plane_stride elements, KAFKA_RAW_F32 or KAFKA_RAW_F64; mask, qa:
 * uint8 rasters (qa = MCD43A2 quality level).
 *   bhr   = (k0*1.0 + k1*0.189184) + k2*(-1.377622)    (rounded step by step, as numpy)
 *   sigma = max(2.5e-3, 0.05 bhr) if qa == 0, max(2.5e-3, 0.07 bhr) if qa == 1, else 0
 *   m = mask[idx];  y = m ? bhr : 0;  w = m ? 1/sigma^2 : 0
 * A pixel with its mask set and qa outside {0,1} gets w = inf, exactly as the
 * reader's 1/R^2 does. */
int kafka_gather_bhr_band(const void* kernels, int kernel_type, int64_t plane_stride,
                          const uint8_t* mask, const uint8_t* qa, const int32_t* idx,
                          int64_t n_pix, double* y, double* w, uint8_t* m_out, void* stream);

/* Observation ingest of the Sentinel-1 reader fused with the gather, replacing the
 * arithmetic of S1Observations.get_band_data
 * (kafka/input_output/Sentinel1_Observations.py:107-130,147-149,176-187).  raw: the
 * sigma0 raster of one polarisation, KAFKA_RAW_F32 or KAFKA_RAW_F64.
 *   m = raw[idx] != missing_value                 (missing_value = -999)
 *   y = m ? raw[idx] : 0
 *   w = m ? 1 / (rel_sigma * raw[idx])^2 : 0      (rel_sigma = 0.05)
 * The product, the square and the reciprocal are rounded in the raster's own type,
 * as the reader's numpy expressions are, and then widened: bit-identical to the
 * reader.  An unmasked zero backscatter gets w = inf, exactly as there. */
int kafka_gather_backscatter_band(const void* raw, int raw_type, const int32_t* idx, int64_t n_pix,
                                  double missing_value, double rel_sigma, double* y, double* w,
                                  uint8_t* m_out, void* stream);

/* Number of pixels whose state differs (bitwise) from pixel 0's: 0 means every
 * pixel holds the same vector, which is what run_emulator's de-duplication
 * detects with a Python set (kafka/inference/utils.py:72-74) -- e.g. the first
 * iteration after a prior reset.  *n_diff is a device int32. */
int kafka_state_uniform(const double* x, int64_t ld, int64_t n_pix, int n_params, int32_t* n_diff,
                        void* stream);

/* Scalar read-back without the copy engine (no counterpart in the reference, which
 * has no device): n_bytes (multiple of 8, <= 64 KiB) from device memory are written
 * by a kernel into PINNED host memory (cudaHostAlloc / cudaHostRegister'ed; pageable
 * memory is refused).  The convergence norm of every iteration (linear_kf.py:293)
 * comes back this way: queued as a memcpy it would wait behind the previous date's
 * output rasters on the device->host engine.  The host may read dst once the work
 * queued on `stream` up to this call has completed (record an event and wait). */
int kafka_read_small(const void* src_dev, void* dst_pinned_host, int64_t n_bytes, void* stream);

/* ---- layout conversion at the boundary -----------------------------------
 * interleaved x[n_params*i+p] (kafka/inference/utils.py:201) <-> planar. */
int kafka_interleaved_to_planar(const double* x_il, int64_t n_pix, int n_params, double* x_pl,
                                int64_t ld, void* stream);
int kafka_planar_to_interleaved(const double* x_pl, int64_t ld, int64_t n_pix, int n_params,
                                double* x_il, void* stream);
/* dense blocks[i][r][q] (what block_diag consumes, kafka/inference/utils.py:329-337)
 * <-> packed planar; unpacking mirrors the lower triangle. */
int kafka_blocks_to_packed(const double* blocks, int64_t n_pix, int n_params, double* a_packed,
                           int64_t ld, void* stream);
int kafka_packed_to_blocks(const double* a_packed, int64_t ld, int64_t n_pix, int n_params,
                           double* blocks, void* stream);
/* main diagonal of the block-diagonal matrix, interleaved like
 * `P_analysis_inverse.diagonal()` (kafka/input_output/observations.py:393). */
int kafka_packed_diagonal(const double* a_packed, int64_t ld, int64_t n_pix, int n_params,
                          double* diag_il, void* stream);

/* ---- a7: Gaussian-process emulator ----------------------------------------
 * Host-side packing of one emulator (fields of gp_emulator.GaussianProcess:
 * inputs[T][D], theta[D+2] = LOG hyper-parameters, invQt[T]) into the table the
 * kernels stream through shared memory.  Table = kafka_gp_table_doubles(T,D)
 * doubles: theta_e[D], centre[D], s_f, nugget, wide, L cbar, then T rows
 * {L zh[D], L e, alpha, L e2, sign} with L = 1024/ln2 (exponents in units of
 * ln2/1024), D padded to even; e2 = e + ln|alpha| - cbar with cbar the mean of
 * ln|alpha| (the linearise kernel forms alpha k in one exponential), sign = 0 or
 * 0x80000000 as an integer in the slot's low word; wide != 0 marks a training
 * set so spread out (a point more than ~1000 length scales from the centre, or
 * weights more than e^100 apart) that the kernels keep their fully range-checked
 * exponential.  Weights must be finite (KAFKA_E_BADARG otherwise). */
int64_t kafka_gp_table_doubles(int n_train, int n_dim);
int kafka_gp_pack_host(const double* inputs_host, const double* theta_host,
                       const double* invqt_host, int n_train, int n_dim, double* table_host);

typedef struct {
    const double* table;             /* device, from kafka_gp_pack_host         */
    int32_t n_train;                 /* T                                        */
    int32_t n_dim;                   /* D                                        */
    int32_t mapper[KAFKA_MAX_DIM];   /* state slot feeding emulator input d
                                        (kafka/inference/utils.py:148-153)       */
} kafka_gp_band;

/* Replaces run_emulator + gp.predict(x, do_unc=False)
 * (kafka/inference/utils.py:68-106,206): for every pixel i and band b evaluates
 * mean h0[b][i] and Jacobian jac[b][d][i] at x[mapper_b[d]][i].  All bands of one
 * call share n_dim.  obs_mask (u8 [b*ld+i], may be NULL) zeroes rows of
 * unobserved pixels (utils.py:199-215).  flags: KAFKA_ROUND_F32. */
int kafka_gp_linearize(const double* x, int64_t ld, int64_t n_pix, const kafka_gp_band* bands_host,
                       int n_bands, const uint8_t* obs_mask, double* h0, double* jac,
                       uint32_t flags, void* stream);

/* gp.predict on free-standing inputs xin[d*ld + i]: mean, optional Jacobian
 * jac[d*ld+i], optional predictive variance (needs inv_q, the T x T inverse of
 * the training covariance, row-major, device) -- the do_unc=True branch that
 * the reference never takes on this path. */
int kafka_gp_predict(const double* xin, int64_t ld, int64_t n_pts, const double* table,
                     int n_train, int n_dim, const double* inv_q, double* mean, double* jac,
                     double* var, void* stream);

/* Second-order ("Hessian") correction of the posterior precision, replacing
 * hessian_correction / hessian_correction_multiband
 * (kafka/inference/kf_tools.py:26-72):
 *     corr_i = sum_b  m_bi * w_bi * innovation_bi * E_b^T (d2 h_b / dx2)(x_i) E_b
 * with E_b the embedding of band b's emulator inputs into the state (its
 * mapper; the reference hard-codes the TIP selector, kf_tools.py:19-23).
 * Writes corr (packed, may be NULL) and/or subtracts it from a_inout (packed,
 * may be NULL): P_analysis_inverse - P_correction (kafka/linear_kf.py:416).
 * w must be gated to 0 where unobserved. */
int kafka_gp_hessian_correction(const double* x, int64_t ld, int64_t n_pix, int n_params,
                                const kafka_gp_band* bands_host, int n_bands, const double* innovations,
                                const double* w, double* corr, double* a_inout, void* stream);

/* gp.hessian (kafka/inference/kf_tools.py:28): hess[(d*D+e)*ld + i]. */
int kafka_gp_hessian(const double* xin, int64_t ld, int64_t n_pts, const double* table,
                     int n_train, int n_dim, double* hess, void* stream);

/* ---- a8/a9/a10: fused normal-equation build + Cholesky solve --------------- */
typedef struct {
    int32_t kind;                      /* KAFKA_BAND_ARRAYS / _LINEAR / _UNIFORM             */
    int32_t n_dim;                     /* ARRAYS, UNIFORM: Jacobian planes of this band      */
    int32_t mapper[KAFKA_MAX_DIM];     /* ARRAYS, UNIFORM: state slot of Jacobian plane d    */
    double row[KAFKA_MAX_PARAMS];      /* LINEAR: operator row; WCM: A,B,C,D,E,cos(theta)    */
    double offset;                     /* LINEAR: H0 = row.x_prev + offset                   */
    const double* h0;                  /* ARRAYS: [ld] device; UNIFORM: 1 value              */
    const double* jac;                 /* ARRAYS: [n_dim][ld] device; UNIFORM: [n_dim]       */
    const double* y;                   /* [ld] observations, 0 where unobserved              */
    const double* w;                   /* [ld] inverse variance, 0 where unobserved          */
} kafka_solve_band;

int64_t kafka_solve_scratch_doubles(int64_t n_pix);

/* One linearised update for every pixel, replacing sort_band_data +
 * variational_kalman_multiband (kafka/inference/solvers.py:81-145) and the
 * convergence norm of do_all_bands (kafka/linear_kf.py:293):
 *     y'_b = y_b + J_b.x_prev - H0_b
 *     A    = sum_b w_b J_b J_b^T + Pinv_f          (packed, written to a_out)
 *     rhs  = sum_b w_b y'_b J_b + Pinv_f x_f
 *     x    = A^-1 rhs   (fp64 Cholesky per pixel; a block that is not positive
 *            definite is solved by Gaussian elimination with partial pivoting,
 *            as the reference's sparse LU solves it, solvers.py:133-134)
 *                                                  (written to x_out)
 *     *norm2 = sum_i sum_p (x - x_prev)^2           (deterministic two-stage sum)
 * Optional outputs (may be NULL): a_out (an iteration that cannot be the last
 * one needs no A), innovations[b*ld+i] = y_b - H0_b
 * (solvers.py:142), fwd[b*ld+i] = J_b.(x - x_f) + H0_b (solvers.py:137),
 * n_bad = number of pixels left at x_f: their block is singular or holds a NaN,
 * or a Water-Cloud band was asked for a state outside the model's domain.
 * `scratch` holds kafka_solve_scratch_doubles(n_pix) doubles.
 * flags: KAFKA_PRIOR_BROADCAST; KAFKA_ROUND_F32 (applies to WCM bands, whose
 * operator the reference also stores in float32); KAFKA_SOLVE_COUNT_LU. */
int kafka_solve_bands(int64_t n_pix, int64_t ld, int n_params, const kafka_solve_band* bands_host,
                      int n_bands, const double* x_prev, const double* x_f, const double* pinv_f,
                      double* x_out, double* a_out, double* innovations, double* fwd,
                      double* norm2, int32_t* n_bad, double* scratch, uint32_t flags,
                      void* stream);

/* One whole iteration of LinearKalman.do_all_bands (kafka/linear_kf.py:258-296)
 * when every band is a GP emulator: relinearise all bands at x_prev
 * (create_prosail/nonlinear_observation_operator, kafka/inference/utils.py:130-219)
 * into the caller's workspaces h0_work[n_bands*ld], jac_work[n_bands*D*ld], then
 * run the update of kafka_solve_bands on them.  y, w: [b*ld+i] planes (w gated to
 * 0 where unobserved); obs_mask as in kafka_gp_linearize (may be NULL).  Same
 * results as kafka_gp_linearize followed by kafka_solve_bands on one stream.
 * The caller loops: swap x_prev/x_out, read *norm2, stop as linear_kf.py:288-305. */
int kafka_assimilate_iter(int64_t n_pix, int64_t ld, int n_params, const kafka_gp_band* gp_bands_host,
                          int n_bands, const double* y, const double* w, const uint8_t* obs_mask,
                          const double* x_prev, const double* x_f, const double* pinv_f,
                          double* h0_work, double* jac_work, double* x_out, double* a_out,
                          double* innovations, double* fwd, double* norm2, int32_t* n_bad,
                          double* scratch, uint32_t flags, void* stream);

/* The whole of LinearKalman.do_all_bands (kafka/linear_kf.py:245-323) for one date when every
 * band is a GP emulator: kafka_assimilate_iter in the reference's loop with the reference's
 * stop rule (norm = ||x - x_prev||_2 / len(x) < tol and n_iter >= min_iterations, or the 26th
 * solve; linear_kf.py:293-304), for images so small that per-call overhead is all there is.
 * x_work: [2][n_params*ld] doubles, the iterates ping-pong between the halves and *x_index
 * says which holds the analysis; scalars: 2 doubles on the device; pinned_host: 26 x 16 bytes of
 * page-locked host memory (every solve's sum of squares and bad-pixel count land there);
 * norms_host: >= 26 doubles, one norm per solve; *n_solves = solves taken; *n_bad_last /
 * *n_bad_max = pixels left at the forecast in the last / the worst iteration.  a_out,
 * innovations and fwd (may be NULL) are written by iterations that can be the last one.
 * x_uniform (may be NULL): n_params contiguous doubles on the device -- the state shared by ALL
 * pixels of x_f (e.g. after a prior reset); the first iteration then evaluates the emulators once
 * for that vector, the reference's single-unique-vector branch (kafka/inference/utils.py:72-74).
 * Synchronises `stream` once per iteration that can end the loop.
 * while_queued (may be NULL): called once, with while_queued_arg, after the first iterations have been
 * queued and before the first wait -- host work the caller wants done while the GPU is busy (the
 * reference's loop has no such slot; here it is where the next date's rasters are read and sent,
 * kafka/linear_kf.py:225-227).  It must not touch `stream`'s pending work or the buffers of this call. */
typedef void (*kafka_host_hook)(void* arg);
int kafka_assimilate_date(int64_t n_pix, int64_t ld, int n_params, const kafka_gp_band* gp_bands_host,
                          int n_bands, const double* y, const double* w, const uint8_t* obs_mask,
                          const double* x_f, const double* pinv_f, double* h0_work, double* jac_work,
                          double* x_work, double* a_out, double* innovations, double* fwd,
                          double* scalars, void* pinned_host, double* scratch, double tol,
                          int min_iterations, uint32_t flags, void* stream, double* norms_host,
                          int* n_solves, int* n_bad_last, int* n_bad_max, int* x_index,
                          const double* x_uniform, kafka_host_hook while_queued, void* while_queued_arg);

/* n host->device copies queued on `stream` by one call: the upload of one date's rasters
 * (observations.get_band_data, kafka/linear_kf.py:225-227, hands over three rasters per band) for
 * images so small that the cost per copy call rivals the transfer.  src_host[k] should be
 * page-locked (then the copies are asynchronous). */
int kafka_upload_rasters(int n, const void* const* src_host, void* const* dst_dev, const int64_t* n_bytes,
                         void* stream);

/* Same update as kafka_solve_bands for callers that cannot allocate device memory
 * and therefore cannot pass `scratch`.  It has no counterpart in the reference
 * (pure Python, no device memory at all).  Typical callers: a plug-in inside a
 * host application that reserved the whole GPU for its own memory pool; a
 * latency-bound loop that must not hit the device-wide synchronisation of
 * cudaMalloc/cudaFree; a thin FFI stub that owns nothing but the input and
 * output arrays it was handed.  The scratch is a fixed array inside the library
 * (KAFKA_STATIC_SCRATCH_DOUBLES doubles = 4 MiB of device memory on every GPU
 * that loads the library, until the process exits), so:
 *   - at most KAFKA_STATIC_MAX_PIXELS pixels per call, else KAFKA_E_TOOLARGE;
 *   - the call neither allocates nor frees anything (no cudaMalloc/cudaFree, no
 *     allocating library routine);
 *   - all calls on a device share the array and it keeps the previous call's
 *     contents; the slice a call uses is zeroed first on the call's stream, and
 *     no result depends on older contents;
 *   - calls that overlap on the GPU (two streams, two host threads) race on the
 *     array: serialise them or use kafka_solve_bands with private scratch;
 *   - the first call on each device looks up the array's device address (one-off
 *     set-up); issue it before any stream capture starts.
 * Arguments and results are exactly those of kafka_solve_bands minus `scratch`. */
#define KAFKA_STATIC_SCRATCH_DOUBLES 524288
#define KAFKA_STATIC_MAX_PIXELS 16758144 /* (524288 - 596) / 4 tiles of 128 pixels */
int kafka_solve_bands_static(int64_t n_pix, int64_t ld, int n_params, const kafka_solve_band* bands_host,
                             int n_bands, const double* x_prev, const double* x_f, const double* pinv_f,
                             double* x_out, double* a_out, double* innovations, double* fwd,
                             double* norm2, int32_t* n_bad, uint32_t flags, void* stream);

/* Water-Cloud SAR forward model and gradient, replacing sar_observation_operator
 * (kafka/observation_operators/sar_forward_model.py:13-113) for inspection and
 * tests: x[0*ld+i] = vegetation descriptor (LAI), x[1*ld+i] = soil moisture;
 * params_host = {A, B, C, D, E, cos(incidence angle)}; sigma0[i] (linear units),
 * grad[0*ld+i], grad[1*ld+i] (may be NULL).  Inside the filter the same maths runs
 * fused in kafka_solve_bands (band kind KAFKA_BAND_WCM). */
int kafka_wcm_forward(const double* x, int64_t ld, int64_t n, const double* params_host, double* sigma0,
                      double* grad, void* stream);

/* ---- a12-a17: advance / propagation ----------------------------------------
 * propagate_information_filter_LAI (kafka/inference/kf_tools.py:292-314),
 * generalised from its hard-coded 7-parameter TIP prior: every slot is reset to
 * prior_mean_host except `carry_slot`, whose value is carried over and whose
 * precision becomes 1/(1/a[slot,slot] + q[slot]); the block is
 * prior_inv_host (n_params^2, row-major) elsewhere.  Block values are rounded to
 * float32 like the reference's block_diag(..., dtype=float32) (:312). */
int kafka_advance_carry_slot(const double* x_a, const double* a_packed, const double* q,
                             int64_t ld, int64_t n_pix, int n_params, int carry_slot,
                             const double* prior_mean_host, const double* prior_inv_host,
                             double* x_f, double* pinv_f, void* stream);
/* propagate_information_filter_approx_SLOW (kf_tools.py:247-289): diagonal
 * m/(1+m q), off-diagonals dropped. */
int kafka_advance_approx(const double* a_packed, const double* q, int64_t ld, int64_t n_pix,
                         int n_params, double* pinv_f, void* stream);
/* propagate_information_filter_SLOW (kf_tools.py:208-245): (I + Pinv Q)^-1 Pinv
 * per block. */
int kafka_advance_exact(const double* a_packed, const double* q, int64_t ld, int64_t n_pix,
                        int n_params, double* pinv_f, int32_t* n_bad, void* stream);
/* propagate_standard_kalman (kf_tools.py:174-205) for diagonal M:
 * x_f = m*x_a, P_f = P_a + diag(q).  m may be NULL (identity). */
int kafka_advance_standard(const double* x_a, const double* p_packed, const double* m,
                           const double* q, int64_t ld, int64_t n_pix, int n_params, double* x_f,
                           double* p_f, void* stream);
/* blend_prior (kf_tools.py:75-96) including its cross-weighted right-hand side
 * b = Pinv_f mu_prior + Cinv_prior x_f (:90).  prior_* may be broadcast (one
 * vector / one packed block) when KAFKA_PRIOR_BROADCAST is set. */
int kafka_blend_prior(const double* prior_mean, const double* prior_inv, const double* x_f,
                      const double* pinv_f, int64_t ld, int64_t n_pix, int n_params, double* x_out,
                      double* c_out, int32_t* n_bad, uint32_t flags, void* stream);

/* ---- e: multi-GPU (one process per GPU, the image sharded by pixel) --------
 * Replaces the reference's fan-out of chunks over dask workers
 * (kafka_test_Py36.py:241-255: client.map(wrapper, chunks) + gather) for a binder
 * that has no torch.distributed: an opaque per-rank context holding an NCCL
 * communicator over NVLink.  Pixels never talk to each other; the one coupling
 * is the scalar convergence norm of kafka/linear_kf.py:293, which the reference
 * computes per chunk and which here is made global by one small all-reduce per
 * iteration.  libnccl.so.2 is opened when the first context is created (the copy
 * already loaded in the process if there is one); status codes >= 1000 are
 * 1000 + ncclResult_t.
 *
 * kafka_comm_unique_id   rank 0 fills id128_host (KAFKA_COMM_ID_BYTES bytes) and hands it to
 *                        the other ranks by whatever means the host application has
 * kafka_ctx_create       collective over all ranks; the communicator is bound to the CUDA
 *                        device that is current in the calling thread
 * kafka_allreduce_norm   in-place SUM over ranks of `count` doubles on the device, queued on
 *                        `stream`: [sum of squares of the step, element count, bad pixels]
 *                        as kafka_solve_bands leaves them -> every rank forms
 *                        sqrt(sums[0]) / sums[1] and takes the same stop decision
 * kafka_gather_outputs   concatenation, in rank order, of every rank's `counts_host[rank]`
 *                        doubles (state planes, diag(P^-1), ...) into recv on every rank:
 *                        the final output gather; shards may differ in size */
#define KAFKA_COMM_ID_BYTES 128
typedef struct kafka_ctx kafka_ctx;
int kafka_comm_unique_id(void* id128_host);
int kafka_ctx_create(int rank, int n_ranks, const void* id128_host, kafka_ctx** ctx);
int kafka_ctx_destroy(kafka_ctx* ctx);
int kafka_ctx_info(const kafka_ctx* ctx, int* rank, int* n_ranks, int* nccl_version);
int kafka_allreduce_norm(kafka_ctx* ctx, double* sums, int count, void* stream);
int kafka_gather_outputs(kafka_ctx* ctx, const double* send, const int64_t* counts_host, double* recv,
                         void* stream);

#ifdef __cplusplus
}
#endif
#endif /* KAFKA_B200_H */
