// kafka_assimilate_iter: ONE iteration of LinearKalman.do_all_bands
// (kafka/linear_kf.py:258-296) behind a single call, for bindings that would
// rather not build the kafka_solve_band descriptors themselves: relinearise every
// emulator band at x_prev (kafka_gp_linearize), then build and solve the normal
// equations and accumulate the convergence norm (kafka_solve_bands).  Nothing is
// computed here that the two entry points do not compute; results are identical
// to calling them one after the other on the same stream.
#include <math.h>

#include "common.cuh"

extern "C" int kafka_assimilate_iter(int64_t n_pix, int64_t ld, int n_params, const kafka_gp_band* gp_bands_host,
                                     int n_bands, const double* y, const double* w, const uint8_t* obs_mask,
                                     const double* x_prev, const double* x_f, const double* pinv_f, double* h0_work,
                                     double* jac_work, double* x_out, double* a_out, double* innovations,
                                     double* fwd, double* norm2, int32_t* n_bad, double* scratch, uint32_t flags,
                                     void* stream) {
    KAFKA_CHECK_ARG(n_bands >= 1 && n_bands <= KAFKA_MAX_BANDS, "kafka_assimilate_iter: n_bands %d outside 1..%d",
                    n_bands, KAFKA_MAX_BANDS);
    KAFKA_CHECK_ARG(gp_bands_host && y && w && h0_work && jac_work,
                    "kafka_assimilate_iter: null band, observation or workspace pointer");
    const int D = gp_bands_host[0].n_dim;
    int rc = kafka_gp_linearize(x_prev, ld, n_pix, gp_bands_host, n_bands, obs_mask, h0_work, jac_work,
                                flags & KAFKA_ROUND_F32, stream);
    if (rc != 0) return rc;
    kafka_solve_band sb[KAFKA_MAX_BANDS];
    for (int b = 0; b < n_bands; ++b) {
        kafka_solve_band& s = sb[b];
        s = kafka_solve_band{};
        s.kind = KAFKA_BAND_ARRAYS;
        s.n_dim = D;
        for (int d = 0; d < D; ++d) s.mapper[d] = gp_bands_host[b].mapper[d];
        s.h0 = h0_work + (int64_t)b * ld;
        s.jac = jac_work + (int64_t)b * D * ld;
        s.y = y + (int64_t)b * ld;
        s.w = w + (int64_t)b * ld;
    }
    return kafka_solve_bands(n_pix, ld, n_params, sb, n_bands, x_prev, x_f, pinv_f, x_out, a_out, innovations, fwd,
                             norm2, n_bad, scratch, flags, stream);
}

// kafka_assimilate_date: the WHOLE iterated linearisation of one observation date
// (LinearKalman.do_all_bands, kafka/linear_kf.py:245-323) behind one call, for images so small
// that the kernels take microseconds and the caller's per-launch overhead is all there is (the
// 128 x 128 .. 256 x 256 chunks the reference's drivers hand to LinearKalman,
// kafka_test_S2.py:202, kafka_test_Py36.py:241).  The loop and the stop rule are the
// reference's (linear_kf.py:253-307): relinearise every band at x_prev, solve,
// norm = ||x - x_prev||_2 / len(x), stop when norm < tol and n_iter >= min_iterations or after
// the 26th solve; the norms come back through page-locked memory written by a kernel
// (kafka_read_small, 16 bytes per solve) and the stream is synchronised once per iteration that
// can end the loop (iterations below min_iterations are queued back to back; before the first wait the
// caller's while_queued hook runs, so its host work hides behind the queued kernels).  The states ping-pong
// between the two halves of x_work; *x_index says which half holds the analysis.  A and the
// innovations are stored only by iterations that can be the last one.
extern "C" int kafka_assimilate_date(int64_t n_pix, int64_t ld, int n_params, const kafka_gp_band* gp_bands_host,
                                     int n_bands, const double* y, const double* w, const uint8_t* obs_mask,
                                     const double* x_f, const double* pinv_f, double* h0_work, double* jac_work,
                                     double* x_work, double* a_out, double* innovations, double* fwd,
                                     double* scalars, void* pinned_host, double* scratch, double tol,
                                     int min_iterations, uint32_t flags, void* stream, double* norms_host,
                                     int* n_solves, int* n_bad_last, int* n_bad_max, int* x_index,
                                     const double* x_uniform, kafka_host_hook while_queued,
                                     void* while_queued_arg) {
    KAFKA_CHECK_ARG(x_work && scalars && pinned_host && norms_host && n_solves && x_index && a_out,
                    "kafka_assimilate_date: null workspace or result pointer");
    KAFKA_CHECK_ARG(n_pix >= 1 && n_params >= 1, "kafka_assimilate_date: empty problem");
    KAFKA_CHECK_ARG(n_bands >= 1 && n_bands <= KAFKA_MAX_BANDS && gp_bands_host, "kafka_assimilate_date: n_bands %d", n_bands);
    cudaStream_t s = (cudaStream_t)stream;
    const double* x_prev = x_f;
    const double n_elems = (double)n_pix * (double)n_params;
    int bad_max = 0, bad = 0, idx = 0, n_read = 0;
    int n_iter = 1;
    for (;; ++n_iter) {
        idx = (n_iter - 1) & 1;
        double* x_new = x_work + (int64_t)idx * n_params * ld;
        const bool may_be_last = n_iter >= min_iterations || n_iter > 25;
        int rc;
        if (n_iter == 1 && x_uniform) {
            // every pixel starts from the same state (x_uniform: its n_params values, contiguous): the
            // emulators are evaluated ONCE, as the reference's run_emulator does for a single unique vector
            // (kafka/inference/utils.py:72-74), and the solve takes that one linearisation for all pixels
            const int D = gp_bands_host[0].n_dim;
            rc = kafka_gp_linearize(x_uniform, 1, 1, gp_bands_host, n_bands, nullptr, h0_work, jac_work,
                                    flags & KAFKA_ROUND_F32, stream);
            if (rc != 0) return rc;
            kafka_solve_band sb[KAFKA_MAX_BANDS];
            for (int b = 0; b < n_bands; ++b) {
                kafka_solve_band& u = sb[b];
                u = kafka_solve_band{};
                u.kind = KAFKA_BAND_UNIFORM;
                u.n_dim = D;
                for (int d = 0; d < D; ++d) u.mapper[d] = gp_bands_host[b].mapper[d];
                u.h0 = h0_work + b;                                   // ld = 1: band b's value at [b], its row at [b*D ..]
                u.jac = jac_work + (int64_t)b * D;
                u.y = y + (int64_t)b * ld;
                u.w = w + (int64_t)b * ld;
            }
            rc = kafka_solve_bands(n_pix, ld, n_params, sb, n_bands, x_prev, x_f, pinv_f, x_new,
                                   may_be_last ? a_out : nullptr, may_be_last ? innovations : nullptr,
                                   may_be_last ? fwd : nullptr, scalars, reinterpret_cast<int32_t*>(scalars + 1), scratch,
                                   flags, stream);
        } else {
            rc = kafka_assimilate_iter(n_pix, ld, n_params, gp_bands_host, n_bands, y, w, obs_mask, x_prev, x_f, pinv_f,
                                       h0_work, jac_work, x_new, may_be_last ? a_out : nullptr,
                                       may_be_last ? innovations : nullptr, may_be_last ? fwd : nullptr, scalars,
                                       reinterpret_cast<int32_t*>(scalars + 1), scratch, flags, stream);
        }
        if (rc != 0) return rc;
        // slot n_iter-1 of the page-locked buffer: [sum of squares | bad pixels] of this solve
        char* slot = static_cast<char*>(pinned_host) + 16 * (n_iter - 1);
        rc = kafka_read_small(scalars, slot, 16, stream);
        if (rc != 0) return rc;
        x_prev = x_new;
        // an iteration below min_iterations cannot end the loop (linear_kf.py:297): its norm is only reported,
        // so the next iteration is queued right behind it and the host waits once for both
        if (n_iter < min_iterations && n_iter <= 25) continue;
        if (while_queued) {                                         // the caller's host work, behind the queued iterations
            while_queued(while_queued_arg);
            while_queued = nullptr;
        }
        KAFKA_CUDA(cudaStreamSynchronize(s));
        double norm = 0.0;
        for (; n_read < n_iter; ++n_read) {
            const volatile double* host = reinterpret_cast<const volatile double*>(static_cast<char*>(pinned_host) + 16 * n_read);
            bad = *reinterpret_cast<const volatile int32_t*>(host + 1);
            bad_max = bad > bad_max ? bad : bad_max;
            norm = sqrt(host[0]) / n_elems;                         // linear_kf.py:293
            norms_host[n_read] = norm;
        }
        if ((norm < tol && n_iter >= min_iterations) || n_iter > 25) break;   // linear_kf.py:297-304
    }
    *n_solves = n_iter;
    *x_index = idx;
    if (n_bad_last) *n_bad_last = bad;
    if (n_bad_max) *n_bad_max = bad_max;
    return 0;
}

// kafka_upload_rasters: `n` host-to-device copies queued on `stream` by ONE call -- for images so small that
// the caller's cost per copy call (a date of 10 bands is 30 rasters) rivals the transfer itself.  Sources
// should be page-locked (the copies are then asynchronous); pageable sources are copied correctly but the
// call waits for them.
extern "C" int kafka_upload_rasters(int n, const void* const* src_host, void* const* dst_dev, const int64_t* n_bytes,
                                    void* stream) {
    KAFKA_CHECK_ARG(n >= 0 && (n == 0 || (src_host && dst_dev && n_bytes)), "kafka_upload_rasters: null pointer");
    for (int k = 0; k < n; ++k) {
        KAFKA_CHECK_ARG(src_host[k] && dst_dev[k] && n_bytes[k] >= 0, "kafka_upload_rasters: bad entry %d", k);
        KAFKA_CUDA(cudaMemcpyAsync(dst_dev[k], src_host[k], (size_t)n_bytes[k], cudaMemcpyHostToDevice, (cudaStream_t)stream));
    }
    return 0;
}
