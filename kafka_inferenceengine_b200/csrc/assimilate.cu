// kafka_assimilate_iter: ONE iteration of LinearKalman.do_all_bands
// (kafka/linear_kf.py:258-296) behind a single call, for bindings that would
// rather not build the kafka_solve_band descriptors themselves: relinearise every
// emulator band at x_prev (kafka_gp_linearize), then build and solve the normal
// equations and accumulate the convergence norm (kafka_solve_bands).  Nothing is
// computed here that the two entry points do not compute; results are identical
// to calling them one after the other on the same stream.
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
