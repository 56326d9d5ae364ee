// kafka_solve_bands_static: the per-pixel update of kafka_solve_bands for callers
// that cannot hand over a scratch buffer because they cannot allocate one.
//
// Who that is.  kafka_solve_bands needs a few kilobytes to a couple of megabytes
// of device scratch (partial sums, counts and one failure bit per pixel)
// and leaves it to the caller to provide it.  Some callers have nowhere to get
// it from:
//   * a plug-in running inside a host application whose own memory manager
//     reserved the whole GPU at start-up, so a cudaMalloc issued later by the
//     plug-in fails or eats into memory the host believes it owns;
//   * a per-date processing loop with a latency budget: cudaMalloc and cudaFree
//     synchronise the device, which stalls every other stream of the process;
//   * a thin foreign-function binding (the ctypes stub a maintainer adds to the
//     reference, a C or Fortran driver) that only ever sees the input and output
//     arrays it was given and has no allocator to size, own and free one more
//     buffer whose length depends on the launch geometry of this library.
// For them the scratch lives in this library instead: a fixed array in device
// global memory, part of the loaded module on every GPU that loads the library
// and held until the process exits.
//
// Rules that follow from sharing one array:
//   * size limit: KAFKA_STATIC_SCRATCH_DOUBLES doubles (4 MiB), enough for
//     KAFKA_STATIC_MAX_PIXELS pixels per call; larger calls are refused with
//     KAFKA_E_TOOLARGE and must use kafka_solve_bands with their own scratch;
//   * the array keeps whatever the previous call left in it.  Nothing in it has
//     to start at zero for a correct result (every slot a call reads is written
//     by that same call first); the slice a call is about to use is zeroed on the
//     call's stream anyway so that stale values can never leak into a result;
//   * two calls that overlap on the GPU (different streams, or different host
//     threads on the same device) write the same slots and race.  Serialise them,
//     or give all but one of them private scratch through kafka_solve_bands;
//   * the first call on each device looks up the device address of the array (and
//     may make the driver load this module).  Make that first call during
//     start-up, before any stream capture begins.
#include <mutex>

#include "common.cuh"

namespace {

// File-scope device array.  In host code this name is only a handle for the
// runtime: its device address must be asked for with cudaGetSymbolAddress.
static __device__ __align__(256) double g_static_scratch[KAFKA_STATIC_SCRATCH_DOUBLES];

// Device code may name the array directly: this kernel zeroes the slice the next
// solve is going to use.
__global__ void clear_static_scratch_kernel(int64_t n) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        g_static_scratch[i] = 0.0;
}

constexpr int kMaxDevices = 64;
double* g_scratch_addr[kMaxDevices] = {nullptr};
std::mutex g_setup_mutex;

// Device address of g_static_scratch on the CURRENT device, looked up once per
// device and remembered.
int static_scratch_for_current_device(double** out) {
    int dev = -1;
    KAFKA_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDevices) {
        kafka::set_error("kafka_solve_bands_static: device ordinal %d outside 0..%d", dev, kMaxDevices - 1);
        return KAFKA_E_UNSUPPORTED;
    }
    std::lock_guard<std::mutex> lock(g_setup_mutex);
    if (g_scratch_addr[dev] == nullptr) {
        double* ptr = nullptr;
        cudaError_t e = cudaGetSymbolAddress((void**)&ptr, g_static_scratch);
        if (e != cudaSuccess || ptr == nullptr) {
            kafka::set_error("kafka_solve_bands_static: cudaGetSymbolAddress failed on device %d: %s", dev,
                             cudaGetErrorString(e));
            return e != cudaSuccess ? (int)e : KAFKA_E_UNSUPPORTED;
        }
        g_scratch_addr[dev] = ptr;
    }
    *out = g_scratch_addr[dev];
    return 0;
}

}  // namespace

extern "C" {

int kafka_solve_bands_static(int64_t n_pix, int64_t ld, int n_params, const kafka_solve_band* bands_host, int n_bands,
                             const double* x_prev, const double* x_f, const double* pinv_f, double* x_out,
                             double* a_out, double* innovations, double* fwd, double* norm2, int32_t* n_bad,
                             uint32_t flags, void* stream) {
    if (n_pix < 0) {
        kafka::set_error("kafka_solve_bands_static: n_pix=%lld is negative", (long long)n_pix);
        return KAFKA_E_BADARG;
    }
    const int64_t need = kafka_solve_scratch_doubles(n_pix);
    if (n_pix > KAFKA_STATIC_MAX_PIXELS || need > KAFKA_STATIC_SCRATCH_DOUBLES) {
        kafka::set_error(
            "kafka_solve_bands_static: %lld pixels need %lld doubles of scratch but the built-in array holds %d "
            "(at most %lld pixels per call); use kafka_solve_bands with caller-provided scratch",
            (long long)n_pix, (long long)need, (int)KAFKA_STATIC_SCRATCH_DOUBLES, (long long)KAFKA_STATIC_MAX_PIXELS);
        return KAFKA_E_TOOLARGE;
    }
    double* scratch = nullptr;
    if (int rc = static_scratch_for_current_device(&scratch)) return rc;
    if (n_pix > 0) {
        const int threads = 256;
        const int blocks = (int)((need + threads - 1) / threads);
        clear_static_scratch_kernel<<<blocks < 1024 ? blocks : 1024, threads, 0, (cudaStream_t)stream>>>(need);
        if (int rc = kafka::launch_status("clear_static_scratch_kernel")) return rc;
    }
    return kafka_solve_bands(n_pix, ld, n_params, bands_host, n_bands, x_prev, x_f, pinv_f, x_out, a_out, innovations,
                             fwd, norm2, n_bad, scratch, flags, stream);
}

}  // extern "C"
