// pb200_noalloc.cu -- pb200_dot_noalloc: the weighted inner product for callers that must not
// allocate device memory.
//
// Who needs this: a host program whose GPU memory is already spoken for -- for example a solver
// or framework that reserved (almost) the whole device for its own pool at start-up, so that a
// cudaMalloc from a library would fail or fragment it -- and time-critical loops in which
// cudaMalloc/cudaFree are unacceptable because they synchronise the whole device. Every other
// entry point of this library obtains its block partials, ticket and result slots from
// cudaMalloc; this one keeps them in a fixed array that is part of the loaded module.
//
// Consequences of that choice, all of which the caller has to know:
//   * the array has a fixed size, so the problem size is capped (PB200_NOALLOC_MAX_BLOCKS reduction
//     blocks, see the entry point); larger requests are refused with PB200_ERR_TOO_LARGE;
//   * every GPU that loads libpoisson_b200.so carries the array (64 KiB + 256 B) until the process
//     exits, whether or not this entry point is ever called;
//   * all calls on one GPU share the one array and it keeps whatever the previous call left in it.
//     The entry point re-zeroes the ticket word before each launch, so calls issued one after the
//     other (same stream, or otherwise ordered) are fine, but two calls whose kernels overlap on the
//     GPU (different streams, different host threads) race on it and return garbage;
//   * the first call on each device looks up the array's device address and caches it. Make that
//     first call outside of any CUDA stream capture.
#include "pb200_solver.h"

#include <mutex>

namespace pb200 {
namespace {

constexpr int NOALLOC_MAX_BLOCKS = 8192;                                   // block partials the array can hold
constexpr size_t NOALLOC_TICKET_OFFSET = sizeof(double) * NOALLOC_MAX_BLOCKS; // ticket word sits behind the partials
constexpr size_t NOALLOC_BYTES = NOALLOC_TICKET_OFFSET + 256;
constexpr int NOALLOC_MAX_DEVICES = 64;

// The scratch itself. In host code this name is only a handle for cudaGetSymbolAddress: it is not a
// device address and must never be handed to a kernel as one.
static __device__ __align__(256) unsigned char g_noalloc_scratch[NOALLOC_BYTES];

std::mutex g_noalloc_mu;
void *g_noalloc_ptr[NOALLOC_MAX_DEVICES] = {nullptr}; // device address of g_noalloc_scratch, per device ordinal

// Same arithmetic and summation order as dot_kernel (pb200_kernels.cuh); the partials, the ticket and
// nothing else live in `scratch`. The last block writes (sum * h1) * h2 (S3:191) to *out.
__global__ void __launch_bounds__(256) dot_noalloc_kernel(int nx, int ny, long long pitch, const double *__restrict__ u,
                                                          const double *__restrict__ v, unsigned char *scratch,
                                                          double h1, double h2, double *out) {
    double *partials = reinterpret_cast<double *>(scratch);
    unsigned int *ticket = reinterpret_cast<unsigned int *>(scratch + NOALLOC_TICKET_OFFSET);
    double acc[1] = {0.0};
    const int lj = 1 + blockIdx.x * blockDim.x + threadIdx.x;
    if (lj <= ny)
        for (int li = 1 + blockIdx.y; li <= nx; li += gridDim.y) {
            const size_t o = (size_t)li * (size_t)pitch + lj;
            acc[0] += u[o] * v[o];
        }
    double tot[1];
    if (grid_reduce<1, 256>(acc, partials, ticket, tot)) out[0] = tot[0] * h1 * h2;
}

// Once per device, with that device current: ask the runtime where this device's copy of the array is.
pb200_status noalloc_scratch_for_current_device(unsigned char **ptr) {
    int dev = -1;
    CK(cudaGetDevice(&dev));
    if (dev < 0 || dev >= NOALLOC_MAX_DEVICES) {
        set_error("pb200_dot_noalloc: device ordinal %d outside the supported 0..%d", dev, NOALLOC_MAX_DEVICES - 1);
        return PB200_ERR_INVALID;
    }
    std::lock_guard<std::mutex> lk(g_noalloc_mu);
    if (!g_noalloc_ptr[dev]) {
        void *p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess || !p) {
            set_error("pb200_dot_noalloc: cudaGetSymbolAddress failed on device %d: %s", dev, cudaGetErrorString(e));
            return PB200_ERR_CUDA;
        }
        g_noalloc_ptr[dev] = p;
    }
    *ptr = static_cast<unsigned char *>(g_noalloc_ptr[dev]);
    return PB200_OK;
}

} // namespace
} // namespace pb200

extern "C" pb200_status pb200_dot_noalloc(const double *dev_u, const double *dev_v, int32_t nx, int32_t ny, int64_t pitch,
                                          double h1, double h2, double *dev_out, void *cuda_stream) {
    using namespace pb200;
    if (!dev_u || !dev_v || !dev_out) { set_error("pb200_dot_noalloc: null pointer"); return PB200_ERR_INVALID; }
    if (nx < 1 || ny < 1 || pitch < (int64_t)ny + 2) { set_error("pb200_dot_noalloc: need nx, ny >= 1 and pitch >= ny+2"); return PB200_ERR_INVALID; }
    const dim3 grid((ny + 255) / 256, nx < 512 ? nx : 512);
    const long long blocks = (long long)grid.x * grid.y;
    if (blocks > NOALLOC_MAX_BLOCKS) {
        set_error("pb200_dot_noalloc: %d x %d needs %lld reduction blocks, the built-in scratch holds %d "
                  "(limit: ceil(ny/256) * min(nx,512) <= %d, e.g. ny <= 4096 for nx >= 512); use pb200_op_dot or the solver",
                  nx, ny, blocks, NOALLOC_MAX_BLOCKS, NOALLOC_MAX_BLOCKS);
        return PB200_ERR_TOO_LARGE;
    }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        set_error("no CUDA device: libpoisson_b200 has no CPU path");
        return PB200_ERR_CUDA;
    }
    unsigned char *scratch = nullptr;
    pb200_status st = noalloc_scratch_for_current_device(&scratch);
    if (st != PB200_OK) return st;
    cudaStream_t stream = static_cast<cudaStream_t>(cuda_stream);
    // the array still holds the previous call's contents: the ticket must start at zero, the partials are
    // all overwritten before they are read
    CK(cudaMemsetAsync(scratch + NOALLOC_TICKET_OFFSET, 0, 256, stream));
    dot_noalloc_kernel<<<grid, 256, 0, stream>>>(nx, ny, pitch, dev_u, dev_v, scratch, h1, h2, dev_out);
    CK(cudaGetLastError());
    return PB200_OK;
}
