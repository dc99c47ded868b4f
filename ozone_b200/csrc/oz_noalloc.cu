// oz_tm_adjoint_static: the reverse sweep for callers that cannot hand the library a workspace.
//
// Who needs this.  oz_tm_adjoint() reduces the gradients of batch-shared inputs (step sizes,
// shared parameters) through per-warp partial sums, and replays checkpoint segments, in a scratch
// buffer the caller must allocate on the device.  Some hosts of this library have no way to do
// that: a plug-in running inside a solver or optimiser that owns the device and forbids
// allocation in its callbacks, a real-time control loop where cudaMalloc's implicit
// synchronisation is not acceptable, or a thin FFI binding (a C or Fortran driver, a ctypes stub)
// that can pass through the input/output pointers it was given but has no allocator of its own.
// For them the scratch lives inside the library image: a fixed array in device memory that every
// GPU which loads liboz_b200.so carries from load until the process exits.
//
// What the caller must know.
//   * Limit: OZ_STATIC_SCRATCH_BYTES (4 MiB).  A problem whose reverse sweep needs more is
//     refused with a message saying how much it needs; use oz_tm_adjoint() with a workspace then.
//   * The first call on each device looks up the array's device address.  Make that first call
//     outside of any stream capture.
//   * All calls on one device share the one array and it keeps whatever the last call left in
//     it.  Nothing in it has to start at zero: every partial sum and every replayed state is
//     written by the sweep before anything reads it, so no clearing pass is enqueued.
//   * Because the array is shared, two calls whose kernels can overlap on the same GPU (different
//     streams, different host threads) race on it.  Serialise them, e.g. issue them on one stream.
//   * No cudaMalloc / cudaFree / allocating library routine is called on this path.
#include <mutex>
#include "oz_common.cuh"

#define OZ_STATIC_SCRATCH_BYTES (4u << 20)

namespace {

// File-scope device array.  In host code this name is only a handle for the runtime's symbol
// table -- NOT a device address; the address comes from cudaGetSymbolAddress below.
static __device__ __align__(256) unsigned char g_static_scratch[OZ_STATIC_SCRATCH_BYTES];

constexpr int kMaxDevices = 64;
std::mutex g_mu;
void* g_scratch_addr[kMaxDevices] = {nullptr};

// Device address of g_static_scratch on the CURRENT device, resolved once per device.
int static_scratch(void** out) {
    int dev = -1;
    OZ_CUDA_CHECK(cudaGetDevice(&dev));
    OZ_REQUIRE(dev >= 0 && dev < kMaxDevices, "device ordinal %d out of range", dev);
    std::lock_guard<std::mutex> lock(g_mu);
    if (!g_scratch_addr[dev]) {
        void* ptr = nullptr;
        OZ_CUDA_CHECK(cudaGetSymbolAddress((void**)&ptr, g_static_scratch));
        OZ_REQUIRE(ptr != nullptr, "cudaGetSymbolAddress returned a null address for the static scratch");
        g_scratch_addr[dev] = ptr;
    }
    *out = g_scratch_addr[dev];
    return 0;
}

}  // namespace

extern "C" {

OZ_API size_t oz_static_scratch_bytes(void) { return (size_t)OZ_STATIC_SCRATCH_BYTES; }

OZ_API int oz_tm_adjoint_static(const oz_problem* pb) {
    OZ_REQUIRE(pb != nullptr, "null problem descriptor");
    const size_t need = oz_tm_adjoint_workspace_bytes(pb);
    if (need == (size_t)-1) return -1;  // message already set by the query
    OZ_REQUIRE(need <= (size_t)OZ_STATIC_SCRATCH_BYTES,
               "oz_tm_adjoint_static: this reverse sweep needs %zu bytes of scratch but the built-in scratch holds "
               "%zu; call oz_tm_adjoint with a caller-provided workspace instead",
               need, (size_t)OZ_STATIC_SCRATCH_BYTES);
    void* scratch = nullptr;
    if (int rc = static_scratch(&scratch)) return rc;
    oz_problem local = *pb;  // the caller's descriptor is not modified
    local.ws = scratch;
    local.ws_bytes = (size_t)OZ_STATIC_SCRATCH_BYTES;
    return oz_tm_adjoint(&local);
}

}  // extern "C"
