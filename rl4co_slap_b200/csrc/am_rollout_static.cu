// rl4co_am_rollout_static - the rollout (K2) with its scratch memory held by the library.
//
// No counterpart in the reference: added for callers that cannot
// allocate device memory around the call (code under CUDA-graph capture, tight serving loops, harnesses
// that forbid cudaMalloc).
//
// The scratch of the rollout path is the per-node table block [n_tables][B*N][128] fp32 that the
// projection (K3, rl4co_am_project_tc) writes and ONLY the rollout kernel (K2, rl4co_am_rollout) reads;
// the regular path has the caller allocate it per chunk (policy.py: torch.empty).  Here it is a file-scope
// static __device__ array of RL4CO_STATIC_TABLE_BYTES (32 MiB) that every GPU loading the library holds
// until the process exits.  The entry point runs K3 into that array and K2 out of it, through the very
// same launchers as the regular entry points, and calls no cudaMalloc / cudaFree / allocating routine.
//
// Host code must not hand the array's NAME to a kernel: on the host `g_static_tables` is a stand-in
// symbol, not a device address.  The device address is fetched once per device with
// cudaGetSymbolAddress (that device current) and cached.
//
// Sharing: every call on a device uses the same array, and it keeps the previous call's contents.
// Nothing in it has to start at zero - K3 overwrites every row K2 goes on to read (rows [0, B*N) of each
// of the n_tables tables) before K2 is launched on the same stream - so there is no clearing pass.
// Calls that overlap on the GPU (two streams, two host threads on one device) RACE on the array:
// serialise them on one stream, or use the regular entry points with caller-owned tables.
#include "common.cuh"

namespace rl4co {
namespace {

constexpr size_t kStaticTableBytes = (size_t)RL4CO_STATIC_TABLE_BYTES;
constexpr int kMaxDevices = 64;

static __device__ __align__(256) unsigned char g_static_tables[kStaticTableBytes];

// Device code may name the array directly.  This kernel is never launched by the library; it exists so that the
// array is referenced from device code of this translation unit (whole-program compilation may drop a static
// __device__ variable nothing on the device refers to, and cudaGetSymbolAddress would then fail).
__global__ void static_tables_probe_kernel(unsigned long long *address_out) {
    if (threadIdx.x == 0 && blockIdx.x == 0) *address_out = (unsigned long long)(uintptr_t)g_static_tables;
}

// device address of g_static_tables per device ordinal (0 = not fetched yet); racing first calls from
// two host threads fetch the same value twice, which is harmless
float *g_static_ptr[kMaxDevices];

int static_tables_address(float **out) {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess || dev < 0 || dev >= kMaxDevices) {
        set_error("am_rollout_static: cudaGetDevice failed or device ordinal %d outside [0, %d): %s", dev, kMaxDevices,
                  cudaGetErrorString(e));
        return 2;
    }
    if (g_static_ptr[dev] == nullptr) {
        void *p = nullptr;
        e = cudaGetSymbolAddress(&p, g_static_tables);      // the CURRENT device's copy of the array
        if (e != cudaSuccess || p == nullptr) {
            set_error("am_rollout_static: cudaGetSymbolAddress(g_static_tables) failed on device %d: %s", dev,
                      cudaGetErrorString(e));
            return 2;
        }
        g_static_ptr[dev] = static_cast<float *>(p);
    }
    *out = g_static_ptr[dev];
    return 0;
}

}  // namespace
}  // namespace rl4co

using namespace rl4co;

extern "C" int rl4co_am_rollout_static(const rl4co_am_rollout_args *args, const float *h, const float *weight_image,
                                       rl4co_stream_t stream) {
    RL4CO_REQUIRE(args != nullptr, "am_rollout_static: null args");
    if (args->B == 0) return 0;
    RL4CO_REQUIRE(args->tables == nullptr,
                  "am_rollout_static: args->tables must be NULL (the tables live in the library's static scratch); "
                  "use rl4co_am_rollout for caller-owned tables");
    RL4CO_REQUIRE(h != nullptr && weight_image != nullptr, "am_rollout_static: null h / weight_image");
    RL4CO_REQUIRE(args->B > 0 && args->N >= 2 && args->n_tables >= 1 && args->n_tables <= 16,
                  "am_rollout_static: bad sizes B=%lld N=%d n_tables=%d", (long long)args->B, args->N, args->n_tables);
    const unsigned long long rows = (unsigned long long)args->B * (unsigned long long)args->N;
    const unsigned long long max_rows = kStaticTableBytes / (128ull * sizeof(float)) / (unsigned long long)args->n_tables;
    RL4CO_REQUIRE(rows <= max_rows,
                  "am_rollout_static: the tables of B=%lld x N=%d nodes x %d tables need %llu bytes, the static scratch "
                  "holds RL4CO_STATIC_TABLE_BYTES = %llu (at most %llu node rows with %d tables); split the batch or use "
                  "rl4co_am_project_tc + rl4co_am_rollout with caller-owned tables",
                  (long long)args->B, args->N, args->n_tables, rows * args->n_tables * 512ull,
                  (unsigned long long)kStaticTableBytes, max_rows, args->n_tables);
    float *tables = nullptr;
    if (int rc = static_tables_address(&tables)) return rc;
    // K3: h [B*N,128] -> tables [n_tables][B*N][128] in the static array (every row K2 reads is overwritten here)
    if (int rc = rl4co_am_project_tc((int64_t)rows, args->n_tables, h, weight_image, tables, stream)) return rc;
    // K2: the regular rollout on the same stream, reading the static array
    rl4co_am_rollout_args a = *args;
    a.tables = tables;
    return rl4co_am_rollout(&a, stream);
}
