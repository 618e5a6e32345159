// htgs_sort_pairs_static: the pair sort of radix_sort.cu for callers that are not allowed to
// allocate device memory.
//
// Who that is: code that runs inside somebody else's memory budget.  A plugin loaded into a
// host application whose device memory is one arena reserved at start-up; a serving process
// whose pool is sized and accounted up front, so an extra cudaMalloc either fails or breaks the
// accounting; a latency-critical loop, where cudaMalloc/cudaFree synchronise the whole device
// and stall every other stream.  Such a caller can still pass the input and output arrays it
// already owns, but it has nowhere to take the sort's scratch (digit histograms + look-back
// words) from.  Here that scratch lives in a fixed array inside the library image instead.
//
// Consequences the caller has to know (also in include/htgs.h):
//   * the array is a fixed HTGS_SORT_STATIC_SCRATCH_BYTES (about 2 MB) and exists on every GPU
//     that loads the library, from load until the process exits;
//   * it bounds the problem size: more than HTGS_SORT_STATIC_MAX_PAIRS pairs is refused;
//   * every call on a device uses the same array.  The part a call needs is cleared at its start,
//     so stale contents of the previous call do no harm, but two calls whose GPU work overlaps
//     (different streams, or different host threads) would overwrite each other's counters:
//     keep calls on one device ordered;
//   * the first call on each device resolves the array's device address (no allocation, but a
//     one-off runtime query); make that first call outside any stream capture.  The address is
//     remembered per device ordinal for the life of the process: after cudaDeviceReset() the
//     library image is reloaded and the remembered address is stale, so a process that resets a
//     device must not call this entry point on it again.
#include "../../include/htgs.h"
#include "common.cuh"
#include <mutex>

namespace {

// Scratch of the sort, inside the library's device image.  In HOST code this name is only a
// handle for the runtime, not an address a kernel could use: the device address is fetched once
// per device with cudaGetSymbolAddress below and that pointer is what the kernels receive.
static __device__ __align__(256) unsigned char g_sort_static_scratch[HTGS_SORT_STATIC_SCRATCH_BYTES];

constexpr int kMaxDevices = 64;
void* g_scratch_addr[kMaxDevices] = {};  // device address of the array, per device ordinal
std::mutex g_setup_mutex;

// Set-up, once per device, with that device current.
int scratch_for_current_device(void** out) {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) {
        htgs::set_error("htgs_sort_pairs_static: cudaGetDevice failed: %s", cudaGetErrorString(e));
        return 1;
    }
    if (dev < 0 || dev >= kMaxDevices) {
        htgs::set_error("htgs_sort_pairs_static: device ordinal %d not supported (max %d)", dev, kMaxDevices - 1);
        return 1;
    }
    std::lock_guard<std::mutex> lock(g_setup_mutex);
    if (g_scratch_addr[dev] == nullptr) {
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_sort_static_scratch);
        if (e != cudaSuccess) {
            htgs::set_error("htgs_sort_pairs_static: cudaGetSymbolAddress failed on device %d: %s", dev,
                            cudaGetErrorString(e));
            return 1;
        }
        if (p == nullptr) {
            htgs::set_error("htgs_sort_pairs_static: cudaGetSymbolAddress returned a null address on device %d", dev);
            return 1;
        }
        g_scratch_addr[dev] = p;
    }
    *out = g_scratch_addr[dev];
    return 0;
}

}  // namespace

extern "C" int htgs_sort_pairs_static(int64_t n, int end_bit, uint64_t* keys_in, uint32_t* values_in, uint64_t* keys_out,
                                      uint32_t* values_out, void* stream) {
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    if (n <= 0) return 0;
    if (n > HTGS_SORT_STATIC_MAX_PAIRS) {
        htgs::set_error("htgs_sort_pairs_static: n=%lld exceeds HTGS_SORT_STATIC_MAX_PAIRS=%lld (the built-in scratch "
                        "array is fixed; use htgs_sort_pairs with caller-provided scratch for larger inputs)",
                        (long long)n, (long long)HTGS_SORT_STATIC_MAX_PAIRS);
        return 1;
    }
    if (htgs::sort_scratch_bytes(n) > (size_t)HTGS_SORT_STATIC_SCRATCH_BYTES) {
        htgs::set_error("htgs_sort_pairs_static: internal scratch array too small for n=%lld", (long long)n);
        return 1;
    }
    if (!keys_in || !values_in || !keys_out || !values_out) {
        htgs::set_error("htgs_sort_pairs_static: null pointer");
        return 1;
    }
    void* scratch = nullptr;
    if (scratch_for_current_device(&scratch)) return 1;
    // sort_pairs clears the histogram / ticket / look-back words this n needs before using them
    // (cudaMemsetAsync on `st`), which is exactly the part that must start at zero.
    if (htgs::sort_pairs(n, end_bit, keys_in, values_in, keys_out, values_out, scratch, st)) return 1;
    if ((htgs::sort_num_passes(end_bit) & 1) == 0) {  // even pass count: the result sits in the input arrays
        HTGS_CUDA_CHECK(cudaMemcpyAsync(keys_out, keys_in, (size_t)n * sizeof(uint64_t), cudaMemcpyDeviceToDevice, st));
        HTGS_CUDA_CHECK(cudaMemcpyAsync(values_out, values_in, (size_t)n * sizeof(uint32_t), cudaMemcpyDeviceToDevice, st));
    }
    return 0;
}
