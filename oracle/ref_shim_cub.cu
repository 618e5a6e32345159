// ref_shim_cub.cu -- TEST / MEASUREMENT INFRASTRUCTURE.  The library sort the reference rasterizer calls,
// cub::DeviceRadixSort::SortPairs<uint64_t, uint32_t>(temp, bytes, keys_in, keys_out, values_in, values_out, n, 0,
// end_bit) (cuda_rasterizer/rasterizer_impl.cu:304-309; temp sized the same way at :187-190), behind an extern "C"
// door so bench.py can time it on the very keys this library's own sort gets ("sort_vs_cub").  CUB comes from the
// CUDA toolkit (it is not vendored by the reference); compiled for sm_100a it instantiates CUB's SM100 onesweep
// policy, i.e. exactly the sort kernel the reference runs on a B200.  Nothing here is part of the product.
#include <cub/cub.cuh>
#include <cuda_runtime.h>
#include <cstdint>

extern "C" {

size_t ref_cub_temp_bytes(int64_t n) {
    size_t bytes = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, bytes, (const uint64_t*)nullptr, (uint64_t*)nullptr, (const uint32_t*)nullptr,
                                    (uint32_t*)nullptr, (int)n);
    return bytes;
}

// Legacy default stream, like the reference.  Inputs are preserved (the non-DoubleBuffer overload).
int ref_cub_sort_pairs(int64_t n, int end_bit, const uint64_t* keys_in, const uint32_t* values_in, uint64_t* keys_out,
                       uint32_t* values_out, void* temp, size_t temp_bytes) {
    cudaError_t e = cub::DeviceRadixSort::SortPairs(temp, temp_bytes, keys_in, keys_out, values_in, values_out, (int)n, 0,
                                                    end_bit);
    return e == cudaSuccess ? 0 : (int)e;
}
}
