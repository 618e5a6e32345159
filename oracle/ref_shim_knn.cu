// ref_shim_knn.cu -- TEST INFRASTRUCTURE.  extern "C" door onto the unmodified
// reference SimpleKNN::knn (simple-knn/simple_knn.cu:186-222), built by
// oracle/Makefile from the sources where they lie.
#include <cuda_runtime.h>
#include "simple_knn.h"

extern "C" void ref_dist2(int P, float* points, float* out) { SimpleKNN::knn(P, (float3*)points, out); }
