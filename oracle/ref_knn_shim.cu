// TEST INFRASTRUCTURE ONLY.  extern "C" shim around the unmodified reference kNN (K/simple_knn.cu:192-228).
#include <cuda_runtime.h>
#include <stdint.h>
#include "simple_knn.h"   // resolved through -I <reference simple-knn dir>
extern "C" int ref_knn(int P, const float* points, float* mean_dists, int32_t* nearest_indices) {
    SimpleKNN::knn(P, (float3*)points, mean_dists, nearest_indices);
    return cudaDeviceSynchronize() == cudaSuccess ? 0 : -1;
}
