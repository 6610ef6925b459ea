// (e) k-nearest neighbours (placeholder translation unit; kernels are added below in a later edit)
#include "common.cuh"
extern "C" {
size_t gsr_knn_temp_bytes(int64_t P) { (void)P; return 0; }
int gsr_knn(int64_t P, int K, const float* points, float* mean_dist2, int32_t* idx, float* dist2, void* tmp,
            size_t tmp_bytes, gsr_stream_t stream) {
    (void)P; (void)K; (void)points; (void)mean_dist2; (void)idx; (void)dist2; (void)tmp; (void)tmp_bytes; (void)stream;
    return GSR_ERR_UNSUPPORTED;
}
}
