// tcgen05 candidate pass of the kNN engine (placeholder until the tensor-core kernel lands).
#include "common.cuh"
#include "knn_internal.cuh"

int knn_cand_tc_launch(cb_ctx*, const float*, int, const float*, const float*, int, int, int, int, int, int,
                       int32_t*, float*, cudaStream_t) {
    return CB_EUNSUPPORTED;
}
double knn_cand_tc_gamma(int) { return 0.0; }
extern "C" const char* cb_knn_variant(const cb_ctx* ctx) { return "simt"; }
