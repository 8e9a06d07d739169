// search_tc.cu — tensor-core (tcgen05/TMEM/TMA) search.  Placeholder until the kernel lands:
// datasets are prepared without the fp16 operand, so the engine dispatch never reaches here.
#include "tm_common.cuh"
int tm_prepare_dataset(ANNkdtree *h, cudaStream_t s) { (void)h; (void)s; return 0; }
int tm_search_tc(ANNkdtree *h, const float *d_q, int64_t nq, int k, int *d_idx, float *d_err, cudaStream_t s) {
    (void)h; (void)d_q; (void)nq; (void)k; (void)d_idx; (void)d_err; (void)s;
    tm_set_error("tensor-core engine not built");
    return TM_EINVAL;
}
