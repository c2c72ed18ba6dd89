// montecarlo_f64.cu — instantiates the fused Monte-Carlo kernels for one element type.
#include "montecarlo.cuh"

namespace svoc {
int montecarlo_f64(const Params &p, const McParams &m, int npad, int threads, size_t smem, int grid, cudaStream_t st) {
    return dispatch_mc<FloatTr<double>>(p, m, npad, threads, smem, grid, st);
}
}  // namespace svoc
