// consensus_f64.cu — instantiates the consensus kernels for one element type
// (kept in its own translation unit so the variants compile in parallel).
#include "consensus_wide2.cuh"

namespace svoc {
int consensus_f64(const Params &p, const Plan &pl, cudaStream_t st) {
    if (pl.variant == 3) return launch_wide2<FloatTr<double>>(p, pl, st);
    return dispatch_npad<FloatTr<double>>(p, pl, st);
}
}  // namespace svoc
