// consensus_i64.cu — instantiates the consensus kernels for one element type
// (kept in its own translation unit so the variants compile in parallel).
#include "consensus_wide2.cuh"

namespace svoc {
int consensus_i64(const Params &p, const Plan &pl, cudaStream_t st) {
    if (pl.variant == 3) return launch_wide2<FixedTr<i64>>(p, pl, st);
    return dispatch_npad<FixedTr<i64>>(p, pl, st);
}
}  // namespace svoc
