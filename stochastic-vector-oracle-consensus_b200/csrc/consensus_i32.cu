// consensus_i32.cu — instantiates the consensus kernels for one element type
// (kept in its own translation unit so the variants compile in parallel).
#include "consensus_mid.cuh"

namespace svoc {
int consensus_i32(const Params &p, const Plan &pl, cudaStream_t st) {
    if (pl.variant == 5) return launch_mid<FixedTr<int32_t>>(p, pl, st);
    if (pl.variant == 4) return launch_wide3<FixedTr<int32_t>>(p, pl, st);
    if (pl.variant == 3) return launch_wide2<FixedTr<int32_t>>(p, pl, st);
    if (pl.variant == 1) return launch_wide<FixedTr<int32_t>>(p, pl, st);
    return dispatch_npad<FixedTr<int32_t>>(p, pl, st);
}
}  // namespace svoc
