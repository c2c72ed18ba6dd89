// consensus_f32.cu — instantiates the consensus kernels for one element type
// (kept in its own translation unit so the variants compile in parallel).
#include "consensus_mid.cuh"

namespace svoc {
int consensus_f32(const Params &p, const Plan &pl, cudaStream_t st) {
    if (pl.variant == 5) return launch_mid<FloatTr<float>>(p, pl, st);
    if (pl.variant == 4) return launch_wide3<FloatTr<float>>(p, pl, st);
    if (pl.variant == 3) return launch_wide2<FloatTr<float>>(p, pl, st);
    if (pl.variant == 1) return launch_wide<FloatTr<float>>(p, pl, st);
    return dispatch_npad<FloatTr<float>>(p, pl, st);
}
}  // namespace svoc
