// svoc_aux.cu — small float helpers of the notebook that sit beside the consensus round.
#include <algorithm>
#include <string>

#include "svoc_common.cuh"

namespace svoc {
int fail(const std::string &m);
void count_launch();

// expected_reliability (ipynb beta :260-267): 1 - 2*mean(|A - median(A)|), where np.median has
// no axis, i.e. the scalar median of the flattened array.  One warp per array; the two middle
// order statistics come from an MSB-first bisection over order-preserving 64-bit keys.
__global__ void __launch_bounds__(128) expected_reliability_kernel(const double *__restrict__ a, i64 n_arrays,
                                                                   int n, double *__restrict__ out) {
    const int lane = threadIdx.x & 31;
    const i64 w = (i64)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (w >= n_arrays) return;
    const double *x = a + (size_t)w * n;
    auto kth = [&](int k) -> u64 {
        u64 prefix = 0;
        for (int bit = 63; bit >= 0; --bit) {
            const u64 cand = prefix | (1ull << bit);
            unsigned cnt = 0;
            for (int i = lane; i < n; i += 32) cnt += okey(x[i]) < cand ? 1u : 0u;
            cnt = __reduce_add_sync(0xffffffffu, cnt);
            if ((int)cnt <= k) prefix = cand;
        }
        return prefix;
    };
    double lo, hi;
    unkey(kth((n - 1) / 2), lo);
    unkey(kth(n / 2), hi);
    const double med = 0.5 * (lo + hi);
    double s = 0;
    for (int i = lane; i < n; i += 32) s += fabs(x[i] - med);
    for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0) out[w] = 1.0 - 2.0 * (s / (double)n);
}

// rank_array (ipynb beta :342-352): stable ascending position by counting, reversed.
__global__ void __launch_bounds__(128) rank_array_kernel(const double *__restrict__ v, i64 n_lists, int n,
                                                         int32_t *__restrict__ ranks) {
    const i64 total = n_lists * n;
    for (i64 e = (i64)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (i64)gridDim.x * blockDim.x) {
        const i64 t = e / n;
        const int i = (int)(e - t * n);
        const double *x = v + (size_t)t * n;
        const double mine = x[i];
        int pos = 0;
        for (int j = 0; j < n; ++j) pos += (x[j] < mine || (x[j] == mine && j < i)) ? 1 : 0;
        ranks[e] = n - 1 - pos;
    }
}
}  // namespace svoc

extern "C" int svoc_rank_array(const double *values, int64_t n_lists, int32_t n, int32_t *ranks, void *stream) {
    using namespace svoc;
    if (n_lists < 0 || n < 1) return fail("bad sizes");
    if (n_lists == 0) return 0;
    if (!values || !ranks) return fail("null pointer");
    const i64 total = n_lists * n;
    const unsigned grid = (unsigned)std::min<i64>((total + 127) / 128, 148 * 16);
    rank_array_kernel<<<grid, 128, 0, reinterpret_cast<cudaStream_t>(stream)>>>(values, n_lists, n, ranks);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(std::string("kernel launch: ") + cudaGetErrorString(e));
    count_launch();
    return 0;
}

extern "C" int svoc_expected_reliability(const double *pred, int64_t n_arrays, int32_t n_values, double *out,
                                         void *stream) {
    using namespace svoc;
    if (n_arrays < 0 || n_values < 1) return fail("bad sizes");
    if (n_arrays == 0) return 0;
    if (!pred || !out) return fail("null pointer");
    const i64 grid = (n_arrays + 3) / 4;
    expected_reliability_kernel<<<(unsigned)grid, 128, 0, reinterpret_cast<cudaStream_t>(stream)>>>(pred, n_arrays,
                                                                                                  n_values, out);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(std::string("kernel launch: ") + cudaGetErrorString(e));
    count_launch();
    return 0;
}
