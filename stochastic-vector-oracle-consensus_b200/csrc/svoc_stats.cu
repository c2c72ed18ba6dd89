// svoc_stats.cu — per-oracle divergence accumulators and reliability histograms over many
// rounds (layout documented in include/svoc.h).  CTA-local shared-memory histograms, then one
// global atomic per touched bin; the buffers are what multi-GPU runs all-reduce.
#include <string>

#include "svoc_common.cuh"

namespace svoc {
int fail(const std::string &m);
void count_launch();

constexpr int HIST = 256;

template <class V, bool FIXED>
__global__ void __launch_bounds__(256) stats_kernel(i64 T, int N, int words, const V *__restrict__ rel1,
                                                    const V *__restrict__ rel2, const int32_t *__restrict__ status,
                                                    const uint32_t *__restrict__ reliable, const V *__restrict__ qr,
                                                    u64 *__restrict__ counts, double *__restrict__ sums) {
    extern __shared__ unsigned char raw[];
    unsigned *h = reinterpret_cast<unsigned *>(raw);          // [16 status][256 rel1][256 rel2][N unreliable]
    const int nh = 16 + 2 * HIST + N;
    for (int k = threadIdx.x; k < nh; k += blockDim.x) h[k] = 0;
    __syncthreads();
    double s1 = 0, s2 = 0, q1 = 0, q2 = 0;
    const i64 stride = (i64)gridDim.x * blockDim.x;
    for (i64 t = (i64)blockIdx.x * blockDim.x + threadIdx.x; t < T; t += stride) {
        const int st = status[t];
        atomicAdd(&h[(st >= 0 && st < 14) ? 2 + st : 15], 1u);
        if (st != 0) continue;
        double a, b;
        if (FIXED) { a = (double)rel1[t] * 1e-6; b = (double)rel2[t] * 1e-6; }
        else { a = (double)rel1[t]; b = (double)rel2[t]; }
        int ba = (int)(a * HIST), bb = (int)(b * HIST);
        ba = ba < 0 ? 0 : (ba > HIST - 1 ? HIST - 1 : ba);
        bb = bb < 0 ? 0 : (bb > HIST - 1 ? HIST - 1 : bb);
        atomicAdd(&h[16 + ba], 1u);
        atomicAdd(&h[16 + HIST + bb], 1u);
        s1 += a; s2 += b; q1 += a * a; q2 += b * b;
        if (reliable) {
            for (int w = 0; w < words; ++w) {
                uint32_t bad = ~reliable[(size_t)t * words + w];
                if (w == words - 1 && (N & 31)) bad &= (1u << (N & 31)) - 1u;
                while (bad) {
                    const int i = __ffs(bad) - 1;
                    bad &= bad - 1;
                    atomicAdd(&h[16 + 2 * HIST + w * 32 + i], 1u);
                }
            }
        }
    }
    // block reduce of the four float sums
    __shared__ double red[4][8];
    double v[4] = {s1, s2, q1, q2};
    for (int k = 0; k < 4; ++k)
        for (int o = 16; o; o >>= 1) v[k] += __shfl_xor_sync(0xffffffffu, v[k], o);
    if ((threadIdx.x & 31) == 0)
        for (int k = 0; k < 4; ++k) red[k][threadIdx.x >> 5] = v[k];
    __syncthreads();
    if (threadIdx.x < 4) {
        double s = 0;
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w) s += red[threadIdx.x][w];
        if (s != 0) atomicAdd(&sums[threadIdx.x], s);
    }
    for (int k = threadIdx.x; k < nh; k += blockDim.x) {
        const unsigned c = h[k];
        if (!c) continue;
        int slot;
        if (k < 16) slot = k;                               // [2..15] status codes
        else if (k < 16 + 2 * HIST) slot = k;               // histograms live at [16..527]
        else slot = SVOC_STATS_COUNTS_BASE + (k - 16 - 2 * HIST);
        atomicAdd(&counts[slot], (u64)c);
        if (k >= 2 && k < 16) atomicAdd(&counts[0], (u64)c);
        if (k == 2) atomicAdd(&counts[1], (u64)c);
    }
}

// per-oracle sum of qr over status-0 rounds: one CTA column-reduces a slab of rounds
template <class V, bool FIXED>
__global__ void __launch_bounds__(256) qr_sum_kernel(i64 T, int N, const int32_t *__restrict__ status,
                                                     const V *__restrict__ qr, u64 *__restrict__ counts,
                                                     double *__restrict__ sums) {
    const i64 per = (T + gridDim.x - 1) / gridDim.x;
    const i64 t0 = (i64)blockIdx.x * per, t1 = min(T, t0 + per);
    for (int i = threadIdx.x; i < N; i += blockDim.x) {
        if (FIXED) {
            u64 s = 0;
            for (i64 t = t0; t < t1; ++t)
                if (status[t] == 0) s += (u64)qr[(size_t)t * N + i];
            if (s) atomicAdd(&counts[SVOC_STATS_COUNTS_BASE + N + i], s);
        } else {
            double s = 0;
            for (i64 t = t0; t < t1; ++t)
                if (status[t] == 0) s += (double)qr[(size_t)t * N + i];
            if (s != 0) atomicAdd(&sums[SVOC_STATS_SUMS_BASE + i], s);
        }
    }
}

template <class V, bool FIXED>
static int run_stats(const svoc_config &c, i64 T, const svoc_outputs &o, int64_t *counts, double *sums, cudaStream_t st) {
    const int N = c.n_oracles, words = (N + 31) / 32;
    const size_t smem = (size_t)(16 + 2 * HIST + N) * sizeof(unsigned);
    const int grid = (int)std::min<i64>((T + 255) / 256, 148 * 4);
    stats_kernel<V, FIXED><<<grid, 256, smem, st>>>(T, N, words, (const V *)o.rel1, (const V *)o.rel2, o.status,
                                                    o.reliable, (const V *)o.qr, (u64 *)counts, sums);
    count_launch();
    if (o.qr) {
        const int g2 = (int)std::min<i64>((T + 63) / 64, 148 * 8);
        qr_sum_kernel<V, FIXED><<<g2, 256, 0, st>>>(T, N, o.status, (const V *)o.qr, (u64 *)counts, sums);
        count_launch();
    }
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(std::string("kernel launch: ") + cudaGetErrorString(e));
    return 0;
}

}  // namespace svoc

extern "C" int svoc_accumulate_stats(const svoc_config *cfg, int64_t n_rounds, const svoc_outputs *out, int64_t *counts,
                                     double *sums, void *stream) {
    using namespace svoc;
    if (!cfg || !out || !counts || !sums) return fail("null argument");
    if (n_rounds < 0) return fail("negative round count");
    if (n_rounds == 0) return 0;
    if (!out->rel1 || !out->rel2 || !out->status) return fail("rel1, rel2 and status are required");
    if (cfg->n_oracles < 1 || cfg->n_oracles > 8192) return fail("n_oracles out of range for the accumulators");
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    switch (cfg->dtype) {
        case SVOC_I32:
        case SVOC_I64: return run_stats<i64, true>(*cfg, n_rounds, *out, counts, sums, st);
        case SVOC_F32: return run_stats<float, false>(*cfg, n_rounds, *out, counts, sums, st);
        case SVOC_F64: return run_stats<double, false>(*cfg, n_rounds, *out, counts, sums, st);
        default: return fail("bad dtype");
    }
}
