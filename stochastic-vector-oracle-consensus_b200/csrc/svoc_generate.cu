// svoc_generate.cu — synthetic oracle predictions on the device (the notebook generators,
// ipynb beta :37-68 and gaussian_distribution_for_tests.ipynb:37-52, on a Philox stream).
//
// Stream layout: key = seed (64 bit); counter = (trial_lo, trial_hi, slot, iteration) where
// slot = oracle*D + dim for component draws and 0xFFFF0000 + oracle for the position keys.
// Everything is a pure function of (seed, global trial, oracle, dim).
#include <string>

#include "generate_device.cuh"

namespace svoc {
int fail(const std::string &m);
void count_launch();

struct GenParams {
    svoc_gen_config c;
    i64 T;
    void *pred;
    uint32_t *honest;
    int rb, words;
    PhiloxKeys rk;
};

template <class Out>
__global__ void __launch_bounds__(256) generate_kernel(const GenParams p) {
    extern __shared__ uint32_t sm[];
    const int N = p.c.n_oracles, D = p.c.dimension, F = p.c.n_failing;
    uint32_t *keys = sm;                                   // [rb][N]
    uint8_t *failing = reinterpret_cast<uint8_t *>(sm + (size_t)p.rb * N);  // [rb][N]
    const i64 r0 = (i64)blockIdx.x * p.rb;
    const int nr = (int)min((i64)p.rb, p.T - r0);
    const int tid = threadIdx.x, nt = blockDim.x;

    for (int w = tid; w < nr * N; w += nt) {
        const int r = w / N, i = w - r * N;
        const u64 t = (u64)(p.c.trial_offset + r0 + r);
        keys[w] = position_key((uint32_t)t, (uint32_t)(t >> 32), i, p.rk);
    }
    __syncthreads();
    for (int w = tid; w < nr * N; w += nt) {
        const int r = w / N, i = w - r * N;
        const uint32_t *kk = keys + (size_t)r * N;
        const uint32_t mine = kk[i];
        int rank = 0;
        for (int j = 0; j < N; ++j) rank += (kk[j] < mine || (kk[j] == mine && j < i)) ? 1 : 0;
        // Bernoulli mode: the position key itself is the oracle's uniform draw
        failing[w] = p.c.failing_prob > 0.0 ? (u01(mine) < (float)p.c.failing_prob ? 1 : 0) : (rank < F ? 1 : 0);
    }
    __syncthreads();
    if (p.honest) {
        for (int w = tid; w < nr * p.words; w += nt) {
            const int r = w / p.words, k = w - r * p.words;
            uint32_t bits = 0;
            for (int b = 0; b < 32 && k * 32 + b < N; ++b) bits |= (uint32_t)(failing[(size_t)r * N + k * 32 + b] == 0) << b;
            p.honest[(size_t)r0 * p.words + w] = bits;
        }
    }
    Out *out = reinterpret_cast<Out *>(p.pred) + (size_t)r0 * N * D;
    fill_components<Out, false>(p.c, failing, N, D, nr, p.c.trial_offset + r0, out, tid, nt, p.rk);
}

}  // namespace svoc

extern "C" int svoc_generate(const svoc_gen_config *cfg, int64_t n_rounds, void *pred, uint32_t *honest, void *stream) {
    using namespace svoc;
    if (!cfg) return fail("null generator config");
    if (n_rounds < 0) return fail("negative round count");
    if (n_rounds == 0) return 0;
    if (!pred) return fail("null prediction buffer");
    if (cfg->n_oracles < 1 || cfg->dimension < 1 || cfg->n_oracles > 65535) return fail("bad generator shape");
    if (cfg->n_failing < 0 || cfg->n_failing > cfg->n_oracles) return fail("n_failing out of range");
    if (!(cfg->failing_prob >= 0.0 && cfg->failing_prob <= 1.0)) return fail("failing_prob must lie in [0, 1]");
    if (cfg->n_params < 1 || cfg->n_params > SVOC_MAX_LAW_PARAMS) return fail("n_params out of range");
    if (cfg->law < SVOC_LAW_BETA || cfg->law > SVOC_LAW_ARCTAN_GAUSSIAN) return fail("unknown law");
    if ((i64)cfg->n_oracles * cfg->dimension >= 0xFFFF0000LL) return fail("n_oracles * dimension too large");
    for (int k = 0; k < cfg->n_params; ++k)
        if ((cfg->law == SVOC_LAW_BETA || cfg->law == SVOC_LAW_KUMARASWAMY) && !(cfg->a[k] > 0 && cfg->b[k] > 0))
            return fail("beta / Kumaraswamy parameters must be > 0");
    GenParams p;
    p.c = *cfg;
    p.rk = philox_round_keys(cfg->seed);
    p.T = n_rounds;
    p.pred = pred;
    p.honest = honest;
    p.words = (cfg->n_oracles + 31) / 32;
    const i64 per_round = (i64)cfg->n_oracles * cfg->dimension;
    int rb = (int)std::max<i64>(1, 4096 / per_round);
    while (rb > 1 && (size_t)rb * cfg->n_oracles * 5 > 40000) rb >>= 1;
    p.rb = rb;
    const size_t smem = (size_t)rb * cfg->n_oracles * 4 + (((size_t)rb * cfg->n_oracles + 3) & ~(size_t)3);
    if (smem > 48 * 1024) return fail("n_oracles too large for the generator");
    const i64 grid = (n_rounds + rb - 1) / rb;
    if (grid > 0x7fffffffLL) return fail("too many rounds for one launch");
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    switch (cfg->dtype) {
        case SVOC_F32: generate_kernel<float><<<(unsigned)grid, 256, smem, st>>>(p); break;
        case SVOC_F64: generate_kernel<double><<<(unsigned)grid, 256, smem, st>>>(p); break;
        case SVOC_I32: generate_kernel<int32_t><<<(unsigned)grid, 256, smem, st>>>(p); break;
        case SVOC_I64: generate_kernel<i64><<<(unsigned)grid, 256, smem, st>>>(p); break;
        default: return fail("bad output dtype");
    }
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(std::string("kernel launch: ") + cudaGetErrorString(e));
    count_launch();
    return 0;
}
