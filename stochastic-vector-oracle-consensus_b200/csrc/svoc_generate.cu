// svoc_generate.cu — synthetic oracle predictions on the device (the notebook generators,
// ipynb beta :37-68 and gaussian_distribution_for_tests.ipynb:37-52, on a Philox stream).
//
// Stream layout: key = seed (64 bit); counter = (trial_lo, trial_hi, slot, iteration) where
// slot = oracle*D + dim for component draws and 0xFFFF0000 + oracle for the position keys.
// Everything is a pure function of (seed, global trial, oracle, dim).
#include <string>

#include "philox.cuh"
#include "svoc_common.cuh"

namespace svoc {
int fail(const std::string &m);
void count_launch();

struct GenParams {
    svoc_gen_config c;
    i64 T;
    void *pred;
    uint32_t *honest;
    int rb, words;
};

__device__ __forceinline__ float normal_from(uint32_t r0, uint32_t r1) {
    // Box-Muller, cosine branch
    const float u = u01(r0), v = u01(r1);
    return sqrtf(-2.0f * __logf(u)) * __cosf(6.2831853071795864f * v);
}

// Marsaglia & Tsang (2000) gamma(alpha, 1) for alpha >= 1; alpha < 1 via gamma(alpha+1) * U^(1/alpha).
__device__ inline float gamma_draw(float alpha, uint32_t t_lo, uint32_t t_hi, uint32_t slot, uint32_t &iter,
                                   uint32_t k0, uint32_t k1) {
    const bool boost = alpha < 1.0f;
    const float al = boost ? alpha + 1.0f : alpha;
    const float d = al - 1.0f / 3.0f, c = rsqrtf(9.0f * d);
    float out = d;
    for (int guard = 0; guard < 64; ++guard) {
        const Philox4 r = philox4x32_10(t_lo, t_hi, slot, iter++, k0, k1);
        const float x = normal_from(r.v[0], r.v[1]);
        float v = 1.0f + c * x;
        if (v <= 0.0f) continue;
        v = v * v * v;
        const float u = u01(r.v[2]);
        const float x2 = x * x;
        if (u < 1.0f - 0.0331f * x2 * x2 || logf(u) < 0.5f * x2 + d * (1.0f - v + logf(v))) {
            out = d * v;
            if (boost) out *= powf(u01(r.v[3]), 1.0f / alpha);
            break;
        }
    }
    return out;
}

__device__ __forceinline__ float beta_mode(float a, float b) { return (a - 1.0f) / (a + b - 2.0f); }  // ipynb beta :80-81

template <class Out>
__device__ __forceinline__ Out convert_out(float x);
template <> __device__ __forceinline__ float convert_out<float>(float x) { return x; }
template <> __device__ __forceinline__ double convert_out<double>(float x) { return (double)x; }
template <> __device__ __forceinline__ int32_t convert_out<int32_t>(float x) { return (int32_t)(x * 1e6f); }     // to_wsad
template <> __device__ __forceinline__ i64 convert_out<i64>(float x) { return (i64)((double)x * 1e6); }

// One prediction component of the notebook generators.
__device__ __forceinline__ float draw_component(const svoc_gen_config &c, bool is_failing, uint32_t t_lo, uint32_t t_hi,
                                                int oracle, int d, uint32_t k0, uint32_t k1) {
    const int pi = d % c.n_params;
    const float a = (float)c.a[pi], b = (float)c.b[pi];
    const uint32_t slot = (uint32_t)(oracle * c.dimension + d);
    uint32_t iter = 0;
    if (is_failing) {
        if (c.failing == SVOC_FAIL_COLLUDING) {
            if (c.law == SVOC_LAW_BETA) return fminf(fmaxf(beta_mode(a, b) + (float)c.shift, 0.0f), 1.0f);
            return a + (float)c.shift * b;
        }
        const Philox4 r = philox4x32_10(t_lo, t_hi, slot, 0x80000000u, k0, k1);
        const float u = u01(r.v[0]);
        if (c.law == SVOC_LAW_BETA) return u;                 // np.random.uniform(0, 1)
        return a + u * (5.0f * b - a);                        // np.random.uniform(mu, 5 sigma)
    }
    if (c.law == SVOC_LAW_BETA) {
        const float ga = gamma_draw(a, t_lo, t_hi, slot, iter, k0, k1);
        iter = 0x40000000u;
        const float gb = gamma_draw(b, t_lo, t_hi, slot, iter, k0, k1);
        return ga / (ga + gb);
    }
    const Philox4 r = philox4x32_10(t_lo, t_hi, slot, 0, k0, k1);
    return a + b * normal_from(r.v[0], r.v[1]);
}

// failing flags of one round: the F smallest position keys (a uniformly random F-subset)
__device__ __forceinline__ uint32_t position_key(uint32_t t_lo, uint32_t t_hi, int oracle, uint32_t k0, uint32_t k1) {
    return philox4x32_10(t_lo, t_hi, 0xFFFF0000u + (uint32_t)oracle, 0, k0, k1).v[0];
}

template <class Out>
__global__ void __launch_bounds__(256) generate_kernel(const GenParams p) {
    extern __shared__ uint32_t sm[];
    const int N = p.c.n_oracles, D = p.c.dimension, F = p.c.n_failing;
    uint32_t *keys = sm;                                   // [rb][N]
    uint8_t *failing = reinterpret_cast<uint8_t *>(sm + (size_t)p.rb * N);  // [rb][N]
    const i64 r0 = (i64)blockIdx.x * p.rb;
    const int nr = (int)min((i64)p.rb, p.T - r0);
    const uint32_t k0 = (uint32_t)p.c.seed, k1 = (uint32_t)(p.c.seed >> 32);
    const int tid = threadIdx.x, nt = blockDim.x;

    for (int w = tid; w < nr * N; w += nt) {
        const int r = w / N, i = w - r * N;
        const u64 t = (u64)(p.c.trial_offset + r0 + r);
        keys[w] = position_key((uint32_t)t, (uint32_t)(t >> 32), i, k0, k1);
    }
    __syncthreads();
    for (int w = tid; w < nr * N; w += nt) {
        const int r = w / N, i = w - r * N;
        const uint32_t *kk = keys + (size_t)r * N;
        const uint32_t mine = kk[i];
        int rank = 0;
        for (int j = 0; j < N; ++j) rank += (kk[j] < mine || (kk[j] == mine && j < i)) ? 1 : 0;
        failing[w] = rank < F ? 1 : 0;
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
    const int per_round = N * D;
    for (int e = tid; e < nr * per_round; e += nt) {
        const int r = e / per_round, rem = e - r * per_round;
        const int i = rem / D, d = rem - i * D;
        const u64 t = (u64)(p.c.trial_offset + r0 + r);
        const float x = draw_component(p.c, failing[(size_t)r * N + i] != 0, (uint32_t)t, (uint32_t)(t >> 32), i, d, k0, k1);
        out[e] = convert_out<Out>(x);
    }
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
    if (cfg->n_params < 1 || cfg->n_params > SVOC_MAX_LAW_PARAMS) return fail("n_params out of range");
    if (cfg->law != SVOC_LAW_BETA && cfg->law != SVOC_LAW_GAUSSIAN) return fail("unknown law");
    if ((i64)cfg->n_oracles * cfg->dimension >= 0xFFFF0000LL) return fail("n_oracles * dimension too large");
    for (int k = 0; k < cfg->n_params; ++k)
        if (cfg->law == SVOC_LAW_BETA && !(cfg->a[k] > 0 && cfg->b[k] > 0)) return fail("beta parameters must be > 0");
    GenParams p;
    p.c = *cfg;
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
