// generate_device.cuh — device-side samplers of the synthetic-workload generator (see svoc_generate.cu):
// Philox4x32-10 counters keyed by (seed, global trial, oracle, dimension); beta via two Marsaglia-Tsang
// gammas; failing positions = the F smallest per-oracle position keys.
#pragma once
#include "philox.cuh"
#include "svoc_common.cuh"

namespace svoc {

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
            float centre;
            if (c.law == SVOC_LAW_BETA) centre = beta_mode(a, b);
            else if (c.law == SVOC_LAW_KUMARASWAMY) centre = powf((a - 1.0f) / (a * b - 1.0f), 1.0f / a);   // ipynb beta :77-78
            else if (c.law == SVOC_LAW_ARCTAN_GAUSSIAN) centre = a;
            else return a + (float)c.shift * b;
            return fminf(fmaxf(centre + (float)c.shift, 0.0f), 1.0f);
        }
        const Philox4 r = philox4x32_10(t_lo, t_hi, slot, 0x80000000u, k0, k1);
        const float u = u01(r.v[0]);
        if (c.law != SVOC_LAW_GAUSSIAN) return u;             // np.random.uniform(0, 1)
        return a + u * (5.0f * b - a);                        // np.random.uniform(mu, 5 sigma)
    }
    if (c.law == SVOC_LAW_KUMARASWAMY) {                      // inverse CDF: x = (1 - (1 - u)^(1/b))^(1/a)
        const Philox4 r = philox4x32_10(t_lo, t_hi, slot, 0, k0, k1);
        return powf(1.0f - powf(1.0f - u01(r.v[0]), 1.0f / b), 1.0f / a);
    }
    if (c.law == SVOC_LAW_ARCTAN_GAUSSIAN) {                  // gaussian_algorithm_demo.ipynb:46-55
        const Philox4 r = philox4x32_10(t_lo, t_hi, slot, 0, k0, k1);
        const float mu = tanf(3.14159265358979f * (a - 0.5f));
        return atanf(mu + b * normal_from(r.v[0], r.v[1])) * 0.318309886183791f + 0.5f;
    }
    if (c.law == SVOC_LAW_BETA) {
        if (a >= 1.0f && b >= 1.0f) {
            // Both gammas (Marsaglia-Tsang) advance together: ONE Philox call per trip gives the two Box-Muller
            // branches (cosine, sine: independent normals) and the two squeeze uniforms.  A trip accepts ~97 % of
            // its attempts, but a warp repeats it while ANY lane still misses a gamma — separate per-gamma
            // retry loops, and the precise logf of the second test, cost a warp three times as much.
            const float da = a - 1.0f / 3.0f, ca = rsqrtf(9.0f * da), db = b - 1.0f / 3.0f, cb = rsqrtf(9.0f * db);
            auto attempt = [](float d, float cc, float x, float u) {
                float v = 1.0f + cc * x;
                if (v <= 0.0f) return -1.0f;
                v = v * v * v;
                const float x2 = x * x;
                if (u < 1.0f - 0.0331f * x2 * x2 || __logf(u) < 0.5f * x2 + d * (1.0f - v + __logf(v))) return d * v;
                return -1.0f;
            };
            float ga = -1.0f, gb = -1.0f;
            for (uint32_t it = 0; it < 64u && (ga < 0.0f || gb < 0.0f); ++it) {
                const Philox4 r = philox4x32_10(t_lo, t_hi, slot, 0x20000000u + it, k0, k1);
                const float rad = sqrtf(-2.0f * __logf(u01(r.v[0])));
                float sn, cs;
                __sincosf(6.2831853071795864f * u01(r.v[1]), &sn, &cs);
                if (ga < 0.0f) ga = attempt(da, ca, rad * cs, u01(r.v[2]));
                if (gb < 0.0f) gb = attempt(db, cb, rad * sn, u01(r.v[3]));
            }
            ga = ga < 0.0f ? da : ga;             // 64 rejections in a row: never in practice
            gb = gb < 0.0f ? db : gb;
            return __fdividef(ga, ga + gb);
        }
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

}  // namespace svoc
