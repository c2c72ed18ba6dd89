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
        // First attempt of both gammas from ONE Philox call: the two Box-Muller branches (cosine, sine) are
        // independent normals, words 2 and 3 the two squeeze uniforms.  Marsaglia-Tsang accepts ~95 % of the
        // attempts for alpha >= 1, so the per-gamma loops below (fresh counters) run for about one draw in ten.
        float ga = -1.0f, gb = -1.0f;
        if (a >= 1.0f && b >= 1.0f) {
            const Philox4 r = philox4x32_10(t_lo, t_hi, slot, 0x20000000u, k0, k1);
            const float rad = sqrtf(-2.0f * __logf(u01(r.v[0])));
            float sn, cs;
            __sincosf(6.2831853071795864f * u01(r.v[1]), &sn, &cs);
            auto attempt = [](float alpha, float x, float u) {
                const float d = alpha - 1.0f / 3.0f, cc = rsqrtf(9.0f * d);
                float v = 1.0f + cc * x;
                if (v <= 0.0f) return -1.0f;
                v = v * v * v;
                const float x2 = x * x;
                if (u < 1.0f - 0.0331f * x2 * x2 || logf(u) < 0.5f * x2 + d * (1.0f - v + logf(v))) return d * v;
                return -1.0f;
            };
            ga = attempt(a, rad * cs, u01(r.v[2]));
            gb = attempt(b, rad * sn, u01(r.v[3]));
        }
        if (ga < 0.0f) ga = gamma_draw(a, t_lo, t_hi, slot, iter, k0, k1);
        if (gb < 0.0f) {
            iter = 0x40000000u;
            gb = gamma_draw(b, t_lo, t_hi, slot, iter, k0, k1);
        }
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
