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

// Division by a block-uniform divisor: one multiply-high instead of the ~20-instruction division sequence.
// m = floor(2^32 / d) + 1 gives the exact quotient while e * d < 2^32; `exact` says whether the caller's range
// allows it (else the plain division runs).
struct UniformDiv {
    uint32_t d, m;
    bool exact;
    __device__ UniformDiv(uint32_t dd, uint32_t max_e) : d(dd), m(dd > 1u ? 0xffffffffu / dd + 1u : 0u),
        exact((u64)max_e * dd < 0x100000000ull) {}
    __device__ __forceinline__ uint32_t operator()(uint32_t e) const {
        if (d <= 1u) return e;
        return exact ? __umulhi(e, m) : e / d;
    }
};

// Every component of `nr` rounds (out[r][i][d], trials trial0 + r) by the threads tid, tid + nt, ... of a CTA.
// `flags` holds one byte per (round, oracle): non-zero = failing (FLAG_IS_HONEST: non-zero = honest).  The values
// are those of draw_component.  For the beta law with a, b >= 1 the lanes of a warp run DECOUPLED: a trip of
// the loop is one attempt (one Philox call) of whatever element the lane is at, and a lane moves to its next
// element as soon as both gammas are accepted.  With draw_component a warp repeats the trip of one element
// until its slowest lane is through: at a = 10, b = 15 a lane needs a second trip 0.5 % of the time and a
// warp 15 % (1.15 trips per element; near a = b = 1, 10 % and 96 %, two trips); decoupled, a warp runs the
// lanes' own average plus the slack of its slowest lane.  Colluding rows need no draw and cost no trip; the
// (round, oracle, dim) of an element comes from multiply-high divisions.
template <class Out, bool FLAG_IS_HONEST>
__device__ __forceinline__ void fill_components(const svoc_gen_config &c, const uint8_t *flags, int N, int D, int nr,
                                                i64 trial0, Out *out, int tid, int nt, const PhiloxKeys &rk) {
    const uint32_t k0 = rk.k[0], k1 = rk.k[1];
    const uint32_t per_round = (uint32_t)N * (uint32_t)D, total = (uint32_t)nr * per_round;
    const UniformDiv by_round(per_round, total), by_dim((uint32_t)D, per_round), by_par((uint32_t)c.n_params, (uint32_t)D);
    bool decoupled = c.law == SVOC_LAW_BETA;
    for (int k = 0; k < c.n_params; ++k) decoupled = decoupled && c.a[k] >= 1.0 && c.b[k] >= 1.0;
    if (!decoupled) {
        for (uint32_t e = tid; e < total; e += nt) {
            const uint32_t r = by_round(e), rem = e - r * per_round, i = by_dim(rem), d = rem - i * (uint32_t)D;
            const u64 t = (u64)(trial0 + r);
            const bool failing = (flags[r * N + i] != 0) != FLAG_IS_HONEST;
            out[e] = convert_out<Out>(draw_component(c, failing, (uint32_t)t, (uint32_t)(t >> 32), (int)i, (int)d, k0, k1));
        }
        return;
    }
    const bool colluding = c.failing == SVOC_FAIL_COLLUDING;
    auto attempt = [](float d, float cc, float x, float u) {          // Marsaglia-Tsang, as in draw_component
        float v = 1.0f + cc * x;
        if (v <= 0.0f) return -1.0f;
        v = v * v * v;
        const float x2 = x * x;
        if (u < 1.0f - 0.0331f * x2 * x2 || __logf(u) < 0.5f * x2 + d * (1.0f - v + __logf(v))) return d * v;
        return -1.0f;
    };
    uint32_t e = tid, it = 0, slot = 0, t_lo = 0, t_hi = 0, cur_pi = 0xffffffffu;
    float da = 0, ca = 0, db = 0, cb = 0, ga = -1.0f, gb = -1.0f, centre = 0;
    bool failing = false;
    // Moves the lane to its next element that needs a draw (colluding rows are written on the way).  Called before the loop and from its accept branch, so the loop body holds the one Philox site —
    // a "fresh element" flag tested at the loop head let the compiler split the body in two copies, and lanes
    // in different copies no longer ran their trips together.
    auto next_element = [&]() {
        while (e < total) {
            const uint32_t r = by_round(e), rem = e - r * per_round, i = by_dim(rem), d = rem - i * (uint32_t)D;
            const uint32_t pi = d - by_par(d) * (uint32_t)c.n_params;
            if (pi != cur_pi) {
                // the constants of this law parameter; a thread mostly stays on one dimension (its elements are
                // nt apart), so the two double -> float conversions and the two rsqrt run once, not per element
                cur_pi = pi;
                const float a = (float)c.a[pi], b = (float)c.b[pi];
                da = a - 1.0f / 3.0f; ca = rsqrtf(9.0f * da); db = b - 1.0f / 3.0f; cb = rsqrtf(9.0f * db);
                if (colluding) centre = fminf(fmaxf(beta_mode(a, b) + (float)c.shift, 0.0f), 1.0f);
            }
            failing = (flags[r * N + i] != 0) != FLAG_IS_HONEST;
            if (failing && colluding) {                               // no draw: mode + shift, clipped
                out[e] = convert_out<Out>(centre);
                e += nt;
                continue;
            }
            const u64 t = (u64)(trial0 + r);
            t_lo = (uint32_t)t; t_hi = (uint32_t)(t >> 32);
            slot = rem;                                               // oracle * D + dim
            ga = -1.0f; gb = -1.0f; it = 0;
            return;
        }
    };
    next_element();
    while (e < total) {
        const Philox4 r = philox4x32_10(t_lo, t_hi, slot, failing ? 0x80000000u : 0x20000000u + it, rk);
        const float rad = sqrtf(-2.0f * __logf(u01(r.v[0])));
        float sn, cs;
        __sincosf(6.2831853071795864f * u01(r.v[1]), &sn, &cs);
        if (ga < 0.0f) ga = attempt(da, ca, rad * cs, u01(r.v[2]));
        if (gb < 0.0f) gb = attempt(db, cb, rad * sn, u01(r.v[3]));
        ++it;
        if (failing || (ga >= 0.0f && gb >= 0.0f) || it == 64u) {
            ga = ga < 0.0f ? da : ga;                                 // 64 rejections in a row: never in practice
            gb = gb < 0.0f ? db : gb;
            float x = __fdividef(ga, ga + gb);
            if (failing) x = u01(r.v[0]);                             // np.random.uniform(0, 1)
            out[e] = convert_out<Out>(x);
            e += nt;
            next_element();
        }
    }
}

// failing flags of one round: the F smallest position keys (a uniformly random F-subset)
__device__ __forceinline__ uint32_t position_key(uint32_t t_lo, uint32_t t_hi, int oracle, const PhiloxKeys &rk) {
    return philox4x32_10(t_lo, t_hi, 0xFFFF0000u + (uint32_t)oracle, 0, rk).v[0];
}

}  // namespace svoc
