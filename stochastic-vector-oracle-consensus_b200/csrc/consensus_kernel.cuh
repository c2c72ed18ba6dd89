// consensus_kernel.cuh — the fused two-pass consensus round on sm_100a.
//
// One kernel does a whole round (first-pass centre, per-oracle divergence, selection of the
// R most-agreeing oracles, second-pass centre, both reliabilities, optional skew/kurtosis)
// for a block of rounds staged once into shared memory by a TMA bulk copy, so every input
// byte is read from HBM exactly once.  Reference semantics restated (paths relative to the
// reference checkout):
//   contract/src/contract.cairo:442-503 (constrained), :370-434 (unconstrained)
//   contract/src/math.cairo:113-126, 170-189, 208-292, 320-363
//   contract/src/sort.cairo:96-102 (tie order), contract.cairo:345-363 (reliable flags)
//   float form: drafts/beta_kumaraswamy_algorithm_demo copy.ipynb :342-371
//
// Two variants of the same skeleton:
//   small<NPAD>  N <= NPAD <= 64: a thread owns one (round, dimension) column in registers
//                and orders it with a Batcher network;
//   generic      any N: columns stay in shared memory (or global when a round does not
//                fit) and order statistics come from an MSB-first bisection over
//                order-preserving keys with counting passes.
#pragma once
#include <limits>
#include <string>

#include "svoc_common.cuh"

namespace svoc {

// ------------------------------------------------------------------------------ parameters
struct Params {
    const void *pred;
    i64 T;
    int N, D, F, R;
    int constrained;
    i64 ms_w;
    double ms_f;
    int centre, centre2, tie, want_moments, flags;
    void *essence1, *qr, *rel1;
    uint32_t *reliable;
    void *consensus, *rel2, *skew, *kurt;
    int32_t *status;
    int32_t *rank;
    int rb;          // rounds per CTA
    int words;       // ceil(N/32)
    int tile_smem;   // 1: rounds staged in shared memory; 0: read from global
    // wide variant (consensus_wide.cuh): parts per column, rows per part, key column stride (bytes),
    // threads per row in the row phases, 1 when every round can be fetched by a TMA bulk copy
    int wP, wRP, wCS, wH, wBulk;
};

// ------------------------------------------------------------------------------ traits
template <class In>
struct FixedTr {
    using in_t = In;
    using val_t = i64;
    using acc_t = u64;
    static constexpr bool fixed = true;
    __device__ static __forceinline__ in_t sentinel() { return std::numeric_limits<In>::max(); }
    __device__ static __forceinline__ val_t half_sum(val_t a, val_t b) { return (a + b) / 2; }  // tdiv
    __device__ static __forceinline__ val_t mean(val_t s, int n) { return s / (val_t)n; }      // tdiv
    __device__ static __forceinline__ acc_t dev(in_t x, val_t c) { return wmul_sq((i64)x - c); }
};
template <class F>
struct FloatTr {
    using in_t = F;
    using val_t = F;
    using acc_t = F;
    static constexpr bool fixed = false;
    __device__ static __forceinline__ in_t sentinel() { return std::numeric_limits<F>::infinity(); }
    __device__ static __forceinline__ val_t half_sum(val_t a, val_t b) { return (F)0.5 * (a + b); }
    __device__ static __forceinline__ val_t mean(val_t s, int n) { return s / (F)n; }
    __device__ static __forceinline__ acc_t dev(in_t x, val_t c) { F d = x - c; return d * d; }
};

// order statistics that make up a centre of n values (math.cairo:113-126 / np.median)
__device__ __forceinline__ void centre_ranks(int kind, int n, int &klo, int &khi) {
    khi = n / 2;
    klo = (kind == SVOC_CENTRE_SMOOTH_MEDIAN_CAIRO) ? n / 2 - 1 : (n - 1) / 2;
    if (klo < 0) klo = 0;  // n < 2 is reported through the status word
}

// The second pass panics on its shape when its centre cannot be formed from R rows: the contract's smooth median
// reads sorted[mid - 1] and sorted[mid] (math.cairo:117: at least two rows), np.median and the mean take one row
// (ipynb beta :369-371 restricted_median over a single kept oracle is valid).  With the contract's own settings
// (constrained -> smooth median, unconstrained -> mean) this is "R < 2 constrained, R < 1 unconstrained".
__device__ __forceinline__ bool second_pass_shape_panic(const Params &p) {
    return p.centre2 == SVOC_CENTRE_SMOOTH_MEDIAN_CAIRO ? p.R < 2 : p.R < 1;
}

// ------------------------------------------------------------------------------ moments
// skewness / kurtosis of one column over the R reliable rows (math.cairo:208-222, 320-363).
// `fe(f)` calls f(x) for every reliable value.
template <class Tr, class ForEach>
__device__ __forceinline__ void column_moments(ForEach fe, int R, typename Tr::val_t &skew,
                                               typename Tr::val_t &kurt, int &stw) {
    if constexpr (Tr::fixed) {
        skew = 0;
        kurt = 0;
        i64 s = 0;
        fe([&](typename Tr::in_t x) { s += (i64)x; });
        const i64 mean = s / (i64)R;                      // math.cairo:240-254
        u64 v = 0;
        fe([&](typename Tr::in_t x) { v += wmul_sq((i64)x - mean); });
        const u64 var = v / (u64)R;                       // math.cairo:208-222
        u64 sd;
        if (!wsqrt(var, sd) || sd == 0 || R <= 2) {       // sqrt(1), wsad_div(., 0), (R-1)(R-2) == 0
            stw = min(stw, STW_SKEW);
            return;
        }
        i128 s3 = 0;
        u128 s4 = 0;
        fe([&](typename Tr::in_t x) {
            const i64 z = wdiv_wide((i64)x - mean, (i64)sd);                        // math.cairo:331
            const u64 az = (u64)(z < 0 ? -z : z);
            const u64 z2 = (u64)(((u128)az * (u128)az + (u128)HALF_WSAD) / (u128)WSAD);
            s3 += ((i128)z2 * (i128)z + (i128)HALF_WSAD) / (i128)WSAD;              // math.cairo:332
            s4 += ((u128)z2 * (u128)z2 + (u128)HALF_WSAD) / (u128)WSAD;            // math.cairo:353-355
        });
        const i128 n = R;
        skew = (i64)((s3 * n) / ((n - 1) * (n - 2)));                               // math.cairo:337
        if (R <= 3) {
            stw = min(stw, STW_KURT);
            return;
        }
        const i128 t1 = ((i128)s4 * n * (n + 1)) / (n - 1);                         // math.cairo:359
        const i128 t2 = (i128)3 * WSAD * (n - 1) * (n - 1);
        kurt = (i64)((t1 - t2) / ((n - 2) * (n - 3)));
    } else {
        using F = typename Tr::val_t;
        F s = 0;
        fe([&](F x) { s += x; });
        const F r = (F)R, mean = s / r;
        F v = 0;
        fe([&](F x) { F d = x - mean; v += d * d; });
        const F sd = sqrt(v / r);
        F s3 = 0, s4 = 0;
        fe([&](F x) { F z = (x - mean) / sd; F z2 = z * z; s3 += z2 * z; s4 += z2 * z2; });
        skew = s3 * r / ((r - 1) * (r - 2));
        kurt = (s4 * r * (r + 1) / (r - 1) - (F)3 * (r - 1) * (r - 1)) / ((r - 2) * (r - 3));
    }
}

// ------------------------------------------------------------------------------ column centre
// small<NPAD>: the column lives in registers.  Returns the centre of the rows whose keep flag
// is set (all rows when keep == nullptr); optionally the moments of the same values.
template <class Tr, int NPAD>
__device__ __forceinline__ typename Tr::val_t column_centre_net(const typename Tr::in_t *col, int stride, int N,
                                                                const uint8_t *keep, int n_valid, int kind,
                                                                bool moments, typename Tr::val_t &skew,
                                                                typename Tr::val_t &kurt, int &stw) {
    using in_t = typename Tr::in_t;
    using val_t = typename Tr::val_t;
    in_t a[NPAD];
    if (NPAD >= 16 && __isShared(col)) {
        // the staged tile (the usual case): rows through the shared-memory window, lds_at.  Rows past N re-read
        // row 0 (an address inside the tile) and become sentinels.  (8-row columns: measured slower, left alone.)
        const uint32_t base = smem_u32(col), step = (uint32_t)stride * (uint32_t)sizeof(in_t);
        constexpr int SURE = NPAD / 2 + 1;     // rows every N of this NPAD has (npad_for: NPAD / 2 < N)
#pragma unroll
        for (int i = 0; i < NPAD; ++i) {
            bool ok = i < SURE || i < N;
            in_t v;
            lds_at(v, base + (ok ? (uint32_t)i : 0u) * step);
            if (ok && keep) ok = keep[i] != 0;
            a[i] = ok ? v : Tr::sentinel();
        }
    } else {
#pragma unroll
        for (int i = 0; i < NPAD; ++i) {
            bool ok = i < N;
            if (ok && keep) ok = keep[i] != 0;
            a[i] = ok ? col[(size_t)i * stride] : Tr::sentinel();
        }
    }
    val_t c;
    if (kind == SVOC_CENTRE_MEAN) {
        val_t s = 0;
#pragma unroll
        for (int i = 0; i < NPAD; ++i) {
            bool ok = i < N;
            if (ok && keep) ok = keep[i] != 0;
            if (ok) s += (val_t)a[i];
        }
        c = Tr::mean(s, n_valid);
    } else {
        network_sort<in_t, NPAD>(a);
        int klo, khi;
        centre_ranks(kind, n_valid, klo, khi);
        c = Tr::half_sum((val_t)pick<in_t, NPAD>(a, klo), (val_t)pick<in_t, NPAD>(a, khi));
    }
    if (moments) {
        // moments are order-free: walk the staged rows again so the register array is never
        // indexed dynamically (that would demote it to local memory)
        auto fe = [&](auto f) {
            for (int i = 0; i < N; ++i)
                if (!keep || keep[i]) f(col[(size_t)i * stride]);
        };
        column_moments<Tr>(fe, n_valid, skew, kurt, stw);
    }
    return c;
}

// generic: k-th smallest (0-based) key among the flagged rows by MSB-first bisection.
template <class Tr>
__device__ __forceinline__ auto kth_key(const typename Tr::in_t *col, int stride, int N, const uint8_t *keep, int k)
    -> decltype(okey(typename Tr::in_t())) {
    using key_t = decltype(okey(typename Tr::in_t()));
    key_t prefix = 0;
    for (int bit = (int)sizeof(key_t) * 8 - 1; bit >= 0; --bit) {
        const key_t cand = prefix | ((key_t)1 << bit);
        int cnt = 0;
        for (int i = 0; i < N; ++i) {
            const bool ok = keep ? keep[i] != 0 : true;
            cnt += (ok && okey(col[(size_t)i * stride]) < cand) ? 1 : 0;
        }
        if (cnt <= k) prefix = cand;  // the k-th key is >= cand
    }
    return prefix;
}

template <class Tr>
__device__ __forceinline__ typename Tr::val_t column_centre_generic(const typename Tr::in_t *col, int stride, int N,
                                                                    const uint8_t *keep, int n_valid, int kind,
                                                                    bool moments, typename Tr::val_t &skew,
                                                                    typename Tr::val_t &kurt, int &stw) {
    using in_t = typename Tr::in_t;
    using val_t = typename Tr::val_t;
    using key_t = decltype(okey(in_t()));
    val_t c;
    if (kind == SVOC_CENTRE_MEAN) {
        val_t s = 0;
        for (int i = 0; i < N; ++i)
            if (!keep || keep[i]) s += (val_t)col[(size_t)i * stride];
        c = Tr::mean(s, n_valid);
    } else {
        int klo, khi;
        centre_ranks(kind, n_valid, klo, khi);
        const key_t k1 = kth_key<Tr>(col, stride, N, keep, klo);
        key_t k2 = k1;
        if (khi != klo) {
            // next order statistic: k1 again when duplicated, else the smallest key above k1
            int cnt_le = 0;
            key_t nxt = ~(key_t)0;
            for (int i = 0; i < N; ++i) {
                if (keep && !keep[i]) continue;
                const key_t kk = okey(col[(size_t)i * stride]);
                cnt_le += kk <= k1 ? 1 : 0;
                if (kk > k1 && kk < nxt) nxt = kk;
            }
            k2 = (cnt_le >= khi + 1) ? k1 : nxt;
        }
        in_t lo, hi;
        unkey(k1, lo);
        unkey(k2, hi);
        c = Tr::half_sum((val_t)lo, (val_t)hi);
    }
    if (moments) {
        auto fe = [&](auto f) {
            for (int i = 0; i < N; ++i)
                if (!keep || keep[i]) f(col[(size_t)i * stride]);
        };
        column_moments<Tr>(fe, n_valid, skew, kurt, stw);
    }
    return c;
}

// ------------------------------------------------------------------------------ reliability
// contract.cairo:436-439 (constrained) / :365-368 (unconstrained), fixed form.
__device__ inline i64 reliability_fixed(u64 avg_qr, const Params &p, int stw_sqrt, int &stw) {
    u64 s;
    if (p.constrained) {
        // the obsolete contract skips the division by the dimension (contract/obsolete/src/contract_nd.cairo:437)
        const u64 v = (p.flags & SVOC_FLAG_OBSOLETE_RELIABILITY) ? avg_qr : avg_qr / (u64)p.D;
        if (!wsqrt(v, s)) { stw = min(stw, stw_sqrt); return 0; }
        return WSAD - 2 * (i64)s;
    }
    if (!wsqrt(avg_qr, s)) { stw = min(stw, stw_sqrt); return 0; }
    if (p.ms_w == 0) { stw = min(stw, STW_SPREAD_ZERO); return 0; }
    const i64 m = (p.ms_w > (i64)s) ? (i64)s : p.ms_w;  // math.cairo:312-318
    return WSAD - wdiv_wide(m, p.ms_w);
}
template <class F>
__device__ inline F reliability_float(F avg_qr, const Params &p) {
    if (p.constrained)
        return (F)1 - (F)2 * sqrt((p.flags & SVOC_FLAG_OBSOLETE_RELIABILITY) ? avg_qr : avg_qr / (F)p.D);
    const F s = sqrt(avg_qr), ms = (F)p.ms_f;
    return (F)1 - (ms > s ? s : ms) / ms;
}

// ------------------------------------------------------------------------------ one block of rounds
// Phases 1-4 for the `nr` rounds staged at `tile` (row-major [nr][N][D]).  Outputs of round r0 + r go
// to index (r0 - ob + r) of the output arrays, so a caller may point them at a per-block scratch.
// Rank of element i among the N unsigned keys qb[stride * j] under (key asc, index desc | asc): the number of
// elements that come before it.  One compare per pair: "equal, but ahead by the index rule" is folded into the
// threshold — element j comes before element i iff key_j < key_i + [j wins a tie against i].  Non-negative
// floats, doubles and int64 order like their bit patterns, so the risks are passed as raw words.
// Both the threshold and the count ride on the carry flag (IADD3 with carry-out, IADD3.X with carry-in): four
// integer instructions and a load per pair.  After sub.cc the flag is the carry of x + ~y + 1, i.e. [x >= y]
// (NOT the borrow), so the loop counts the elements that do NOT come before element i.
template <class UT, int stride>
__device__ __forceinline__ int rank_of(const UT *qb, int N, int i, bool high) {
    const UT mine = qb[stride * i];
    int ge = 0;
    auto step = [&](UT v, int x, int y) {                  // ge += v >= mine + [x >= y]
        if constexpr (sizeof(UT) == 4) {
            asm("{\n\t.reg .u32 t, thr;\n\tsub.cc.u32 t, %2, %3;\n\taddc.u32 thr, %4, 0;\n\t"
                "sub.cc.u32 t, %1, thr;\n\taddc.u32 %0, %0, 0;\n\t}"
                : "+r"(ge) : "r"((uint32_t)v), "r"(x), "r"(y), "r"((uint32_t)mine));
        } else {
            asm("{\n\t.reg .u32 t, tl, th;\n\tsub.cc.u32 t, %3, %4;\n\taddc.cc.u32 tl, %5, 0;\n\taddc.u32 th, %6, 0;\n\t"
                "sub.cc.u32 t, %1, tl;\n\tsubc.cc.u32 t, %2, th;\n\taddc.u32 %0, %0, 0;\n\t}"
                : "+r"(ge)
                : "r"((uint32_t)v), "r"((uint32_t)((u64)v >> 32)), "r"(x), "r"(y), "r"((uint32_t)mine),
                  "r"((uint32_t)((u64)mine >> 32)));
        }
    };
    if (high) {                                            // the tie goes to the higher index: j >= i + 1
#pragma unroll 4
        for (int j = 0; j < N; ++j) step(qb[stride * j], j, i + 1);
    } else {                                               // ... to the lower index: i >= j + 1
#pragma unroll 4
        for (int j = 0; j < N; ++j) step(qb[stride * j], i, j + 1);
    }
    return N - ge;
}

template <class Tr, int NPAD>
__device__ __forceinline__ void consensus_block(const Params &p, const typename Tr::in_t *tile, i64 r0, int nr, i64 ob,
                                                typename Tr::val_t *c1, typename Tr::acc_t *qr_s, int *stw_s,
                                                uint8_t *keep_s) {
    using in_t = typename Tr::in_t;
    using val_t = typename Tr::val_t;
    using acc_t = typename Tr::acc_t;
    const int N = p.N, D = p.D, R = p.R;
    const int tid = threadIdx.x, nt = blockDim.x;
    const size_t round_elems = (size_t)N * D;

    // ---- phase 1: first-pass centre of every (round, dimension) column
    for (int c = tid; c < nr * D; c += nt) {
        const int r = c / D, d = c - r * D;
        const in_t *col = tile + (size_t)r * round_elems + d;
        val_t sk, ku;
        int stw = STW_OK;
        val_t e1;
        if constexpr (NPAD > 0)
            e1 = column_centre_net<Tr, NPAD>(col, D, N, nullptr, N, p.centre, false, sk, ku, stw);
        else
            e1 = column_centre_generic<Tr>(col, D, N, nullptr, N, p.centre, false, sk, ku, stw);
        c1[c] = e1;
        if (p.essence1) reinterpret_cast<val_t *>(p.essence1)[(size_t)(r0 - ob) * D + c] = e1;
    }
    __syncthreads();

    // ---- phase 2: per-oracle quadratic risk against the first-pass centre (math.cairo:225-238)
    for (int w = tid; w < nr * N; w += nt) {
        const int r = w / N, i = w - r * N;
        const in_t *row = tile + (size_t)r * round_elems + (size_t)i * D;
        const val_t *cc = c1 + (size_t)r * D;
        acc_t q = 0;
        bool bad = false;
        if constexpr (Tr::fixed) {
            int d = i % D;  // rotated start: conflict-free shared-memory reads across rows (integer sums are order-free)
            for (int k = 0; k < D; ++k) {
                const in_t x = row[d];
                if (p.constrained) bad |= (x < 0) | ((i64)x > WSAD);          // math.cairo:298-310
                else bad |= ((i64)x > ENGINE_RANGE) | ((i64)x < -ENGINE_RANGE);
                q += Tr::dev(x, cc[d]);
                d = (d + 1 == D) ? 0 : d + 1;
            }
        } else {
            // float form: EVERY row is summed in the same (natural) order, so that identical rows — colluders — get
            // bit-identical risks and a cut inside their group is decided by the tie rule alone, as in the reference
            // (np.linalg.norm per row, ipynb beta :358-360).  16-byte loads where the row allows it keep the
            // shared-memory conflicts of the unrotated walk low.
            constexpr int V = 16 / (int)sizeof(in_t);
            if (D % V == 0 && (reinterpret_cast<uintptr_t>(row) & 15) == 0) {
                for (int d0 = 0; d0 < D; d0 += V) {
                    const int4 raw = *reinterpret_cast<const int4 *>(row + d0);
                    in_t xv[V];
                    memcpy(xv, &raw, 16);
#pragma unroll
                    for (int k = 0; k < V; ++k) q += Tr::dev(xv[k], cc[d0 + k]);
                }
            } else {
                for (int d = 0; d < D; ++d) q += Tr::dev(row[d], cc[d]);
            }
        }
        qr_s[w] = q;
        if (p.qr) reinterpret_cast<acc_t *>(p.qr)[(size_t)(r0 - ob) * N + w] = q;
        if (bad) atomicMin(&stw_s[r], p.constrained ? STW_INTERVAL : STW_ENGINE_RANGE);
    }
    __syncthreads();

    // ---- phase 3a: reliable flags — rank under (value asc, index desc | asc)  (sort.cairo:96-102,
    //      contract.cairo:345-363)
    {
        const bool high = p.tie == SVOC_TIE_HIGH_INDEX_KEPT;
        // constrained fixed form: every qr is at most D * 1e6, so when that fits 32 bits the low words decide
        bool low32 = false;
        if constexpr (Tr::fixed) low32 = p.constrained && (u64)D * (u64)WSAD < 0xffffffffull;
        for (int w = tid; w < nr * N; w += nt) {
            const int r = w / N, i = w - r * N;
            const acc_t *q = qr_s + (size_t)r * N;
            int rank;
            if (low32) rank = rank_of<uint32_t, 2>(reinterpret_cast<const uint32_t *>(q), N, i, high);
            else if constexpr (sizeof(acc_t) == 4) rank = rank_of<uint32_t, 1>(reinterpret_cast<const uint32_t *>(q), N, i, high);
            else rank = rank_of<u64, 1>(reinterpret_cast<const u64 *>(q), N, i, high);
            keep_s[w] = rank < R ? 1 : 0;
            if (p.rank) p.rank[(size_t)(r0 - ob) * N + w] = rank;
        }
    }
    __syncthreads();

    // ---- phase 3b: reliabilities and the packed reliable mask.  Two scalar tasks per round (first- and
    //      second-pass reliability: a sum and a Newton square root each), handed to threads spread over
    //      the warps so that no single warp carries all the square roots.
    {
        const int ntask = 2 * nr;
        const int spread = max(1, nt / max(1, ntask));     // > 1 when there are fewer tasks than threads
        for (int task = (tid % spread == 0) ? tid / spread : ntask; task < ntask; task += nt) {
            const int r = task >> 1, second = task & 1;
            const acc_t *q = qr_s + (size_t)r * N;
            const uint8_t *kp = keep_s + (size_t)r * N;
            acc_t sum = 0;
            if (second) { for (int i = 0; i < N; ++i) if (kp[i]) sum += q[i]; }
            else { for (int i = 0; i < N; ++i) sum += q[i]; }
            int stw = STW_OK;
            const bool shape2 = second_pass_shape_panic(p);
            if (!second && N < 2) stw = min(stw, STW_FIRST_SHAPE);
            if (second && shape2) stw = min(stw, STW_SECOND_SHAPE);
            const int cnt = second ? R : N;
            if constexpr (Tr::fixed) {
                i64 rel = 0;
                if (!(second && shape2)) {
                    rel = reliability_fixed(sum / (u64)cnt, p, second ? STW_SQRT_REL2 : STW_SQRT_REL1, stw);
                    if (rel < 0 || rel > WSAD) stw = min(stw, second ? STW_REL2_RANGE : STW_REL1_RANGE);
                }
                reinterpret_cast<i64 *>(second ? p.rel2 : p.rel1)[r0 - ob + r] = rel;
            } else {
                val_t rel = 0;
                if (!(second && shape2)) {
                    rel = reliability_float<val_t>(sum / (val_t)cnt, p);
                    if (!(rel >= 0 && rel <= 1)) stw = min(stw, second ? STW_REL2_RANGE : STW_REL1_RANGE);
                }
                reinterpret_cast<val_t *>(second ? p.rel2 : p.rel1)[r0 - ob + r] = rel;
            }
            if (stw != STW_OK) atomicMin(&stw_s[r], stw);
        }
    }
    if (p.reliable) {
        for (int w = tid; w < nr * p.words; w += nt) {
            const int r = w / p.words, k = w - r * p.words;
            const uint8_t *kp = keep_s + (size_t)r * N;
            uint32_t bits = 0;
            for (int b = 0; b < 32 && k * 32 + b < N; ++b) bits |= (uint32_t)(kp[k * 32 + b] != 0) << b;
            p.reliable[(size_t)(r0 - ob) * p.words + w] = bits;
        }
    }

    // ---- phase 4: second-pass centre over the reliable rows (+ moments)
    const bool shape2 = second_pass_shape_panic(p);
    for (int c = tid; c < nr * D; c += nt) {
        const int r = c / D, d = c - r * D;
        const in_t *col = tile + (size_t)r * round_elems + d;
        const uint8_t *kp = keep_s + (size_t)r * N;
        val_t sk = 0, ku = 0, cons = 0;
        int stw = STW_OK;
        if (!shape2) {
            if constexpr (NPAD > 0)
                cons = column_centre_net<Tr, NPAD>(col, D, N, kp, R, p.centre2, p.want_moments != 0, sk, ku, stw);
            else
                cons = column_centre_generic<Tr>(col, D, N, kp, R, p.centre2, p.want_moments != 0, sk, ku, stw);
        }
        reinterpret_cast<val_t *>(p.consensus)[(size_t)(r0 - ob) * D + c] = cons;
        if (p.want_moments) {
            if (p.skew) reinterpret_cast<val_t *>(p.skew)[(size_t)(r0 - ob) * D + c] = sk;
            if (p.kurt) reinterpret_cast<val_t *>(p.kurt)[(size_t)(r0 - ob) * D + c] = ku;
        }
        if (stw != STW_OK) atomicMin(&stw_s[r], stw);
    }
    __syncthreads();
    for (int r = tid; r < nr; r += nt) {
        const int w = stw_s[r];
        p.status[r0 - ob + r] = (w == STW_OK) ? SVOC_ST_OK : (w & 0xff);
    }
}

// ------------------------------------------------------------------------------ the kernel
// Resident CTAs per SM the compiler must make room for: three for the 64-row networks (168 registers; four, 128
// registers with some spills, for int32 fixed point: +7 % on config 3), eight for
// the fp32 32-row kernel (64 registers instead of 113: +20 % on the config-5 shape — latency hiding beats the
// few extra moves), no constraint elsewhere (64-bit arithmetic spills under a tight register cap).
template <class Tr, int NPAD>
constexpr int small_min_ctas() {
    return NPAD == 64 ? ((Tr::fixed && sizeof(typename Tr::in_t) == 4) ? 4 : 3) : ((NPAD == 32 && !Tr::fixed && sizeof(typename Tr::acc_t) == 4) ? 8 : 1);
}
template <class Tr, int NPAD>
__global__ void __launch_bounds__(NPAD == 0 ? 256 : 128, small_min_ctas<Tr, NPAD>()) consensus_kernel(const Params p) {
    using in_t = typename Tr::in_t;
    using val_t = typename Tr::val_t;
    using acc_t = typename Tr::acc_t;
    extern __shared__ __align__(128) unsigned char smem_raw[];

    const int N = p.N, D = p.D, R = p.R;
    const i64 r0 = (i64)blockIdx.x * p.rb;
    const int nr = (int)min((i64)p.rb, p.T - r0);  // rounds in this CTA
    const int tid = threadIdx.x, nt = blockDim.x;
    const size_t round_elems = (size_t)N * D;

    // shared-memory carve-up
    size_t off = 0;
    in_t *tile_s = reinterpret_cast<in_t *>(smem_raw);
    if (p.tile_smem) off += ((size_t)p.rb * round_elems * sizeof(in_t) + 15) & ~(size_t)15;
    val_t *c1 = reinterpret_cast<val_t *>(smem_raw + off);
    off += ((size_t)p.rb * D * sizeof(val_t) + 15) & ~(size_t)15;
    acc_t *qr_s = reinterpret_cast<acc_t *>(smem_raw + off);
    off += ((size_t)p.rb * N * sizeof(acc_t) + 15) & ~(size_t)15;
    int *stw_s = reinterpret_cast<int *>(smem_raw + off);
    off += ((size_t)p.rb * sizeof(int) + 15) & ~(size_t)15;
    u64 *bar = reinterpret_cast<u64 *>(smem_raw + off);
    off += 16;
    uint8_t *keep_s = smem_raw + off;

    const in_t *gsrc = reinterpret_cast<const in_t *>(p.pred) + (size_t)r0 * round_elems;
    const in_t *tile = p.tile_smem ? tile_s : gsrc;

    // ---- phase 0: stage the rounds (one TMA bulk copy when 16-byte aligned)
    for (int r = tid; r < p.rb; r += nt) stw_s[r] = STW_OK;
    if (p.tile_smem) {
        const size_t bytes = (size_t)nr * round_elems * sizeof(in_t);
        const bool bulk = ((reinterpret_cast<uintptr_t>(gsrc) & 15) == 0) && ((bytes & 15) == 0) && bytes < (1u << 20);
        if (bulk) {
            if (tid == 0) {
                mbar_init(bar, 1);
                mbar_expect_tx(bar, (uint32_t)bytes);
                bulk_g2s(tile_s, gsrc, (uint32_t)bytes, bar);
            }
            __syncthreads();
            mbar_wait(bar, 0);
        } else {
            for (size_t e = tid; e < (size_t)nr * round_elems; e += nt) tile_s[e] = gsrc[e];
        }
    }
    __syncthreads();

    consensus_block<Tr, NPAD>(p, tile, r0, nr, 0, c1, qr_s, stw_s, keep_s);
}

// ------------------------------------------------------------------------------ launch
struct Plan {
    int npad, rb, threads, tile_smem;
    size_t smem;
    int variant;  // 0: small<NPAD> / generic, 1: wide
    long long grid;
    int wP, wRP, wCS, wH;
};

int fail(const std::string &m);
void count_launch();

template <class Tr, int NPAD>
static int launch(const Params &p, const Plan &pl, cudaStream_t st) {
    auto kern = consensus_kernel<Tr, NPAD>;
    if (pl.smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl.smem);
        if (e != cudaSuccess) return fail(std::string("cudaFuncSetAttribute: ") + cudaGetErrorString(e));
    }
    const i64 grid = (p.T + pl.rb - 1) / pl.rb;
    if (grid > 0x7fffffffLL) return fail("too many rounds for one launch");
    kern<<<(unsigned)grid, pl.threads, pl.smem, st>>>(p);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(std::string("kernel launch: ") + cudaGetErrorString(e));
    count_launch();
    return 0;
}

template <class Tr>
int dispatch_npad(const Params &p, const Plan &pl, cudaStream_t st) {
    switch (pl.npad) {
        case 8: return launch<Tr, 8>(p, pl, st);
        case 16: return launch<Tr, 16>(p, pl, st);
        case 32: return launch<Tr, 32>(p, pl, st);
        case 64: return launch<Tr, 64>(p, pl, st);
        default: return launch<Tr, 0>(p, pl, st);
    }
}

// one translation unit per element type (consensus_<dtype>.cu) instantiates these
int consensus_i32(const Params &p, const Plan &pl, cudaStream_t st);
int consensus_i64(const Params &p, const Plan &pl, cudaStream_t st);
int consensus_f32(const Params &p, const Plan &pl, cudaStream_t st);
int consensus_f64(const Params &p, const Plan &pl, cudaStream_t st);

}  // namespace svoc
