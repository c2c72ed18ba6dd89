// consensus_col.cuh — the `column` variant of the fused consensus round for 64 < N <= 512 oracles and
// 32-bit elements (fp32 / int32): the kernel config 4 (N=256, D=128, fp32) runs.
//
// Work decomposition (chosen after profiling the earlier `wide` variant, where partial-count exchanges,
// shared-memory atomics and ~40 barriers per round cost more than the arithmetic):
//   * one THREAD owns one whole column (all N rows of one dimension): selection needs no cross-thread
//     communication at all — no partial sums, no atomics, no barriers;
//   * one GROUP of G = ceil32(D) threads owns one round; a CTA holds several groups that run
//     independent rounds and synchronise only inside their group (named barriers), so one round's
//     latency-bound stretches are filled by the others;
//   * the round is never staged: its three sweeps (min/max, keys, row divergences) read global memory
//     with the lanes of a warp on consecutive columns (128-byte coalesced); the first sweep comes from
//     HBM, the other two from L2, and the group's next round is prefetched into L2
//     (cp.async.bulk.prefetch.L2).  Shared memory only holds the packed keys and small per-round state.
// Selection per column (same idea as `wide`): exact min/max -> monotone 7-bit keys packed 4 per word in
// a transposed, bank-conflict-free scratch; MSB-first bisection with SWAR counting (one add + one
// signed IDP.4A per four keys); the rows of the target bucket (and the next two when needed) are
// collected and ranked exactly; windows that are too crowded are re-keyed over their own value range.
// Reference semantics: see consensus_kernel.cuh (contract.cairo:442-503 etc.).
#pragma once
#include "consensus_wide.cuh"

namespace svoc {

constexpr int COL_THREADS = 512;
constexpr int COL_CAP = 32;   // candidates ranked in place; larger windows are re-keyed

struct ColCtx {
    int N, NP, D, LD;      // rows, rows padded to 32, dimensions, leading dimension of the candidate lists
    int c;                 // this thread's column
    uint8_t *keys;         // this column's packed keys (NP bytes, 16-byte aligned)
};

template <class Tr, bool FIRST, int DT, int NPT>
__device__ __forceinline__ void col_build_keys(const ColCtx &w, const typename Tr::in_t *col, const uint8_t *keep,
                                               typename Tr::in_t lo, typename Tr::in_t hi,
                                               const typename WideKey<Tr>::Scale &sc) {
    const int WD = DT ? DT : w.D;
    const int NP = NPT ? NPT : w.NP;
    const typename Tr::in_t *ptr = col;
#pragma unroll 2
    for (int g = 0; g < NP; g += 16) {
        unsigned f[16];
        if (!keep && g + 16 <= w.N) {
#pragma unroll
            for (int k = 0; k < 16; ++k) f[k] = wide_key<Tr, FIRST>(ptr[k * WD], lo, hi, sc);
        } else {
#pragma unroll
            for (int k = 0; k < 16; ++k) {
                const int row = g + k;
                f[k] = 127u;
                if (row < w.N && (!keep || keep[row])) f[k] = wide_key<Tr, FIRST>(ptr[k * WD], lo, hi, sc);
            }
        }
        *reinterpret_cast<uint4 *>(w.keys + g) =
            make_uint4(pack4(f[0], f[1], f[2], f[3]), pack4(f[4], f[5], f[6], f[7]), pack4(f[8], f[9], f[10], f[11]),
                       pack4(f[12], f[13], f[14], f[15]));
        ptr += 16 * WD;
    }
}

// Centre of the flagged rows of this thread's column (all rows when keep == nullptr).  Entirely
// thread-local.  `refined` tells the caller that the level-0 keys were overwritten.
template <class Tr, int DT, int NPT>
__device__ __forceinline__ typename Tr::val_t col_centre(const ColCtx &w, const typename Tr::in_t *col,
                                                         const uint8_t *keep, int n_valid, int kind, bool col_const,
                                                         bool col_fallback, typename Tr::in_t const_value,
                                                         typename Tr::in_t *list, bool &refined) {
    using in_t = typename Tr::in_t;
    using val_t = typename Tr::val_t;
    using Key = WideKey<Tr>;
    const int WD = DT ? DT : w.D;
    const int NP = NPT ? NPT : w.NP;
    const int LD = DT ? ((DT + 31) & ~31) : w.LD;
    refined = false;
    if (kind == SVOC_CENTRE_MEAN) {
        val_t s = 0;
        for (int i = 0; i < w.N; ++i)
            if (!keep || keep[i]) s += (val_t)col[i * WD];
        return Tr::mean(s, n_valid);
    }
    if (col_const) return Tr::half_sum((val_t)const_value, (val_t)const_value);
    if (col_fallback) {
        val_t sk, ku;
        int stw = STW_OK;
        return column_centre_generic<Tr>(col, WD, w.N, keep, n_valid, kind, false, sk, ku, stw);
    }
    int klo, khi;
    centre_ranks(kind, n_valid, klo, khi);
    in_t *lc = list + w.c;
    in_t v_lo = const_value, v_hi = const_value;

    for (;;) {
        // ---- bisection over the 7 key bits for rank klo
        unsigned prefix = 0;
        int L = 0, LE = n_valid;
        const int sum_keys = key_byte_sum<NPT>(w.keys, NP);
#pragma unroll
        for (int bit = 6; bit >= 0; --bit) {
            const unsigned cand = prefix | (1u << bit);
            const int lt = NP - (int)count_ge<NPT>(w.keys, NP, cand, sum_keys);   // sentinels are never below a pivot
            if (lt <= klo) { prefix = cand; L = lt; }
            else LE = lt;   // the last rejected pivot is prefix + 1, so LE ends as #(key <= prefix)
        }
        // ---- collect the window: the bucket of rank klo alone when rank khi is inside it, else three buckets
        const unsigned span = (khi < LE) ? 1u : 3u;
        const unsigned up = prefix + span > 127u ? 127u : prefix + span;
        const unsigned c1b = (128u - prefix) * 0x01010101u, c2b = (128u - up) * 0x01010101u;
        uint32_t kmin = 0xffffffffu, kmax = 0u;
        int m = 0;
        // First the row indices of the window (no loads), then their values with four loads in flight.
        // Row index k lives in the upper half of this column's list (two per 32-bit slot, slots 16..31),
        // value k in slot k: a value only ever overwrites row indices that were already consumed.
        auto row_slot = [&](int k) -> unsigned short & {
            return reinterpret_cast<unsigned short *>(&lc[(COL_CAP / 2 + (k >> 1)) * LD])[k & 1];
        };
        auto window_flags = [&](int g) {
            const uint4 a = *reinterpret_cast<const uint4 *>(w.keys + g);
            const uint4 b = *reinterpret_cast<const uint4 *>(w.keys + g + 16);
            unsigned acc = (a.x + c1b) & ~(a.x + c2b) & 0x80808080u;
            acc |= ((a.y + c1b) & ~(a.y + c2b) & 0x80808080u) >> 1;
            acc |= ((a.z + c1b) & ~(a.z + c2b) & 0x80808080u) >> 2;
            acc |= ((a.w + c1b) & ~(a.w + c2b) & 0x80808080u) >> 3;
            acc |= ((b.x + c1b) & ~(b.x + c2b) & 0x80808080u) >> 4;
            acc |= ((b.y + c1b) & ~(b.y + c2b) & 0x80808080u) >> 5;
            acc |= ((b.z + c1b) & ~(b.z + c2b) & 0x80808080u) >> 6;
            acc |= ((b.w + c1b) & ~(b.w + c2b) & 0x80808080u) >> 7;
            return acc;
        };
#pragma unroll 2
        for (int g = 0; g < NP; g += 32) {
            unsigned acc = window_flags(g);
            while (acc) {
                const int pos = __ffs(acc) - 1;
                acc &= acc - 1;
                if (m < COL_CAP) row_slot(m) = (unsigned short)(g + 4 * (7 - (pos & 7)) + (pos >> 3));
                ++m;
            }
        }
        if (m <= COL_CAP) {
            int k = 0;
            for (; k + 4 <= m; k += 4) {
                in_t xs[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) xs[u] = col[(int)row_slot(k + u) * WD];
#pragma unroll
                for (int u = 0; u < 4; ++u) lc[(k + u) * LD] = xs[u];
            }
            for (; k < m; ++k) {
                const in_t x = col[(int)row_slot(k) * WD];
                lc[k * LD] = x;
            }
        } else {
            // crowded window (rare): only its value range is needed for the re-keying
            for (int g = 0; g < NP; g += 32) {
                unsigned acc = window_flags(g);
                while (acc) {
                    const int pos = __ffs(acc) - 1;
                    acc &= acc - 1;
                    const uint32_t kx = okey(col[(g + 4 * (7 - (pos & 7)) + (pos >> 3)) * WD]);
                    kmin = kx < kmin ? kx : kmin;
                    kmax = kx > kmax ? kx : kmax;
                }
            }
        }
        const int j1 = klo - L, j2 = khi - L;
        bool rescue = false;
        in_t thr = const_value;
        if (m <= COL_CAP) {
            // lt(v) = #{u < v}; the value of rank j is the candidate with the largest lt <= j
            int b1 = -1, b2 = -1, b3 = -1;
            for (int a = 0; a < m; ++a) {
                const in_t v = lc[a * LD];
                int rk = 0;
#pragma unroll 4
                for (int b = 0; b < m; ++b) rk += lc[b * LD] < v ? 1 : 0;
                if (rk <= j1 && rk > b1) { b1 = rk; v_lo = v; }
                if (rk <= j2 && rk > b2) { b2 = rk; v_hi = v; }
                if (rk > b3) { b3 = rk; thr = v; }
            }
            rescue = j2 >= m;
        } else {
            in_t a, b;
            unkey(kmin, a);
            unkey(kmax, b);
            typename Key::Scale sc;
            if (!(b > a)) {               // the whole window is one value
                v_lo = a;
                v_hi = a;
                thr = a;
                rescue = j2 >= m;
            } else if (!Key::make(a, b, sc)) {
                val_t sk, ku;
                int stw = STW_OK;
                return column_centre_generic<Tr>(col, WD, w.N, keep, n_valid, kind, false, sk, ku, stw);
            } else {                       // finer keys over the window's value range, then try again
                col_build_keys<Tr, false, DT, NPT>(w, col, keep, a, b, sc);
                refined = true;
                continue;
            }
        }
        if (rescue) {                      // rank khi lies above the window: smallest flagged value above it
            uint32_t best = 0xffffffffu;
            for (int row = 0; row < w.N; ++row) {
                if (keep && !keep[row]) continue;
                const in_t x = col[row * WD];
                if (x > thr) { const uint32_t kx = okey(x); best = kx < best ? kx : best; }
            }
            unkey(best, v_hi);
        }
        break;
    }
    return Tr::half_sum((val_t)v_lo, (val_t)v_hi);
}

__device__ __forceinline__ void group_sync(int id, int count) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(count) : "memory");
}

// G = threads per group (a multiple of 32, >= D); groups per CTA = blockDim.x / G.
template <class Tr, int DT, int NPT>
__global__ void __launch_bounds__(COL_THREADS, 1) consensus_col_kernel(const Params p) {
    using in_t = typename Tr::in_t;
    using val_t = typename Tr::val_t;
    using acc_t = typename Tr::acc_t;
    using Key = WideKey<Tr>;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int N = p.N, D = DT ? DT : p.D, R = p.R;
    const int NP = NPT ? NPT : ((N + 31) & ~31);
    const int G = p.wP;                 // threads per group
    const int CS = p.wCS;               // key column stride (bytes)
    const int ngroups = p.wH;           // groups (rounds in flight) per CTA
    const int tid = threadIdx.x;
    const int grp = tid / G, t = tid - grp * G;   // group and thread-in-group
    if (grp >= ngroups) return;         // not a member of any group barrier
    const int bar_id = grp;             // named barrier of the group (__syncthreads is never used here)
    auto al = [](size_t x) { return (x + 15) & ~(size_t)15; };

    // per-group shared memory
    unsigned char *base = smem_raw + (size_t)grp * p.wRP;    // wRP carries the per-group byte count
    size_t off = 0;
    uint8_t *keysT = base + off;                                off += al((size_t)D * CS);
    in_t *list = reinterpret_cast<in_t *>(base + off);          off += al((size_t)COL_CAP * G * sizeof(in_t));
    val_t *c1 = reinterpret_cast<val_t *>(base + off);          off += al((size_t)D * sizeof(val_t));
    acc_t *qr_s = reinterpret_cast<acc_t *>(base + off);        off += al((size_t)N * sizeof(acc_t));
    acc_t *red_s = reinterpret_cast<acc_t *>(base + off);       off += al((size_t)2 * (G / 32) * sizeof(acc_t));
    uint32_t *maskw = reinterpret_cast<uint32_t *>(base + off); off += al((size_t)(NP / 32) * 4);
    int *stw_s = reinterpret_cast<int *>(base + off);           off += 16;
    uint8_t *keep_s = base + off;

    ColCtx w;
    w.N = N; w.NP = NP; w.D = D; w.LD = G;
    w.c = t;
    w.keys = keysT + (size_t)t * CS;
    const bool has_col = t < D;
    const size_t round_elems = (size_t)N * D;
    const uint32_t round_bytes = (uint32_t)(round_elems * sizeof(in_t));
    const bool shape2 = p.constrained ? (R < 2) : (R < 1);
    const i64 stride = (i64)gridDim.x * ngroups;

    for (i64 r = (i64)blockIdx.x * ngroups + grp; r < p.T; r += stride) {
        const in_t *gsrc = reinterpret_cast<const in_t *>(p.pred) + (size_t)r * round_elems;
        if (t == 0) {
            *stw_s = STW_OK;
            if (p.wBulk && p.tile_smem && r + stride < p.T)    // pull this group's next round towards L2 (only when it fits)
                asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(gsrc + (size_t)stride * round_elems), "r"(round_bytes) : "memory");
        }
        const in_t *col = gsrc + t;
        // ---- phase 1: this column's first-pass centre
        bool col_const = false, col_fb = false, refined1 = false, refined2 = false;
        in_t lo = 0, hi = 0;
        typename Key::Scale sc{};
        if (has_col) {
            // Key scale from the first 64 rows only: values outside it clamp to the end buckets, which keeps
            // the keys monotone (a crowded end bucket is re-keyed like any other).  A degenerate sample
            // falls back to the exact sweep, which also decides whether the column is constant.
            lo = std::numeric_limits<in_t>::max();
            hi = std::numeric_limits<in_t>::lowest();
            const int ns = N < 64 ? N : 64;
            {
                const in_t *ptr = col;
#pragma unroll 16
                for (int row = 0; row < ns; ++row, ptr += D) {
                    const in_t x = *ptr;
                    lo = vmin(lo, x);
                    hi = vmax(hi, x);
                }
            }
            if (!(hi > lo)) {
                const in_t *ptr = col + (size_t)ns * D;
#pragma unroll 16
                for (int row = ns; row < N; ++row, ptr += D) {
                    const in_t x = *ptr;
                    lo = vmin(lo, x);
                    hi = vmax(hi, x);
                }
            }
            col_const = !(hi > lo);
            col_fb = !col_const && !Key::make(lo, hi, sc);
            if (!col_const && !col_fb) col_build_keys<Tr, true, DT, NPT>(w, col, nullptr, lo, hi, sc);
            const val_t e1 = col_centre<Tr, DT, NPT>(w, col, nullptr, N, p.centre, col_const, col_fb, lo, list, refined1);
            c1[t] = e1;
            if (p.essence1) reinterpret_cast<val_t *>(p.essence1)[(size_t)r * D + t] = e1;
        }
        group_sync(bar_id, G);

        // ---- phase 2: quadratic risk of the rows this thread owns (t, t + G, ...), 32 bytes per load pair
        for (int i = t; i < N; i += G) {
            const in_t *row = gsrc + (size_t)i * D;
            acc_t q = 0;
            bool bad = false;
            auto one = [&](in_t x, val_t c) {
                if constexpr (Tr::fixed) {
                    if (p.constrained) bad |= (x < 0) | ((i64)x > WSAD);
                    else bad |= ((i64)x > ENGINE_RANGE) | ((i64)x < -ENGINE_RANGE);
                }
                q += Tr::dev(x, c);
            };
            if ((D & 7) == 0 && p.wBulk) {
#pragma unroll 4
                for (int d = 0; d < D; d += 8) {
                    const int4 xa = *reinterpret_cast<const int4 *>(row + d);
                    const int4 xb = *reinterpret_cast<const int4 *>(row + d + 4);
                    in_t x[8];
                    memcpy(&x[0], &xa, 16);
                    memcpy(&x[4], &xb, 16);
                    val_t cc[8];
                    if constexpr (sizeof(val_t) == 4) {
                        const int4 ca = *reinterpret_cast<const int4 *>(c1 + d), cb = *reinterpret_cast<const int4 *>(c1 + d + 4);
                        memcpy(&cc[0], &ca, 16);
                        memcpy(&cc[4], &cb, 16);
                    } else {
#pragma unroll
                        for (int k = 0; k < 8; k += 2) {
                            const int4 cv = *reinterpret_cast<const int4 *>(c1 + d + k);
                            memcpy(&cc[k], &cv, 16);
                        }
                    }
#pragma unroll
                    for (int k = 0; k < 8; ++k) one(x[k], cc[k]);
                }
            } else {
                for (int d = 0; d < D; ++d) one(row[d], c1[d]);
            }
            qr_s[i] = q;
            if (p.qr) reinterpret_cast<acc_t *>(p.qr)[(size_t)r * N + i] = q;
            if (bad) atomicMin(stw_s, p.constrained ? STW_INTERVAL : STW_ENGINE_RANGE);
        }
        group_sync(bar_id, G);

        // ---- phase 3: rank of this thread's rows under (qr asc, index desc | asc); reliable <=> rank < R.
        //      The rows of one warp are consecutive, so every oracle below the warp's first row (above its
        //      last row) needs one compare, strict or not by the tie rule; only the warp's own 32 rows
        //      need the index test.  qr is read with 128-bit broadcast loads.
        {
            const bool high = p.tie == SVOC_TIE_HIGH_INDEX_KEPT;
            for (int i0 = t & ~31; i0 < N; i0 += G) {
                const int i = i0 + (t & 31);
                const bool ok = i < N;
                const acc_t mine = ok ? qr_s[i] : qr_s[0];
                const int i1 = min(N, i0 + 32);
                int rank = 0;
                auto count = [&](int j0, int j1, bool strict) {
                    int c = 0, j = j0;
                    if constexpr (sizeof(acc_t) == 4) {
                        for (; j < j1 && (j & 3); ++j) c += (strict ? qr_s[j] < mine : qr_s[j] <= mine) ? 1 : 0;
                        if (strict) {
                            for (; j + 4 <= j1; j += 4) {
                                const float4 v = *reinterpret_cast<const float4 *>(reinterpret_cast<const float *>(qr_s) + j);
                                c += (v.x < mine ? 1 : 0) + (v.y < mine ? 1 : 0) + (v.z < mine ? 1 : 0) + (v.w < mine ? 1 : 0);
                            }
                        } else {
                            for (; j + 4 <= j1; j += 4) {
                                const float4 v = *reinterpret_cast<const float4 *>(reinterpret_cast<const float *>(qr_s) + j);
                                c += (v.x <= mine ? 1 : 0) + (v.y <= mine ? 1 : 0) + (v.z <= mine ? 1 : 0) + (v.w <= mine ? 1 : 0);
                            }
                        }
                    }
                    for (; j < j1; ++j) c += (strict ? qr_s[j] < mine : qr_s[j] <= mine) ? 1 : 0;
                    return c;
                };
                rank = count(0, i0, high) + count(i1, N, !high);
                {   // the warp's own 32 rows: ties decided by index
                    int lt = 0, eq = 0;
                    if (high) {
                        for (int j = i0; j < i1; ++j) { const acc_t q = qr_s[j]; lt += q < mine ? 1 : 0; eq += (q == mine) & (j > i) ? 1 : 0; }
                    } else {
                        for (int j = i0; j < i1; ++j) { const acc_t q = qr_s[j]; lt += q < mine ? 1 : 0; eq += (q == mine) & (j < i) ? 1 : 0; }
                    }
                    rank += lt + eq;
                }
                if (ok) {
                    keep_s[i] = rank < R ? 1 : 0;
                    if (p.rank) p.rank[(size_t)r * N + i] = rank;
                }
            }
        }
        group_sync(bar_id, G);
        // ---- mask words, reliabilities
        for (int k = t; k < NP / 32; k += G) {
            uint32_t bits = 0;
            for (int b = 0; b < 32 && k * 32 + b < N; ++b) bits |= (uint32_t)(keep_s[k * 32 + b] != 0) << b;
            maskw[k] = bits;
            if (p.reliable && k < p.words) p.reliable[(size_t)r * p.words + k] = bits;
        }
        {
            acc_t s1 = 0, s2 = 0;
            for (int i = t; i < N; i += G) {
                const acc_t q = qr_s[i];
                s1 += q;
                if (keep_s[i]) s2 += q;
            }
            for (int o = 16; o; o >>= 1) {
                s1 += __shfl_xor_sync(0xffffffffu, s1, o);
                s2 += __shfl_xor_sync(0xffffffffu, s2, o);
            }
            if ((t & 31) == 0) { red_s[t >> 5] = s1; red_s[(G >> 5) + (t >> 5)] = s2; }
        }
        group_sync(bar_id, G);
        if (t == 0) {
            acc_t s1 = 0, s2 = 0;
            for (int k = 0; k < (G >> 5); ++k) { s1 += red_s[k]; s2 += red_s[(G >> 5) + k]; }
            int stw = STW_OK;
            if constexpr (Tr::fixed) {
                const i64 rel1 = reliability_fixed(s1 / (u64)N, p, STW_SQRT_REL1, stw);
                if (rel1 < 0 || rel1 > WSAD) stw = min(stw, STW_REL1_RANGE);
                i64 rel2 = 0;
                if (shape2) stw = min(stw, STW_SECOND_SHAPE);
                else {
                    rel2 = reliability_fixed(s2 / (u64)R, p, STW_SQRT_REL2, stw);
                    if (rel2 < 0 || rel2 > WSAD) stw = min(stw, STW_REL2_RANGE);
                }
                reinterpret_cast<i64 *>(p.rel1)[r] = rel1;
                reinterpret_cast<i64 *>(p.rel2)[r] = rel2;
            } else {
                const val_t rel1 = reliability_float<val_t>(s1 / (val_t)N, p);
                if (!(rel1 >= 0 && rel1 <= 1)) stw = min(stw, STW_REL1_RANGE);
                val_t rel2 = 0;
                if (shape2) stw = min(stw, STW_SECOND_SHAPE);
                else {
                    rel2 = reliability_float<val_t>(s2 / (val_t)R, p);
                    if (!(rel2 >= 0 && rel2 <= 1)) stw = min(stw, STW_REL2_RANGE);
                }
                reinterpret_cast<val_t *>(p.rel1)[r] = rel1;
                reinterpret_cast<val_t *>(p.rel2)[r] = rel2;
            }
            if (stw != STW_OK) atomicMin(stw_s, stw);
            if (!p.want_moments) {   // every contribution to the status is in: publish it now
                const int sw = *stw_s;
                p.status[r] = (sw == STW_OK) ? SVOC_ST_OK : (sw & 0xff);
            }
        }
        // ---- phase 4: second-pass centre over the reliable rows
        if (has_col) {
            val_t cons = 0;
            if (!shape2) {
                if (!col_const && !col_fb && p.centre2 != SVOC_CENTRE_MEAN) {
                    if (refined1) {
                        col_build_keys<Tr, true, DT, NPT>(w, col, keep_s, lo, hi, sc);
                    } else {
                        for (int k = 0; k < NP / 32; ++k) {      // dropped rows -> sentinel key
                            unsigned drop = ~maskw[k];
                            if (k * 32 + 32 > N) drop &= (k * 32 < N) ? ((1u << (N - k * 32)) - 1u) : 0u;
                            while (drop) {
                                const int b = __ffs(drop) - 1;
                                drop &= drop - 1;
                                w.keys[k * 32 + b] = 127;
                            }
                        }
                    }
                }
                cons = col_centre<Tr, DT, NPT>(w, col, keep_s, R, p.centre2, col_const, col_fb, lo, list, refined2);
            }
            reinterpret_cast<val_t *>(p.consensus)[(size_t)r * D + t] = cons;
            if (p.want_moments) {
                val_t sk = 0, ku = 0;
                int stw = STW_OK;
                if (!shape2) {
                    auto fe = [&](auto f) {
                        for (int i = 0; i < N; ++i)
                            if (keep_s[i]) f(col[(size_t)i * D]);
                    };
                    column_moments<Tr>(fe, R, sk, ku, stw);
                }
                if (p.skew) reinterpret_cast<val_t *>(p.skew)[(size_t)r * D + t] = sk;
                if (p.kurt) reinterpret_cast<val_t *>(p.kurt)[(size_t)r * D + t] = ku;
                if (stw != STW_OK) atomicMin(stw_s, stw);
            }
        }
        // Without moments nothing after this point touches the group's shared state before the next round's
        // first barrier, so fast threads may run ahead into the next round's (thread-local) first phase.
        if (p.want_moments) {
            group_sync(bar_id, G);
            if (t == 0) {
                const int sw = *stw_s;
                p.status[r] = (sw == STW_OK) ? SVOC_ST_OK : (sw & 0xff);
            }
        }
    }
}

// bytes of shared memory one group needs (must match the carve-up above)
inline size_t col_group_smem(int N, int D, int G, int CS, size_t es, size_t vs) {
    auto al = [](size_t x) { return (x + 15) & ~(size_t)15; };
    const int NP = (N + 31) & ~31;
    size_t s = 0;
    s += al((size_t)D * CS);
    s += al((size_t)COL_CAP * G * es);
    s += al((size_t)D * vs);
    s += al((size_t)N * vs);
    s += al((size_t)2 * (G / 32) * vs);
    s += al((size_t)(NP / 32) * 4);
    s += 16;
    s += al((size_t)N);
    return (s + 127) & ~(size_t)127;
}

template <class Tr>
int launch_col(const Params &p, const Plan &pl, cudaStream_t st) {
    void (*kern)(const Params) = consensus_col_kernel<Tr, 0, 0>;
    if (p.D == 128 && p.N == 256) kern = consensus_col_kernel<Tr, 128, 256>;
    else if (p.D == 128) kern = consensus_col_kernel<Tr, 128, 0>;
    else if (p.D == 64) kern = consensus_col_kernel<Tr, 64, 0>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl.smem);
    if (e != cudaSuccess) return fail(std::string("cudaFuncSetAttribute: ") + cudaGetErrorString(e));
    kern<<<(unsigned)pl.grid, COL_THREADS, pl.smem, st>>>(p);
    e = cudaGetLastError();
    if (e != cudaSuccess) return fail(std::string("kernel launch: ") + cudaGetErrorString(e));
    count_launch();
    return 0;
}

}  // namespace svoc
