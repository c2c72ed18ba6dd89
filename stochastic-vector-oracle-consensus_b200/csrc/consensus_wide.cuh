// consensus_wide.cuh — the `wide` variant of the fused consensus round: 64 < N <= 512 oracles,
// 32-bit elements (fp32 / int32).  This is the kernel config 4 (N=256, D=128, fp32) runs.
//
// One persistent CTA (512 threads, one per SM) processes one round at a time out of shared
// memory: the N x D tile arrives by a single TMA bulk copy; a thread owns one (column, row-part)
// unit, P parts per column.  Order statistics are found without sorting:
//   1. per column, exact min/max -> every element gets a monotone 7-bit key (0..126, 127 is the
//      sentinel of padded / dropped rows), packed 4 per word in a transposed, bank-conflict-free
//      scratch (keysT[column][row], column stride == 16 bytes mod 128);
//   2. MSB-first bisection over the 7 key bits: counting "key >= pivot" is SWAR — one add puts
//      the comparison of four keys into the byte MSBs, one AND isolates them, one IDP.4A
//      accumulates — 3 ALU ops per 4 elements, two LDS.128 per 32 elements;
//   3. the rows whose key falls in the bucket of the wanted rank (and the next bucket) are
//      collected (a few per column), insertion-sorted, and the two order statistics read off.
//      Anything unusual (more than WIDE_CAP candidates, degenerate scale) falls back to the
//      exact generic bisection for that column only.
// The second pass reuses the keys: dropped rows are re-keyed to the sentinel.
// Reference semantics: see consensus_kernel.cuh (contract.cairo:442-503 etc.).
#pragma once
#include "consensus_kernel.cuh"

namespace svoc {

constexpr int WIDE_THREADS = 512;
constexpr int WIDE_CAP = 48;

// monotone map of an element onto [0, 1] for key building
template <class Tr>
struct WideKey;
template <>
struct WideKey<FloatTr<float>> {
    struct Scale { float inv, off; };
    __device__ static __forceinline__ bool make(float lo, float hi, Scale &s) {
        s.inv = 1.0f / (hi - lo);
        s.off = -lo * s.inv;
        return isfinite(s.inv) && isfinite(s.off);
    }
    __device__ static __forceinline__ float t(float x, const Scale &s) { return __saturatef(fmaf(x, s.inv, s.off)); }
};
template <>
struct WideKey<FixedTr<int32_t>> {
    struct Scale { int32_t lo; float inv; };
    __device__ static __forceinline__ bool make(int32_t lo, int32_t hi, Scale &s) {
        s.lo = lo;
        s.inv = 1.0f / (float)((uint32_t)hi - (uint32_t)lo);
        return true;
    }
    __device__ static __forceinline__ float t(int32_t x, const Scale &s) {
        // values below the scale's origin clamp to 0 (the unsigned difference would wrap)
        return x < s.lo ? 0.0f : __saturatef((float)((uint32_t)x - (uint32_t)s.lo) * s.inv);
    }
};

// 64-bit elements: the scale is applied in double precision, the key arithmetic stays fp32 (conversion to fp32
// rounds monotonically, and a key only has to be monotone)
template <>
struct WideKey<FloatTr<double>> {
    struct Scale { double inv, off; };
    __device__ static __forceinline__ bool make(double lo, double hi, Scale &s) {
        s.inv = 1.0 / (hi - lo);
        s.off = -lo * s.inv;
        return isfinite(s.inv) && isfinite(s.off);
    }
    __device__ static __forceinline__ float t(double x, const Scale &s) { return __saturatef((float)fma(x, s.inv, s.off)); }
};
template <>
struct WideKey<FixedTr<i64>> {
    struct Scale { i64 lo; double inv; };
    __device__ static __forceinline__ bool make(i64 lo, i64 hi, Scale &s) {
        s.lo = lo;
        s.inv = 1.0 / (double)((u64)hi - (u64)lo);
        return true;
    }
    __device__ static __forceinline__ float t(i64 x, const Scale &s) {
        return x < s.lo ? 0.0f : __saturatef((float)((double)((u64)x - (u64)s.lo) * s.inv));
    }
};

__device__ __forceinline__ double vmin(double a, double b) { return fmin(a, b); }
__device__ __forceinline__ double vmax(double a, double b) { return fmax(a, b); }
__device__ __forceinline__ i64 vmin(i64 a, i64 b) { return a < b ? a : b; }
__device__ __forceinline__ i64 vmax(i64 a, i64 b) { return a > b ? a : b; }
__device__ __forceinline__ float vmin(float a, float b) { return fminf(a, b); }
__device__ __forceinline__ float vmax(float a, float b) { return fmaxf(a, b); }
__device__ __forceinline__ int32_t vmin(int32_t a, int32_t b) { return min(a, b); }
__device__ __forceinline__ int32_t vmax(int32_t a, int32_t b) { return max(a, b); }

// barrier among the threads that share this unit's columns
__device__ __forceinline__ void unit_sync(int id, int count) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(count) : "memory");
}
__device__ __forceinline__ int unit_sync_or(int id, int count, int pred) {
    int r;
    asm volatile(
        "{\n"
        ".reg .pred p, q;\n"
        "setp.ne.s32 p, %1, 0;\n"
        "barrier.red.or.pred q, %2, %3, p;\n"
        "selp.s32 %0, 1, 0, q;\n"
        "}\n"
        : "=r"(r)
        : "r"(pred), "r"(id), "r"(count)
        : "memory");
    return r;
}

__device__ __forceinline__ unsigned pack4(unsigned a, unsigned b, unsigned c, unsigned d) {
    return __byte_perm(__byte_perm(a, b, 0x0040), __byte_perm(c, d, 0x0040), 0x5410);
}

// Counting "key >= cand" over packed byte keys (rp multiple of 32, keys <= 127, 1 <= cand <= 127).
// y = key + (128 - cand) sets the byte MSB exactly when key >= cand and never carries.  Summing the
// bytes of y as SIGNED values subtracts 256 for every set MSB, so with the plain byte sum known
//     #(key >= cand) = (sum_keys + rp * (128 - cand) - signed_sum(y)) / 256
// which costs one add and one IDP.4A per four keys.
template <int RPT>
__device__ __forceinline__ int key_byte_sum(const uint8_t *keys, int rp_rt) {
    const int rp = RPT ? RPT : rp_rt;
    unsigned acc = 0;
#pragma unroll
    for (int g = 0; g < rp; g += 16) {
        const uint4 a = *reinterpret_cast<const uint4 *>(keys + g);
        acc = __dp4a(a.x, 0x01010101u, acc);
        acc = __dp4a(a.y, 0x01010101u, acc);
        acc = __dp4a(a.z, 0x01010101u, acc);
        acc = __dp4a(a.w, 0x01010101u, acc);
    }
    return (int)acc;
}

template <int RPT>
__device__ __forceinline__ unsigned count_ge(const uint8_t *keys, int rp_rt, unsigned cand, int sum_keys) {
    const int rp = RPT ? RPT : rp_rt;
    // The add is written as VIADDMNMX (max(a + b, 0) on unsigned words is a + b): that instruction lives on the
    // integer ALU pipe, while the plain add is compiled to IMAD.IADD on the FMA pipe, which IDP.4A already fills.
    const unsigned cb = (128u - cand) * 0x01010101u;
    int acc0 = 0, acc1 = 0, acc2 = 0, acc3 = 0;
#pragma unroll
    for (int g = 0; g < rp; g += 32) {
        const uint4 a = *reinterpret_cast<const uint4 *>(keys + g);
        const uint4 b = *reinterpret_cast<const uint4 *>(keys + g + 16);
        acc0 = __dp4a((int)__viaddmax_u32(a.x, cb, 0u), 0x01010101, acc0);
        acc1 = __dp4a((int)__viaddmax_u32(a.y, cb, 0u), 0x01010101, acc1);
        acc2 = __dp4a((int)__viaddmax_u32(a.z, cb, 0u), 0x01010101, acc2);
        acc3 = __dp4a((int)__viaddmax_u32(a.w, cb, 0u), 0x01010101, acc3);
        acc0 = __dp4a((int)__viaddmax_u32(b.x, cb, 0u), 0x01010101, acc0);
        acc1 = __dp4a((int)__viaddmax_u32(b.y, cb, 0u), 0x01010101, acc1);
        acc2 = __dp4a((int)__viaddmax_u32(b.z, cb, 0u), 0x01010101, acc2);
        acc3 = __dp4a((int)__viaddmax_u32(b.w, cb, 0u), 0x01010101, acc3);
    }
    return (unsigned)(sum_keys + rp * (int)(128u - cand) - ((acc0 + acc1) + (acc2 + acc3))) >> 8;
}

struct WideCtx {
    int N, D, P, RP, CS;
    int c, part;        // this thread's unit
    bool has_unit;
    uint8_t *mykeys;    // keysT + c*CS + part*RP
    int *partial;       // [2][P][D]
    int *cand_cnt;      // [D]
    void *list;         // [WIDE_CAP][D] in_t
    int *flag;          // [D] 0 resolved, 1 needs the minimum above `thr`, 2 needs a finer re-keying
    void *thr, *rlo, *rhi;          // [D] in_t each
    int *best;                      // [3][D] packed (rank << 8 | slot) maxima: lower, upper statistic, window maximum
    int bar_id, bar_count;          // barrier of this unit's slab (32 columns x P parts), or the CTA barrier
    bool slab;                      // slab mode: threads without a unit skip the selection entirely
    uint32_t *wmin, *wmax, *resc;   // [D] order-preserving keys: window min / max, rescue minimum
};

// 7-bit key of one element.  Level 0: the scale spans the column's exact [min, max] and keys use
// 0..126.  Refinement levels: the scale spans a sub-range; 0 = below it, 126 = above it, 1..125 inside.
// KTOP = the largest key (126 for 7-bit keys; the mid kernel's 64-row columns use 5-bit keys, KTOP = 30).
template <class Tr, bool FIRST, int KTOP = 126>
__device__ __forceinline__ unsigned wide_key(typename Tr::in_t x, typename Tr::in_t lo, typename Tr::in_t hi,
                                             const typename WideKey<Tr>::Scale &sc) {
    const float t = WideKey<Tr>::t(x, sc);
    if (FIRST) return __float_as_uint(__fmaf_rn(t, (float)KTOP, 8388608.0f));
    const unsigned k = __float_as_uint(__fmaf_rn(t, (float)(KTOP - 2), 8388609.0f));   // 1 + round((KTOP - 2) t) in the low byte
    return x < lo ? 0u : (x > hi ? (unsigned)KTOP : k);
}

template <class Tr, bool FIRST, int DT, int RPT>
__device__ __forceinline__ void wide_build_keys(const WideCtx &w, const typename Tr::in_t *tile, const uint8_t *keep,
                                                typename Tr::in_t lo, typename Tr::in_t hi,
                                                const typename WideKey<Tr>::Scale &sc) {
    const int WD = DT ? DT : w.D;
    const int WRP = RPT ? RPT : w.RP;
    const int row_lo = w.part * WRP;
    const typename Tr::in_t *ptr = tile + row_lo * WD + w.c;
#pragma unroll 2
    for (int g = 0; g < WRP; g += 16) {
        unsigned f[16];
        if (!keep && row_lo + g + 16 <= w.N) {   // whole group valid: no per-row checks
#pragma unroll
            for (int k = 0; k < 16; ++k) f[k] = wide_key<Tr, FIRST>(ptr[k * WD], lo, hi, sc);
        } else {
#pragma unroll
            for (int k = 0; k < 16; ++k) {
                const int row = row_lo + g + k;
                f[k] = 127u;
                if (row < w.N && (!keep || keep[row])) f[k] = wide_key<Tr, FIRST>(ptr[k * WD], lo, hi, sc);
            }
        }
        ptr += 16 * WD;
        *reinterpret_cast<uint4 *>(w.mykeys + g) =
            make_uint4(pack4(f[0], f[1], f[2], f[3]), pack4(f[4], f[5], f[6], f[7]), pack4(f[8], f[9], f[10], f[11]),
                       pack4(f[12], f[13], f[14], f[15]));
    }
}

// Centre of the flagged rows of every column (all rows when keep == nullptr).  Must be called by
// all threads of the CTA (it synchronises); the result is valid in the part-0 thread of a column.
template <class Tr, int DT, int PT, int RPT>
__device__ __forceinline__ typename Tr::val_t wide_centre(const WideCtx &w, const typename Tr::in_t *tile,
                                                          const uint8_t *keep, int n_valid, int kind, bool col_const,
                                                          bool col_fallback, typename Tr::in_t const_value, int &buf,
                                                          bool &refined) {
    using in_t = typename Tr::in_t;
    using val_t = typename Tr::val_t;
    using Key = WideKey<Tr>;
    const int WD = DT ? DT : w.D;
    const int WP = PT ? PT : w.P;
    const int WRP = RPT ? RPT : w.RP;
    refined = false;
    val_t centre = 0;
    if (kind == SVOC_CENTRE_MEAN) {
        if (w.has_unit && w.part == 0) {
            val_t s = 0;
            for (int i = 0; i < w.N; ++i)
                if (!keep || keep[i]) s += (val_t)tile[i * WD + w.c];
            centre = Tr::mean(s, n_valid);
        }
        return centre;
    }
    if (w.slab && !w.has_unit) return centre;   // not a member of any slab barrier
    int klo, khi;
    centre_ranks(kind, n_valid, klo, khi);
    in_t *list = reinterpret_cast<in_t *>(w.list);
    in_t *thr = reinterpret_cast<in_t *>(w.thr), *rlo = reinterpret_cast<in_t *>(w.rlo), *rhi = reinterpret_cast<in_t *>(w.rhi);
    const int slots = WP * WRP;
    const int row_lo = w.part * WRP, row_hi = min(w.N, row_lo + WRP);
    const int pstride = WP * WD;                        // one buffer of partial counts
    int *my_partial = w.partial + w.part * WD + w.c + buf * pstride;     // this pass's slot ...
    int *my_partial_alt = w.partial + w.part * WD + w.c + (buf ^ 1) * pstride;   // ... and the next pass's
    const int *col_partial = w.partial + w.c + buf * pstride;
    const int *col_partial_alt = w.partial + w.c + (buf ^ 1) * pstride;
    int *best1 = w.best + w.c, *best2 = w.best + WD + w.c, *best3 = w.best + 2 * WD + w.c;
    const in_t *lc = list + w.c;
    bool done = !w.has_unit || col_const;
    in_t v_lo = const_value, v_hi = const_value;
    bool have_centre = false;

    for (;;) {
        const bool active = !done;
        // ---- bisection over the 7 key bits for rank klo
        unsigned prefix = 0;
        int L = 0, LE = n_valid;
        const int sum_keys = active ? key_byte_sum<RPT>(w.mykeys, WRP) : 0;
#pragma unroll
        for (int bit = 6; bit >= 0; --bit) {
            const unsigned cand = prefix | (1u << bit);
            if (active) *my_partial = (int)count_ge<RPT>(w.mykeys, WRP, cand, sum_keys);
            unit_sync(w.bar_id, w.bar_count);
            if (active) {
                int tot = 0;
#pragma unroll
                for (int q = 0; q < WP; ++q) tot += col_partial[q * WD];
                const int lt = slots - tot;  // sentinels (127) are never below a pivot
                if (lt <= klo) { prefix = cand; L = lt; }
                else LE = lt;    // the last rejected pivot is prefix + 1: LE = #(key <= prefix)
            }
            buf ^= 1;
            { int *t = my_partial; my_partial = my_partial_alt; my_partial_alt = t; }
            { const int *t = col_partial; col_partial = col_partial_alt; col_partial_alt = t; }
        }
        // ---- collect the rows whose key lies in the window: the bucket of rank klo alone when rank khi
        //      is inside it too, else that bucket and the next two
        if (active) {
            const unsigned span = (khi < LE) ? 1u : 3u;
            const unsigned up = prefix + span > 127u ? 127u : prefix + span;
            const unsigned c1b = (128u - prefix) * 0x01010101u, c2b = (128u - up) * 0x01010101u;
            uint32_t kmin = 0xffffffffu, kmax = 0u;
#pragma unroll
            for (int g = 0; g < WRP; g += 32) {
                const uint4 a = *reinterpret_cast<const uint4 *>(w.mykeys + g);
                const uint4 b = *reinterpret_cast<const uint4 *>(w.mykeys + g + 16);
                unsigned acc = (a.x + c1b) & ~(a.x + c2b) & 0x80808080u;
                acc |= ((a.y + c1b) & ~(a.y + c2b) & 0x80808080u) >> 1;
                acc |= ((a.z + c1b) & ~(a.z + c2b) & 0x80808080u) >> 2;
                acc |= ((a.w + c1b) & ~(a.w + c2b) & 0x80808080u) >> 3;
                acc |= ((b.x + c1b) & ~(b.x + c2b) & 0x80808080u) >> 4;
                acc |= ((b.y + c1b) & ~(b.y + c2b) & 0x80808080u) >> 5;
                acc |= ((b.z + c1b) & ~(b.z + c2b) & 0x80808080u) >> 6;
                acc |= ((b.w + c1b) & ~(b.w + c2b) & 0x80808080u) >> 7;
                while (acc) {
                    const int pos = __ffs(acc) - 1;
                    acc &= acc - 1;
                    const int row = row_lo + g + 4 * (7 - (pos & 7)) + (pos >> 3);
                    const in_t x = tile[row * WD + w.c];
                    const uint32_t kx = okey(x);
                    kmin = kx < kmin ? kx : kmin;
                    kmax = kx > kmax ? kx : kmax;
                    const int slot = atomicAdd(&w.cand_cnt[w.c], 1);
                    if (slot < WIDE_CAP) list[slot * WD + w.c] = x;
                }
            }
            if (kmin <= kmax) {
                atomicMin(&w.wmin[w.c], kmin);
                atomicMax(&w.wmax[w.c], kmax);
            }
        }
        unit_sync(w.bar_id, w.bar_count);
        // ---- rank the window's rows: the P threads of a column take every P-th slot, four at a time.
        //      lt(v) = #{u < v}; the value of rank j is the candidate with the largest lt <= j.
        int flag = 0;
        const int m = active ? w.cand_cnt[w.c] : 0;   // rows in the window (not capped)
        const int j1 = klo - L, j2 = khi - L;
        const bool rankable = active && !col_fallback && m <= WIDE_CAP;
        if (rankable) {
            for (int base = w.part; base < m; base += 4 * WP) {
                const int a0 = base, a1 = base + WP, a2 = base + 2 * WP, a3 = base + 3 * WP;
                const in_t v0 = lc[a0 * WD];
                const in_t v1 = a1 < m ? lc[a1 * WD] : v0, v2 = a2 < m ? lc[a2 * WD] : v0, v3 = a3 < m ? lc[a3 * WD] : v0;
                int r0 = 0, r1 = 0, r2 = 0, r3 = 0;
#pragma unroll 2
                for (int b = 0; b < m; ++b) {
                    const in_t u = lc[b * WD];
                    r0 += u < v0 ? 1 : 0;
                    r1 += u < v1 ? 1 : 0;
                    r2 += u < v2 ? 1 : 0;
                    r3 += u < v3 ? 1 : 0;
                }
                auto publish = [&](int rk, int slot) {
                    if (rk <= j1) atomicMax(best1, (rk << 8) | slot);
                    if (rk <= j2) atomicMax(best2, (rk << 8) | slot);
                    if (j2 >= m) atomicMax(best3, (rk << 8) | slot);
                };
                publish(r0, a0);
                if (a1 < m) publish(r1, a1);
                if (a2 < m) publish(r2, a2);
                if (a3 < m) publish(r3, a3);
            }
            if (j2 >= m) flag = 1;
        } else if (active && w.part == 0) {
            if (col_fallback) {
                val_t sk, ku;
                int stw = STW_OK;
                centre = column_centre_generic<Tr>(tile + w.c, WD, w.N, keep, n_valid, kind, false, sk, ku, stw);
                have_centre = true;
            } else {
                in_t a, b;
                unkey(w.wmin[w.c], a);
                unkey(w.wmax[w.c], b);
                typename Key::Scale sc;
                if (!(b > a)) {              // the whole window is one value
                    v_lo = a;
                    v_hi = a;
                    have_centre = j2 < m;
                    centre = Tr::half_sum((val_t)a, (val_t)a);
                    if (j2 >= m) { flag = 1; thr[w.c] = a; }
                } else if (!Key::make(a, b, sc)) {
                    val_t sk, ku;
                    int stw = STW_OK;
                    centre = column_centre_generic<Tr>(tile + w.c, WD, w.N, keep, n_valid, kind, false, sk, ku, stw);
                    have_centre = true;
                } else {
                    flag = 2;
                    rlo[w.c] = a;
                    rhi[w.c] = b;
                }
            }
        }
        if (active && w.part == 0) {
            w.flag[w.c] = flag;
            w.wmin[w.c] = 0xffffffffu;
            w.wmax[w.c] = 0u;
            w.resc[w.c] = 0xffffffffu;
        }
        const int any = unit_sync_or(w.bar_id, w.bar_count, flag);
        if (active && w.part == 0) {
            w.cand_cnt[w.c] = 0;   // everyone has read it; the next collect comes after more barriers
            if (rankable) {
                v_lo = lc[(*best1 & 0xff) * WD];
                if (j2 < m) v_hi = lc[(*best2 & 0xff) * WD];
                else thr[w.c] = lc[(*best3 & 0xff) * WD];
                *best1 = -1;
                *best2 = -1;
                *best3 = -1;
            }
        }
        if (!any) break;            // common case: every column resolved
        unit_sync(w.bar_id, w.bar_count);   // thr / ranges written by the part-0 threads are visible
        const int f = active ? w.flag[w.c] : 0;
        if (f == 1) {                          // minimum of the flagged rows above the window
            const in_t t0 = thr[w.c];
            uint32_t best = 0xffffffffu;
            const in_t *ptr = tile + row_lo * WD + w.c;
            for (int row = row_lo; row < row_hi; ++row, ptr += WD) {
                if (keep && !keep[row]) continue;
                const in_t x = *ptr;
                if (x > t0) { const uint32_t kx = okey(x); best = kx < best ? kx : best; }
            }
            if (best != 0xffffffffu) atomicMin(&w.resc[w.c], best);
        } else if (f == 2) {                   // finer keys over the window's value range
            const in_t a = rlo[w.c], b = rhi[w.c];
            typename Key::Scale sc;
            Key::make(a, b, sc);
            wide_build_keys<Tr, false, DT, RPT>(w, tile, keep, a, b, sc);
            refined = true;
        }
        const int more = unit_sync_or(w.bar_id, w.bar_count, f == 2);
        if (f == 1 && w.part == 0) unkey(w.resc[w.c], v_hi);
        if (f != 2) done = true;
        if (!more) break;
    }
    if (w.has_unit && w.part == 0 && !have_centre) centre = Tr::half_sum((val_t)v_lo, (val_t)v_hi);
    return centre;
}

// GT = 1: the tile is NOT staged in shared memory; the three sweeps read the round straight from global
// memory (first sweep from HBM, the later ones from L2, the next round prefetched into L2), which
// leaves room for two CTAs — two rounds in flight — per SM.
template <class Tr, int DT, int PT, int NT, int GT, int RPT>
__global__ void __launch_bounds__(NT, GT ? 2 : 1) consensus_wide_kernel(const Params p) {
    using in_t = typename Tr::in_t;
    using val_t = typename Tr::val_t;
    using acc_t = typename Tr::acc_t;
    using Key = WideKey<Tr>;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int N = p.N, D = DT ? DT : p.D, R = p.R, P = PT ? PT : p.wP, RP = RPT ? RPT : p.wRP, CS = p.wCS, H = p.wH;
    const int tid = threadIdx.x;
    auto al = [](size_t x) { return (x + 15) & ~(size_t)15; };

    size_t off = 0;
    in_t *tile_s = reinterpret_cast<in_t *>(smem_raw);          off += GT ? 0 : al((size_t)N * D * sizeof(in_t));
    val_t *c1 = reinterpret_cast<val_t *>(smem_raw + off);      off += al((size_t)D * sizeof(val_t));
    acc_t *qr_s = reinterpret_cast<acc_t *>(smem_raw + off);    off += al((size_t)N * sizeof(acc_t));
    acc_t *qrp = reinterpret_cast<acc_t *>(smem_raw + off);     off += al((size_t)H * N * sizeof(acc_t));
    u64 *bar = reinterpret_cast<u64 *>(smem_raw + off);         off += 16;
    uint8_t *keysT = smem_raw + off;                            off += al((size_t)D * CS);
    in_t *list = reinterpret_cast<in_t *>(smem_raw + off);      off += al((size_t)WIDE_CAP * D * sizeof(in_t));
    in_t *smin = reinterpret_cast<in_t *>(smem_raw + off);      off += al((size_t)P * D * sizeof(in_t));
    in_t *smax = reinterpret_cast<in_t *>(smem_raw + off);      off += al((size_t)P * D * sizeof(in_t));
    int *partial = reinterpret_cast<int *>(smem_raw + off);     off += al((size_t)2 * P * D * sizeof(int));
    int *cand_cnt = reinterpret_cast<int *>(smem_raw + off);    off += al((size_t)D * sizeof(int));
    int *rankp = reinterpret_cast<int *>(smem_raw + off);       off += al((size_t)H * N * sizeof(int));
    uint32_t *maskw = reinterpret_cast<uint32_t *>(smem_raw + off); off += al((size_t)(P * RP / 32 + 1) * 4);
    int *stw_s = reinterpret_cast<int *>(smem_raw + off);       off += 16;
    int *flag_s = reinterpret_cast<int *>(smem_raw + off);      off += al((size_t)D * 4);
    in_t *thr_s = reinterpret_cast<in_t *>(smem_raw + off);     off += al((size_t)D * sizeof(in_t));
    in_t *rlo_s = reinterpret_cast<in_t *>(smem_raw + off);     off += al((size_t)D * sizeof(in_t));
    in_t *rhi_s = reinterpret_cast<in_t *>(smem_raw + off);     off += al((size_t)D * sizeof(in_t));
    uint32_t *wmin_s = reinterpret_cast<uint32_t *>(smem_raw + off); off += al((size_t)D * 4);
    uint32_t *wmax_s = reinterpret_cast<uint32_t *>(smem_raw + off); off += al((size_t)D * 4);
    uint32_t *resc_s = reinterpret_cast<uint32_t *>(smem_raw + off); off += al((size_t)D * 4);
    int *best_s = reinterpret_cast<int *>(smem_raw + off);      off += al((size_t)3 * D * 4);
    uint8_t *keep_s = smem_raw + off;

    WideCtx w;
    w.N = N; w.D = D; w.P = P; w.RP = RP; w.CS = CS;
    w.has_unit = tid < D * P;
    w.c = w.has_unit ? tid % D : 0;
    w.part = w.has_unit ? tid / D : 0;
    w.mykeys = keysT + (size_t)w.c * CS + (size_t)w.part * RP;
    w.partial = partial;
    w.cand_cnt = cand_cnt;
    w.list = list;
    w.flag = flag_s; w.thr = thr_s; w.rlo = rlo_s; w.rhi = rhi_s;
    w.wmin = wmin_s; w.wmax = wmax_s; w.resc = resc_s;
    w.best = best_s;
    // slab mode: the P warps that own the same 32 columns synchronise among themselves
    w.slab = (D % 32 == 0) && (D / 32 <= 15);
    w.bar_id = w.slab ? 1 + (w.c >> 5) : 0;
    w.bar_count = w.slab ? 32 * P : NT;

    const size_t round_elems = (size_t)N * D;
    const uint32_t round_bytes = (uint32_t)(round_elems * sizeof(in_t));
    if (tid == 0) mbar_init(bar, 1);
    for (int k = tid; k < D; k += NT) {
        cand_cnt[k] = 0;
        wmin_s[k] = 0xffffffffu;
        wmax_s[k] = 0u;
        resc_s[k] = 0xffffffffu;
        flag_s[k] = 0;
        best_s[k] = -1;
        best_s[D + k] = -1;
        best_s[2 * D + k] = -1;
    }
    __syncthreads();
    uint32_t parity = 0;
    int buf = 0;
    const int row_lo = w.part * RP, row_hi = min(N, (w.part + 1) * RP);
    const bool shape2 = second_pass_shape_panic(p);

    for (i64 r = blockIdx.x; r < p.T; r += gridDim.x) {
        const in_t *gsrc = reinterpret_cast<const in_t *>(p.pred) + (size_t)r * round_elems;
        // ---- phase 0: stage the round (or, GT: only pull the NEXT round towards L2)
        const in_t *tile = GT ? gsrc : tile_s;
        if (GT) {
            if (tid == 0 && p.wBulk && r + gridDim.x < p.T)
                asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(gsrc + (size_t)gridDim.x * round_elems), "r"(round_bytes) : "memory");
        } else if (p.wBulk) {
            if (tid == 0) {
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic reads of the old tile before the async write
                mbar_expect_tx(bar, round_bytes);
                bulk_g2s(tile_s, gsrc, round_bytes, bar);
            }
            mbar_wait(bar, parity);
            parity ^= 1;
        } else {
            for (size_t e = tid; e < round_elems; e += NT) tile_s[e] = gsrc[e];
        }
        if (tid == 0) *stw_s = STW_OK;
        __syncthreads();

        // ---- phase 1a: exact column min / max
        if (w.has_unit) {
            in_t lo = std::numeric_limits<in_t>::max(), hi = std::numeric_limits<in_t>::lowest();
            const in_t *ptr = tile + row_lo * D + w.c;
#pragma unroll 8
            for (int row = row_lo; row < row_hi; ++row, ptr += D) {
                const in_t x = *ptr;
                lo = vmin(lo, x);
                hi = vmax(hi, x);
            }
            smin[w.part * D + w.c] = lo;
            smax[w.part * D + w.c] = hi;
        }
        __syncthreads();
        // ---- phase 1b: 7-bit keys of this unit's rows
        bool col_const = false, col_fb = false;
        in_t cval = 0, lo0 = 0, hi0 = 0;
        typename Key::Scale sc0{};
        if (w.has_unit) {
            in_t lo = smin[w.c], hi = smax[w.c];
            for (int q = 1; q < P; ++q) {
                const in_t a = smin[q * D + w.c], b = smax[q * D + w.c];
                lo = vmin(lo, a);
                hi = vmax(hi, b);
            }
            col_const = !(hi > lo);
            cval = lo;
            typename Key::Scale sc;
            col_fb = !col_const && !Key::make(lo, hi, sc);
            if (!col_const && !col_fb) {
                wide_build_keys<Tr, true, DT, RPT>(w, tile, nullptr, lo, hi, sc);
            } else {
                for (int g = 0; g < RP; g += 16) *reinterpret_cast<uint4 *>(w.mykeys + g) = make_uint4(0, 0, 0, 0);
            }
            lo0 = lo; hi0 = hi; sc0 = sc;
        }
        // ---- phase 1c: first-pass centre
        bool refined1 = false, refined2 = false;
        const val_t e1 = wide_centre<Tr, DT, PT, RPT>(w, tile, nullptr, N, p.centre, col_const, col_fb, cval, buf, refined1);
        if (w.has_unit && w.part == 0) {
            c1[w.c] = e1;
            if (p.essence1) reinterpret_cast<val_t *>(p.essence1)[(size_t)r * D + w.c] = e1;
        }
        __syncthreads();

        // ---- phase 2: quadratic risk of every row (H threads per row, each a slice of the dimensions).
        //      128-bit shared loads, start rotated by 16 bytes per row so that the 8 lanes of a
        //      quarter-warp hit 8 different bank groups.
        if (tid < N * H) {
            const int i = tid % N, h = tid / N;
            const int dh = (((D + H - 1) / H) + 3) & ~3, dlo = h * dh, dhi = min(D, dlo + dh);
            acc_t q = 0;
            bool bad = false;
            auto one = [&](in_t x, val_t c) {
                if constexpr (Tr::fixed) {
                    if (p.constrained) bad |= (x < 0) | ((i64)x > WSAD);
                    else bad |= ((i64)x > ENGINE_RANGE) | ((i64)x < -ENGINE_RANGE);
                }
                q += Tr::dev(x, c);
            };
            if (dlo < dhi) {
                const in_t *row = tile + i * D;
                if ((D & 3) == 0) {
                    const int ng = (dhi - dlo) >> 2;          // groups of 4 dimensions
                    int g = i % ng;
                    for (int k = 0; k < ng; ++k) {
                        const int d = dlo + 4 * g;
                        const int4 xv = *reinterpret_cast<const int4 *>(row + d);
                        in_t x0, x1, x2, x3;
                        memcpy(&x0, &xv.x, 4); memcpy(&x1, &xv.y, 4); memcpy(&x2, &xv.z, 4); memcpy(&x3, &xv.w, 4);
                        one(x0, c1[d]); one(x1, c1[d + 1]); one(x2, c1[d + 2]); one(x3, c1[d + 3]);
                        g = (g + 1 == ng) ? 0 : g + 1;
                    }
                } else {
                    int d = dlo + i % (dhi - dlo);
                    for (int k = dlo; k < dhi; ++k) {
                        one(row[d], c1[d]);
                        d = (d + 1 == dhi) ? dlo : d + 1;
                    }
                }
            }
            qrp[h * N + i] = q;
            if (bad) atomicMin(stw_s, p.constrained ? STW_INTERVAL : STW_ENGINE_RANGE);
        }
        __syncthreads();
        if (tid < N) {
            acc_t q = 0;
            for (int h = 0; h < H; ++h) q += qrp[h * N + tid];
            qr_s[tid] = q;
            if (p.qr) reinterpret_cast<acc_t *>(p.qr)[(size_t)r * N + tid] = q;
        }
        __syncthreads();

        // ---- phase 3a: rank of every row under (qr asc, index desc | asc).  For the rows of one warp the
        //      oracles below the warp's first row and above its last row need a single compare each
        //      (strict or non-strict by side); only the warp's own index range needs the tie rule.
        {
            const bool mine_ok = tid < N * H;
            const int i = mine_ok ? tid % N : 0, h = mine_ok ? tid / N : 0;
            const int jh = (N + H - 1) / H, jlo = h * jh, jhi = mine_ok ? min(N, jlo + jh) : jlo;
            const acc_t mine = qr_s[i];
            const int imin = __reduce_min_sync(0xffffffffu, i), imax = __reduce_max_sync(0xffffffffu, i);
            const bool high = p.tie == SVOC_TIE_HIGH_INDEX_KEPT;
            int rank = 0;
            const int a_end = min(jhi, max(jlo, imin));        // [jlo, a_end): every j < i
            const int b_beg = max(jlo, min(jhi, imax + 1));    // [b_beg, jhi): every j > i
            // one compare per pair; 128-bit broadcast loads where the range allows
            auto count_lt = [&](int j0, int j1) {
                int c = 0, j = j0;
                if constexpr (sizeof(acc_t) == 4) {
                    for (; j < j1 && (j & 3); ++j) c += qr_s[j] < mine ? 1 : 0;
                    for (; j + 4 <= j1; j += 4) {
                        const float4 v = *reinterpret_cast<const float4 *>(reinterpret_cast<const float *>(qr_s) + j);
                        c += (v.x < mine ? 1 : 0) + (v.y < mine ? 1 : 0) + (v.z < mine ? 1 : 0) + (v.w < mine ? 1 : 0);
                    }
                }
                for (; j < j1; ++j) c += qr_s[j] < mine ? 1 : 0;
                return c;
            };
            auto count_le = [&](int j0, int j1) {
                int c = 0, j = j0;
                if constexpr (sizeof(acc_t) == 4) {
                    for (; j < j1 && (j & 3); ++j) c += qr_s[j] <= mine ? 1 : 0;
                    for (; j + 4 <= j1; j += 4) {
                        const float4 v = *reinterpret_cast<const float4 *>(reinterpret_cast<const float *>(qr_s) + j);
                        c += (v.x <= mine ? 1 : 0) + (v.y <= mine ? 1 : 0) + (v.z <= mine ? 1 : 0) + (v.w <= mine ? 1 : 0);
                    }
                }
                for (; j < j1; ++j) c += qr_s[j] <= mine ? 1 : 0;
                return c;
            };
            if (high) {
                rank = count_lt(jlo, a_end) + count_le(b_beg, jhi);
                for (int j = a_end; j < b_beg; ++j) rank += (qr_s[j] < mine || (qr_s[j] == mine && j > i)) ? 1 : 0;
            } else {
                rank = count_le(jlo, a_end) + count_lt(b_beg, jhi);
                for (int j = a_end; j < b_beg; ++j) rank += (qr_s[j] < mine || (qr_s[j] == mine && j < i)) ? 1 : 0;
            }
            if (mine_ok) rankp[h * N + i] = rank;
        }
        __syncthreads();
        {
            // keep flags + packed mask words (one ballot per warp of rows)
            const int slots = P * RP;
            bool keepme = false;
            if (tid < N) {
                int rank = 0;
                for (int h = 0; h < H; ++h) rank += rankp[h * N + tid];
                keepme = rank < R;
                keep_s[tid] = keepme ? 1 : 0;
                if (p.rank) p.rank[(size_t)r * N + tid] = rank;
            }
            if (tid < slots || tid < ((N + 31) & ~31)) {
                const unsigned bits = __ballot_sync(0xffffffffu, keepme);
                if ((tid & 31) == 0) {
                    maskw[tid >> 5] = bits;
                    if (p.reliable && (tid >> 5) < p.words) p.reliable[(size_t)r * p.words + (tid >> 5)] = bits;
                }
            }
        }
        __syncthreads();
        // ---- phase 3b: reliabilities (warp 0) while the other warps re-key the dropped rows
        if (tid < 32) {
            acc_t s1 = 0, s2 = 0;
            for (int i = tid; i < N; i += 32) {
                const acc_t q = qr_s[i];
                s1 += q;
                if (keep_s[i]) s2 += q;
            }
            for (int o = 16; o; o >>= 1) {
                s1 += __shfl_xor_sync(0xffffffffu, s1, o);
                s2 += __shfl_xor_sync(0xffffffffu, s2, o);
            }
            if (tid == 0) {
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
            }
        }
        // level-0 keys of the kept rows, sentinel for the dropped ones (the first pass may have left
        // refined keys behind, so the unit is re-keyed rather than patched)
        if (w.has_unit && !col_const && !col_fb) {
            if (refined1) {
                wide_build_keys<Tr, true, DT, RPT>(w, tile, keep_s, lo0, hi0, sc0);
            } else {
                for (int g = 0; g < RP; g += 32) {   // dropped rows of this unit -> sentinel key
                    const int base = row_lo + g;
                    unsigned drop = ~maskw[base >> 5];
                    if (base + 32 > N) drop &= (base < N) ? ((1u << (N - base)) - 1u) : 0u;
                    while (drop) {
                        const int b = __ffs(drop) - 1;
                        drop &= drop - 1;
                        w.mykeys[g + b] = 127;
                    }
                }
            }
        }
        // ---- phase 4: second-pass centre over the reliable rows (+ moments)
        val_t cons = 0;
        if (!shape2) cons = wide_centre<Tr, DT, PT, RPT>(w, tile, keep_s, R, p.centre2, col_const, col_fb, cval, buf, refined2);
        if (w.has_unit && w.part == 0) {
            reinterpret_cast<val_t *>(p.consensus)[(size_t)r * D + w.c] = cons;
            if (p.want_moments) {
                val_t sk = 0, ku = 0;
                int stw = STW_OK;
                if (!shape2) {
                    const in_t *col = tile + w.c;
                    auto fe = [&](auto f) {
                        for (int i = 0; i < N; ++i)
                            if (keep_s[i]) f(col[(size_t)i * D]);
                    };
                    column_moments<Tr>(fe, R, sk, ku, stw);
                }
                if (p.skew) reinterpret_cast<val_t *>(p.skew)[(size_t)r * D + w.c] = sk;
                if (p.kurt) reinterpret_cast<val_t *>(p.kurt)[(size_t)r * D + w.c] = ku;
                if (stw != STW_OK) atomicMin(stw_s, stw);
            }
        }
        __syncthreads();  // every read of the tile is done: the next round may overwrite it
        if (tid == 0) {
            const int sw = *stw_s;
            p.status[r] = (sw == STW_OK) ? SVOC_ST_OK : (sw & 0xff);
        }
    }
}

// shared-memory footprint of the wide variant (must match the carve-up above)
inline size_t wide_smem_bytes(int N, int D, int P, int RP, int CS, int H, size_t es, size_t vs, bool global_tile = false) {
    auto al = [](size_t x) { return (x + 15) & ~(size_t)15; };
    size_t s = 0;
    if (!global_tile) s += al((size_t)N * D * es);
    s += al((size_t)D * vs);
    s += al((size_t)N * vs);
    s += al((size_t)H * N * vs);
    s += 16;
    s += al((size_t)D * CS);
    s += al((size_t)WIDE_CAP * D * es);
    s += 2 * al((size_t)P * D * es);
    s += al((size_t)2 * P * D * 4);
    s += al((size_t)D * 4);
    s += al((size_t)H * N * 4);
    s += al((size_t)(P * RP / 32 + 1) * 4);
    s += 16;
    s += al((size_t)D * 4) + 3 * al((size_t)D * es) + 3 * al((size_t)D * 4) + al((size_t)3 * D * 4);
    s += al((size_t)N);
    return s;
}

template <class Tr>
int launch_wide(const Params &p, const Plan &pl, cudaStream_t st) {
    // `wide` is the fallback of `wide2` (rounds that are not 16-byte aligned): the tile is staged in shared
    // memory.  One instantiation with compile-time strides for the 128-dimension shape, one generic.
    void (*kern)(const Params) = consensus_wide_kernel<Tr, 0, 0, WIDE_THREADS, 0, 0>;
    if (p.D == 128 && p.wP == 4 && p.wRP == 64) kern = consensus_wide_kernel<Tr, 128, 4, WIDE_THREADS, 0, 64>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl.smem);
    if (e != cudaSuccess) return fail(std::string("cudaFuncSetAttribute: ") + cudaGetErrorString(e));
    kern<<<(unsigned)pl.grid, pl.threads, pl.smem, st>>>(p);
    e = cudaGetLastError();
    if (e != cudaSuccess) return fail(std::string("kernel launch: ") + cudaGetErrorString(e));
    count_launch();
    return 0;
}

}  // namespace svoc
