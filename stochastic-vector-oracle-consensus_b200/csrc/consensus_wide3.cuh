// consensus_wide3.cuh — the `wide3` variant of the fused consensus round: 64 < N <= 256 oracles, D = 128
// dimensions, 32-bit elements (fp32 / int32).  This is the kernel BASELINE config 4 runs.
//
// It keeps the selection method of `wide2` (monotone 7-bit keys, SWAR bisection, exact ranking of a small
// window; risks selected by one warp) and changes where the round lives, after the round-1 profiles showed
// `wide2` waiting on L2 round trips (every sweep after the first re-read the round from L2) and spending a
// fifth of its instructions on loads, address arithmetic and shared-memory key traffic:
//
//   * A round is split by COLUMNS over a cluster of two CTAs (64 columns each, 256 threads): two CTAs — halves
//     of two different rounds — share an SM and fill each other's waits.
//   * Each CTA's half round (256 rows x 64 columns, 64 KB) is fetched by two 2-D TMA tensor-map copies
//     (cp.async.bulk.tensor, 128-byte swizzle, completing on an mbarrier) into shared memory and STAYS there for
//     the whole round: the key build reads it column-wise (a thread owns rows 8k + 2*part + {0,1} of one column;
//     with the swizzle the 32 lanes of a warp hit 32 different banks), the risk sweep row-wise with 128-bit
//     loads (a quarter-warp reads one 128-byte line), and the few window candidates of a column are fetched by
//     row index.  Nothing in the round touches global memory or L2 again except the outputs.  The next round
//     is pulled into L2 while this one is computed (cp.async.bulk.prefetch.tensor) and copied into shared memory
//     as soon as the last thread has made its last read (an "empty" mbarrier), so the copy is an L2 hit whose
//     latency the co-resident CTA covers.
//   * Keys (16 packed words per thread) live in registers: the bisection has no memory traffic at all.
//   * The per-oracle risks need all 128 columns: each CTA sums its 64 columns (a transposing shuffle tree over
//     the 16 lanes that share a row leaves every thread with ONE row total), sends its 256 partial sums into
//     the PEER's shared memory with st.async, which completes on the peer's mbarrier — no cluster barrier, no
//     fence — and both CTAs form partial(rank 0) + partial(rank 1), bit-identical, and select the same rows
//     redundantly.  Buffers and barriers are double-buffered by round parity.
//
// A first version kept the round in REGISTERS (64 values per thread, 128 registers, the landing buffer freed
// right after the key build): parity-green but 12.2 M rounds/s against 14.2 M for wide2 — window candidates had
// to be extracted with a predicated store per element (registers cannot be indexed), which cost as much as the
// loads it saved (profiles/r02b_*).
//
// Anything this kernel does not cover (moments, rank output, mean centres, other D, misaligned rounds) is
// dispatched to `wide2`.  Reference semantics: see consensus_kernel.cuh (contract.cairo:442-503 etc.).
#pragma once
#include <cuda.h>

#include "consensus_wide2.cuh"

namespace svoc {

constexpr int W3_COLS = 64;          // columns per CTA
#ifndef W3_DEFAULT_LANES
#define W3_DEFAULT_LANES 4
#endif
constexpr int W3_D = 128;            // columns per round (two CTAs)
constexpr int W3_ROWS = 256;         // rows of a landing box (rows >= N are ignored)
constexpr int W3_CAP = 16;           // candidate slots per column (one key bucket; fuller buckets are re-keyed)
constexpr int W3_LSTRIDE = 20;       // words per column in the candidate list: 16 slots + 4 sentinels; 20 * c mod 32 takes
                                     // 8 different multiples of 4 for the 8 columns of a warp: conflict-free 128-bit reads
constexpr int W3_BOX_BYTES = 32 * W3_ROWS * 4;     // one 32-column box
constexpr int W3_BOX_WORDS = W3_BOX_BYTES / 4;

// Geometry for P lanes per column (P = 4: 256 threads, 64 values per thread, up to 128 registers; P = 8: 512 threads,
// 32 values per thread, 64 registers).  A thread owns, of every group of 8 rows, the E = 8 / P consecutive rows
// E*part .. E*part + E - 1: with the 128-byte swizzle (16-byte chunk index XOR row mod 8) the 32 lanes of a warp —
// 32 / P adjacent columns x P parts — then read 32 different banks.
template <int P>
struct W3Geo {
    static constexpr int LANES = P;
    static constexpr int THREADS = W3_COLS * P;
    static constexpr int WARPS = THREADS / 32;
    static constexpr int CW = 32 / P;              // columns per warp
    static constexpr int NV = W3_ROWS / P;         // values per thread
    static constexpr int NKW = NV / 4;             // packed key words per thread
    static constexpr int E = 8 / P;                // consecutive rows per thread in a group of 8
    static constexpr int MW = (NKW + 7) / 8;       // 32-bit words of a thread's row mask
    static constexpr int RT = THREADS / 16;        // rows per trip of the risk sweep (16 lanes share a row)
    static constexpr int NT = W3_ROWS / RT;        // trips
    static constexpr int SLOTS = W3_ROWS;          // keys per column, sentinels included
    static constexpr int KBITS = 7;                // key bits: keys 0..126, 127 = sentinel
    static constexpr int NEXT = 2;                 // buckets above the median's searched for the next value
};

struct W3Ctx {
    int N;
    int part;            // which rows of every 8-row group
    unsigned cmask;      // the lanes of this thread's column
    int lane_base;       // lane of part 0 of this column
};

template <int P>
__device__ __forceinline__ int w3_row(int j, int part) {
    constexpr int E = W3Geo<P>::E;
    return 8 * (j / E) + E * part + (j % E);
}

// The lanes of a column are lane_base + CW * k, k < LANES: sums and reductions over them by xor shuffles
template <class G, class T>
__device__ __forceinline__ T sel_col_sum(const W3Ctx &w, T v) {
    if (G::LANES >= 2) v += __shfl_xor_sync(w.cmask, v, G::CW);
    if (G::LANES >= 4) v += __shfl_xor_sync(w.cmask, v, 2 * G::CW);
    if (G::LANES >= 8) v += __shfl_xor_sync(w.cmask, v, 4 * G::CW);
    return v;
}
template <class G, class T, class Op>
__device__ __forceinline__ T sel_col_reduce(const W3Ctx &w, T v, Op op) {
    if (G::LANES >= 2) v = op(v, __shfl_xor_sync(w.cmask, v, G::CW));
    if (G::LANES >= 4) v = op(v, __shfl_xor_sync(w.cmask, v, 2 * G::CW));
    if (G::LANES >= 8) v = op(v, __shfl_xor_sync(w.cmask, v, 4 * G::CW));
    return v;
}

// ---- cluster / DSMEM / TMA primitives
__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ uint32_t map_to_peer(uint32_t smem_addr, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(smem_addr), "r"(rank));
    return r;
}
// store into the peer CTA's shared memory; the bytes are counted on the peer's mbarrier when they have landed
__device__ __forceinline__ void st_async_peer(uint32_t addr, float v, uint32_t bar) {
    asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.b32 [%0], %1, [%2];" ::"r"(addr), "r"(__float_as_uint(v)), "r"(bar) : "memory");
}
__device__ __forceinline__ void st_async_peer(uint32_t addr, u64 v, uint32_t bar) {
    asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.b64 [%0], %1, [%2];" ::"r"(addr), "l"(v), "r"(bar) : "memory");
}
__device__ __forceinline__ void st_async_peer(uint32_t addr, uint32_t v, uint32_t bar) {
    asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.b32 [%0], %1, [%2];" ::"r"(addr), "r"(v), "r"(bar) : "memory");
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive(u64 *bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// one box of the round: 32 columns x 256 rows starting at (col0, row0) of the [T*N][128] tensor
__device__ __forceinline__ void tma_load_box(void *dst, const CUtensorMap *map, int col0, int row0, u64 *bar) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(
            smem_u32(dst)),
        "l"(map), "r"(col0), "r"(row0), "r"(smem_u32(bar))
        : "memory");
}
__device__ __forceinline__ void tma_prefetch_box(const CUtensorMap *map, int col0, int row0) {
    asm volatile("cp.async.bulk.prefetch.tensor.2d.L2.global.tile [%0, {%1, %2}];" ::"l"(map), "r"(col0), "r"(row0) : "memory");
}

// the exact, slow column centre as a real call: it is reached only by degenerate columns, and inlining its
// call sites would add a sixth to the kernel's code
template <class Tr>
__device__ __noinline__ typename Tr::val_t sel_exact_centre(const typename Tr::in_t *gcol, int gstride, int N, const uint8_t *keep,
                                                            int n_valid, int kind) {
    typename Tr::val_t sk, ku;
    int stw = STW_OK;
    return column_centre_generic<Tr>(gcol, gstride, N, keep, n_valid, kind, false, sk, ku, stw);
}

// A thread's values in the swizzled tile, by value index j: a0 is the shared-memory address of its first row, a1 of
// its second row of the same group of 8 (E = 2 only; the two differ by one 128-byte line and the change of the
// swizzled chunk).  Explicit ld.shared: through a struct of generic pointers the compiler no longer saw the address
// space and emitted generic loads (6 % slower).
__device__ __forceinline__ void lds32(float &v, uint32_t addr) { asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(addr)); }
__device__ __forceinline__ void lds32(int32_t &v, uint32_t addr) { asm volatile("ld.shared.s32 %0, [%1];" : "=r"(v) : "r"(addr)); }
template <int P, class in_t>
struct W3Vals {
    uint32_t a0, a1;
    int part;
    __device__ __forceinline__ in_t operator()(int j) const {
        constexpr int E = W3Geo<P>::E;
        in_t v;
        if (E == 1) lds32(v, a0 + (uint32_t)j * 1024u);
        else lds32(v, ((j & 1) ? a1 : a0) + (uint32_t)(j >> 1) * 1024u);
        return v;
    }
    __device__ __forceinline__ int row(int j) const { return w3_row<P>(j, part); }
};

// ---- keys in registers
// FROM_REGS: the values are passed in x (first build of a round); otherwise they are read from the tile.
// ONLY_VALID: rows whose current key is the sentinel (dropped / padded) keep it.
template <class Tr, class G, bool FIRST, bool FULL, bool ONLY_VALID, bool FROM_REGS, class V>
__device__ __forceinline__ void sel_build_keys(const typename Tr::in_t *x, const V val, unsigned (&kw)[G::NKW], int N,
                                               typename Tr::in_t lo, typename Tr::in_t hi,
                                               const typename WideKey<Tr>::Scale &sc) {
#pragma unroll
    for (int q = 0; q < G::NKW; ++q) {
        unsigned f[4];
        const unsigned old = kw[q];
#pragma unroll
        for (int b = 0; b < 4; ++b) {
            const int j = 4 * q + b;
            const typename Tr::in_t xv = FROM_REGS ? x[j] : val(j);
            f[b] = wide_key<Tr, FIRST, (1 << G::KBITS) - 2>(xv, lo, hi, sc);        // pack4 takes the low byte
            if (!FULL && val.row(j) >= N) f[b] = 127u;
            if (ONLY_VALID && ((old >> (8 * b)) & 0xffu) == 127u) f[b] = 127u;
        }
        kw[q] = pack4(f[0], f[1], f[2], f[3]);
    }
}

template <int NKW>
__device__ __forceinline__ int w3_key_sum(const unsigned (&kw)[NKW]) {
    unsigned a0 = 0, a1 = 0;
#pragma unroll
    for (int q = 0; q < NKW; q += 2) {
        a0 = __dp4a(kw[q], 0x01010101u, a0);
        a1 = __dp4a(kw[q + 1], 0x01010101u, a1);
    }
    return (int)(a0 + a1);
}

// #(key >= cand) among this thread's keys — see count_ge in consensus_wide.cuh for the SWAR identity
template <int NKW>
__device__ __forceinline__ int w3_count_ge(const unsigned (&kw)[NKW], unsigned cand, int sum_keys) {
    const unsigned cb = (128u - cand) * 0x01010101u;
    int a0 = 0, a1 = 0, a2 = 0, a3 = 0;
#pragma unroll
    for (int q = 0; q < NKW; q += 4) {
        a0 = __dp4a((int)__viaddmax_u32(kw[q], cb, 0u), 0x01010101, a0);
        a1 = __dp4a((int)__viaddmax_u32(kw[q + 1], cb, 0u), 0x01010101, a1);
        a2 = __dp4a((int)__viaddmax_u32(kw[q + 2], cb, 0u), 0x01010101, a2);
        a3 = __dp4a((int)__viaddmax_u32(kw[q + 3], cb, 0u), 0x01010101, a3);
    }
    return (int)((unsigned)(sum_keys + 4 * NKW * (int)(128u - cand) - ((a0 + a1) + (a2 + a3))) >> 8);
}

template <class T> __device__ __forceinline__ T w3_never_less();
template <> __device__ __forceinline__ float w3_never_less<float>() { return __int_as_float(0x7f800000); }
template <> __device__ __forceinline__ int32_t w3_never_less<int32_t>() { return 0x7fffffff; }

// Centre of the valid rows of this thread's column (valid = key != sentinel: all rows < N in the first pass, the
// kept rows in the second).  Executed by the four lanes of the column together; every one returns the result.
// src0 / src1: this thread's rows in the tile (w3_value).  gcol / keep serve only the rare exact fallback, which
// walks the column in global memory.  lcw: this column's candidate list (W3_LSTRIDE words).
template <class Tr, class G, bool FULL, class V>
__device__ __forceinline__ typename Tr::val_t sel_centre(const W3Ctx &w, const V val, unsigned (&kw)[G::NKW],
                                                         const typename Tr::in_t *gcol, int gstride, const uint8_t *keep, int n_valid,
                                                        int kind, bool col_const, bool col_fallback,
                                                        typename Tr::in_t const_value, typename Tr::in_t *lcw,
                                                        bool &refined) {
    using in_t = typename Tr::in_t;
    using val_t = typename Tr::val_t;
    using Key = WideKey<Tr>;
    constexpr int NKW = G::NKW, MW = G::MW;
    refined = false;
    if (col_const) return Tr::half_sum((val_t)const_value, (val_t)const_value);
    if (col_fallback) return sel_exact_centre<Tr>(gcol, gstride, w.N, keep, n_valid, kind);
    int klo, khi;
    centre_ranks(kind, n_valid, klo, khi);
    in_t v_lo = const_value, v_hi = const_value;

    for (;;) {
        // ---- bisection over the 7 key bits for rank klo: own 64 keys, then a shuffle sum over the column's lanes
        unsigned prefix = 0;
        int L = 0, LE = n_valid;
        const int sum_keys = w3_key_sum(kw);
#pragma unroll
        for (int bit = G::KBITS - 1; bit >= 0; --bit) {
            const unsigned cand = prefix | (1u << bit);
            const int ge = sel_col_sum<G>(w, w3_count_ge(kw, cand, sum_keys));
            const int lt = G::SLOTS - ge;        // sentinels (127) are never below a pivot
            if (lt <= klo) { prefix = cand; L = lt; }
            else LE = lt;                         // the last rejected pivot is prefix + 1: LE = #(key <= prefix)
        }
        // ---- resolving the bucket.  One bit per value: the byte MSBs of a key word are gathered into a nibble by
        //      one multiply (bits 7, 15, 23, 31 -> 28..31), 16 nibbles make the 64-bit mask of this thread's rows.
        //      w1 = rows whose key is `prefix`.
        const unsigned cA = (128u - prefix) * 0x01010101u;            // byte MSB set <=> key >= prefix
        const unsigned cB = cA - 0x01010101u;                         //              <=> key >= prefix + 1
        unsigned w1[MW] = {};
#pragma unroll
        for (int q = 0; q < NKW; ++q) {
            const unsigned fl = (kw[q] + cA) & ~(kw[q] + cB) & 0x80808080u;
            w1[q >> 3] += ((fl * 0x00204081u) >> 28) << (4 * (q & 7));
        }
        // the two fetch loops below take TWO rows per trip (two independent loads in flight): a trip is a chain of
        // find-first-set, address, load and compare, and the loops are latency-bound
        bool rescue = false;
        in_t thr = const_value;
        if (khi >= LE) {
            // Rank klo is the LAST of its bucket and rank khi the first value above it: v_lo = the bucket's maximum,
            // v_hi = the minimum of the next G::NEXT buckets (w2) — no ranking, no candidate list; an empty w2 (rare)
            // leaves the scan over all valid values above v_lo.
            const unsigned up = prefix + 1u + G::NEXT > 127u ? 127u : prefix + 1u + G::NEXT;
            const unsigned cC = (128u - up) * 0x01010101u;
            unsigned w2[MW] = {};
#pragma unroll
            for (int q = 0; q < NKW; ++q) {
                const unsigned fl = (kw[q] + cB) & ~(kw[q] + cC) & 0x80808080u;
                w2[q >> 3] += ((fl * 0x00204081u) >> 28) << (4 * (q & 7));
            }
            in_t vmx = std::numeric_limits<in_t>::lowest(), vmn = w3_never_less<in_t>();
#pragma unroll
            for (int h = 0; h < MW; ++h) {
                unsigned bits = w1[h];
                while (bits) {
                    const int pos = __ffs(bits) - 1;
                    bits &= bits - 1;
                    const int pos2 = bits ? __ffs(bits) - 1 : pos;       // an idle second slot re-reads the first row
                    bits &= bits - 1;
                    const in_t xa = val(32 * h + pos), xb = val(32 * h + pos2);
                    vmx = vmax(vmx, vmax(xa, xb));
                }
                bits = w2[h];
                while (bits) {
                    const int pos = __ffs(bits) - 1;
                    bits &= bits - 1;
                    const int pos2 = bits ? __ffs(bits) - 1 : pos;
                    bits &= bits - 1;
                    const in_t xa = val(32 * h + pos), xb = val(32 * h + pos2);
                    vmn = vmin(vmn, vmin(xa, xb));
                }
            }
            v_lo = sel_col_reduce<G>(w, vmx, [](in_t a, in_t b) { return vmax(a, b); });
            v_hi = sel_col_reduce<G>(w, vmn, [](in_t a, in_t b) { return vmin(a, b); });
            rescue = v_hi == w3_never_less<in_t>();
            thr = v_lo;
        } else {
            // both order statistics lie in the bucket, which holds m = LE - L values
            const int m = LE - L;
            const int j1 = klo - L, j2 = khi - L;
            if (m <= W3_CAP) {
                // slot offset of this part = candidates of the lower parts
                int mine = 0;
#pragma unroll
                for (int h = 0; h < MW; ++h) mine += __popc(w1[h]);
                int off = 0;
#pragma unroll
                for (int q = 0; q < G::LANES - 1; ++q) {
                    const int cq = __shfl_sync(w.cmask, mine, w.lane_base + G::CW * q);
                    if (q < w.part) off += cq;
                }
                // values of this part's bucket rows -> slots off.., fetched from the tile by row index
                int k = off;
#pragma unroll
                for (int h = 0; h < MW; ++h) {
                    unsigned bits = w1[h];
                    while (bits) {
                        const int pos = __ffs(bits) - 1;
                        bits &= bits - 1;
                        lcw[k++] = val(32 * h + pos);
                    }
                }
                if (w.part == 0) {                // four sentinels behind the last candidate: the ranking reads 4 slots a trip
                    const in_t big = w3_never_less<in_t>();
                    lcw[m] = big; lcw[m + 1] = big; lcw[m + 2] = big; lcw[m + 3] = big;
                }
                __syncwarp(w.cmask);
                // rank this part's share of the slots: lt(v) = #{u < v}; rank j's value has the largest lt <= j
                int b1 = -1, b2 = -1;
                for (int a = w.part; a < m; a += G::LANES) {
                    const in_t v = lcw[a];
                    int rr = 0;
#pragma unroll 1
                    for (int b = 0; b < m; b += 4) {
                        const int4 raw = *reinterpret_cast<const int4 *>(lcw + b);
                        in_t u[4];
                        memcpy(u, &raw, 16);
#pragma unroll
                        for (int s4 = 0; s4 < 4; ++s4) count_lt(rr, u[s4], v);
                    }
                    const int code = (rr << 8) | a;       // largest (rank, slot) code with rank <= j1 / j2
                    if (rr <= j1 && code > b1) b1 = code;
                    if (rr <= j2 && code > b2) b2 = code;
                }
                auto imax = [](int a, int b) { return a > b ? a : b; };
                b1 = sel_col_reduce<G>(w, b1, imax);
                b2 = sel_col_reduce<G>(w, b2, imax);
                v_lo = lcw[b1 & 0xff];
                v_hi = lcw[b2 & 0xff];
                __syncwarp(w.cmask);              // the list may be overwritten by the next selection
            } else {
                // crowded bucket (ties, clusters): its value range decides between "all one value" and a finer re-keying
                uint32_t kmin = 0xffffffffu, kmax = 0u;
#pragma unroll
                for (int h = 0; h < MW; ++h) {
                    unsigned bits = w1[h];
                    while (bits) {
                        const int pos = __ffs(bits) - 1;
                        bits &= bits - 1;
                        const uint32_t kx = okey(val(32 * h + pos));
                        kmin = kx < kmin ? kx : kmin;
                        kmax = kx > kmax ? kx : kmax;
                    }
                }
                kmin = sel_col_reduce<G>(w, kmin, [](uint32_t a, uint32_t b) { return a < b ? a : b; });
                kmax = sel_col_reduce<G>(w, kmax, [](uint32_t a, uint32_t b) { return a > b ? a : b; });
                in_t a, b;
                unkey(kmin, a);
                unkey(kmax, b);
                typename Key::Scale sc;
                if (!(b > a)) {                   // the whole bucket is one value
                    v_lo = a;
                    v_hi = a;
                } else if (!Key::make(a, b, sc)) {
                    return sel_exact_centre<Tr>(gcol, gstride, w.N, keep, n_valid, kind);
                } else {
                    sel_build_keys<Tr, G, false, FULL, true, false>(nullptr, val, kw, w.N, a, b, sc);
                    refined = true;
                    continue;
                }
            }
        }
        if (rescue) {                             // rank khi lies above the window: smallest valid value above it
            uint32_t best = 0xffffffffu;
#pragma unroll
            for (int q = 0; q < NKW; ++q) {
#pragma unroll
                for (int b = 0; b < 4; ++b) {
                    const bool valid = ((kw[q] >> (8 * b)) & 0xffu) != 127u;
                    const in_t xv = val(4 * q + b);
                    if (valid && xv > thr) { const uint32_t kx = okey(xv); best = kx < best ? kx : best; }
                }
            }
            best = sel_col_reduce<G>(w, best, [](uint32_t a, uint32_t b) { return a < b ? a : b; });
            unkey(best, v_hi);
        }
        break;
    }
    return Tr::half_sum((val_t)v_lo, (val_t)v_hi);
}

struct W3Smem {
    // offsets in bytes from the 1024-aligned base
    size_t tile, list, c1, theirs, xflag, lflag, qr, wq, wi, ww, maskw, patch, misc, keep, bar, total;
};
__host__ __device__ inline W3Smem w3_smem_layout(size_t vs) {
    W3Smem s{};
    size_t off = 0;
    auto take = [&](size_t bytes, size_t align) { off = (off + align - 1) & ~(align - 1); const size_t at = off; off += bytes; return at; };
    s.tile = take(2 * W3_BOX_BYTES, 1024);
    s.list = take((size_t)W3_COLS * W3_LSTRIDE * 4, 16);
    s.c1 = take((size_t)W3_COLS * vs, 16);                // first-pass centre of the CTA's columns
    s.theirs = take((size_t)2 * W3_ROWS * vs, 16);        // [parity][row]: the peer CTA's 64-column partial risks
    s.xflag = take(2 * 16 * 4, 16);                       // [parity][warp]: the peer's input-range flags
    s.lflag = take(2 * 16 * 4, 16);                       // [parity][warp]: this CTA's
    s.qr = take((size_t)W3_ROWS * vs, 16);
    s.wq = take((size_t)W2_RCAP * vs, 16);
    s.wi = take((size_t)W2_RCAP * 4, 16);
    s.ww = take(64, 16);
    s.maskw = take(16 * 4, 16);
    s.patch = take(W3_ROWS, 16);                          // [part][value]: 0x7f where the row is dropped
    s.misc = take(16, 16);                                // stw[2], sel
    s.keep = take(W3_ROWS, 16);
    s.bar = take(48, 16);                                 // full, xbar[2], (spare), arrival counter
    s.total = off;
    return s;
}

// Resident CTAs per SM the compiler makes room for.  P = 4 lanes per column: three CTAs of 256 threads at 80 registers
// for fp32, whose layout stays under 75 KB (228 KB per SM, 1 KB reserved per CTA), two at 128 registers for the
// fixed form with its 64-bit risks.  P = 8: two CTAs of 512 threads at 64 registers.
__host__ __device__ constexpr int w3_ctas_per_sm(int P, size_t vs) { return P == 8 ? 2 : (vs == 4 ? 3 : 2); }

template <class Tr, int P, bool FULL>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(W3Geo<P>::THREADS, w3_ctas_per_sm(P, sizeof(typename Tr::acc_t)))
consensus_wide3_kernel(const Params p, const __grid_constant__ CUtensorMap tmap) {
    using in_t = typename Tr::in_t;
    using val_t = typename Tr::val_t;
    using acc_t = typename Tr::acc_t;
    using Key = WideKey<Tr>;
    using G = W3Geo<P>;
    constexpr int NV = G::NV, NKW = G::NKW, E = G::E, CW = G::CW, RT = G::RT, NT = G::NT;
    extern __shared__ __align__(1024) unsigned char smem_w3[];     // the 128-byte swizzle wants the boxes 1024-aligned
    unsigned char *const smem_raw = smem_w3;
    const W3Smem S = w3_smem_layout(sizeof(acc_t));
    const int N = FULL ? W3_ROWS : p.N, R = p.R;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t rank = cluster_ctarank(), peer = rank ^ 1u;
    const i64 cluster_id = blockIdx.x >> 1, n_clusters = gridDim.x >> 1;

    in_t *tile_s = reinterpret_cast<in_t *>(smem_raw + S.tile);
    in_t *list = reinterpret_cast<in_t *>(smem_raw + S.list);
    val_t *c1_s = reinterpret_cast<val_t *>(smem_raw + S.c1);
    acc_t *theirs_s = reinterpret_cast<acc_t *>(smem_raw + S.theirs);
    uint32_t *xflag_s = reinterpret_cast<uint32_t *>(smem_raw + S.xflag);
    uint32_t *lflag_s = reinterpret_cast<uint32_t *>(smem_raw + S.lflag);
    acc_t *qr_s = reinterpret_cast<acc_t *>(smem_raw + S.qr);
    acc_t *wq = reinterpret_cast<acc_t *>(smem_raw + S.wq);
    int *wi = reinterpret_cast<int *>(smem_raw + S.wi);
    unsigned *ww = reinterpret_cast<unsigned *>(smem_raw + S.ww);
    uint32_t *maskw = reinterpret_cast<uint32_t *>(smem_raw + S.maskw);
    uint8_t *patch_s = smem_raw + S.patch;
    int *stw_s = reinterpret_cast<int *>(smem_raw + S.misc);        // [2] by round parity
    int *sel_s = stw_s + 2;
    uint8_t *keep_s = smem_raw + S.keep;
    u64 *full_bar = reinterpret_cast<u64 *>(smem_raw + S.bar);
    u64 *xbar = full_bar + 1;                                        // [2] by round parity
    unsigned *done_s = reinterpret_cast<unsigned *>(full_bar + 4);   // warps that have made their last read of the tile

    // this thread's column unit: column c of the CTA's 64, rows 8k + E*part + {0 .. E-1}
    const int cq = lane % CW;
    W3Ctx w;
    w.N = N;
    w.part = lane / CW;
    w.lane_base = cq;
    w.cmask = (P == 8 ? 0x11111111u : 0x01010101u) << cq;
    const int part = w.part;
    const int c = warp * CW + cq;                 // column within the CTA
    const int gc = (int)rank * W3_COLS + c;       // column within the round
    in_t *lcw = list + c * W3_LSTRIDE;
    // tile address of this thread's first row, and the step to its next row inside a group of 8 (128-byte swizzle: the
    // 16-byte chunk index of a row is XOR-ed with the row number mod 8, which is E*part + e for every row read here)
    const int c32 = c & 31;
    const in_t *src0 = tile_s + (c >> 5) * W3_BOX_WORDS + (E * part) * 32 + ((((c32 >> 2) ^ (E * part)) << 2) | (c32 & 3));
    const in_t *src1 = E == 1 ? src0 : tile_s + (c >> 5) * W3_BOX_WORDS + (E * part + 1) * 32 + ((((c32 >> 2) ^ (E * part + 1)) << 2) | (c32 & 3));
    const W3Vals<P, in_t> val{smem_u32(src0), smem_u32(src1), part};
    // this thread's row unit in the risk sweep: 4-column chunk rh of row RT*t + rg (t = 0 .. NT-1); after the shuffle
    // tree the thread holds the total of row RT * (rh >> (4 - log2 NT)) + rg
    const int rh = lane & 15, rg = warp * 2 + (lane >> 4);
    const in_t *rsrc = tile_s + (rh >> 3) * W3_BOX_WORDS + rg * 32 + (((rh & 7) ^ (rg & 7)) << 2);
    const int my_row = RT * (NT == 16 ? rh : rh >> 1) + rg;
    const bool row_owner = NT == 16 || (rh & 1) == 0;     // NT = 8: two lanes end up with the same row total

    const uint32_t tile_bytes = 2 * W3_BOX_BYTES;
    constexpr uint32_t XBYTES = W3_ROWS * sizeof(acc_t) + (Tr::fixed ? G::WARPS * 4 : 0);
    if (tid == 0) {
        if (smem_u32(tile_s) & 1023u) __trap();      // the swizzle pattern is a function of the address bits
        mbar_init(full_bar, 1);
        *done_s = 0u;
        mbar_init(xbar, 1);
        mbar_init(xbar + 1, 1);
        stw_s[0] = STW_OK;
        mbar_expect_tx(full_bar, tile_bytes);
        const int row0 = (int)(cluster_id * p.N);
        tma_load_box(tile_s, &tmap, (int)rank * W3_COLS, row0, full_bar);
        tma_load_box(tile_s + W3_BOX_WORDS, &tmap, (int)rank * W3_COLS + 32, row0, full_bar);
    }
    __syncthreads();
    // the peer must be running (its barriers initialised) before anything is sent into its shared memory
    cluster_sync_all();

    constexpr int NS = W3_ROWS / 32;
    const bool shape2 = second_pass_shape_panic(p);
    const bool high = p.tie == SVOC_TIE_HIGH_INDEX_KEPT;
    const uint32_t theirs_remote = map_to_peer(smem_u32(theirs_s), peer);
    const uint32_t xflag_remote = map_to_peer(smem_u32(xflag_s), peer);
    const uint32_t xbar_remote = map_to_peer(smem_u32(xbar), peer);

    uint32_t it = 0;
    for (i64 r = cluster_id; r < p.T; r += n_clusters, ++it) {
        const uint32_t par = it & 1u;
        const bool has_next = r + n_clusters < p.T;
        const in_t *gcol = reinterpret_cast<const in_t *>(p.pred) + (size_t)r * p.N * W3_D + gc;
        if (tid == 0) {
            stw_s[par ^ 1u] = STW_OK;                              // the next round's status word
            mbar_expect_tx(xbar + par, XBYTES);                    // this round's exchange
            if (has_next) {                                        // pull the next round towards L2 meanwhile
                const int row0 = (int)((r + n_clusters) * p.N);
                tma_prefetch_box(&tmap, (int)rank * W3_COLS, row0);
                tma_prefetch_box(&tmap, (int)rank * W3_COLS + 32, row0);
            }
        }
        // ---- phase 1a: the tile has landed; key scale from a 64-row sample (rows 0..63: the first 16 values of
        //      every part), then the keys
        mbar_wait_sleepy(full_bar, par);
        bool col_const = false, col_fb = false, refined1 = false;
        in_t lo = std::numeric_limits<in_t>::max(), hi = std::numeric_limits<in_t>::lowest();
        typename Key::Scale sc{};
        unsigned kw[NKW];
        {
            in_t x[NV];
#pragma unroll
            for (int j = 0; j < NV; ++j) x[j] = val(j);
#pragma unroll
            for (int j = 0; j < 64 / P; ++j) {       // rows 0..63
                lo = vmin(lo, x[j]);
                hi = vmax(hi, x[j]);
            }
            lo = sel_col_reduce<G>(w, lo, [](in_t a, in_t b) { return vmin(a, b); });
            hi = sel_col_reduce<G>(w, hi, [](in_t a, in_t b) { return vmax(a, b); });
            if (!(hi > lo)) {                         // degenerate sample: exact sweep, which also detects a constant column
#pragma unroll
                for (int j = 0; j < NV; ++j) {
                    if (FULL || w3_row<P>(j, part) < N) {
                        lo = vmin(lo, x[j]);
                        hi = vmax(hi, x[j]);
                    }
                }
                lo = sel_col_reduce<G>(w, lo, [](in_t a, in_t b) { return vmin(a, b); });
                hi = sel_col_reduce<G>(w, hi, [](in_t a, in_t b) { return vmax(a, b); });
            }
            col_const = !(hi > lo);
            col_fb = !col_const && !Key::make(lo, hi, sc);
#pragma unroll
            for (int q = 0; q < NKW; ++q) kw[q] = 0u;
            if (!col_const && !col_fb) sel_build_keys<Tr, G, true, FULL, false, true>(x, val, kw, N, lo, hi, sc);
        }
        // ---- the two passes share ONE inlined copy of the centre code: pass 0 = first-pass centre, risks, row
        //      selection, key patching; pass 1 = second-pass centre
#pragma unroll 1
        for (int pass = 0;; ++pass) {
            bool refined = false;
            const val_t ctr = sel_centre<Tr, G, FULL>(w, val, kw, gcol, W3_D, pass ? keep_s : nullptr, pass ? R : N,
                                                  pass ? p.centre2 : p.centre, col_const, col_fb, lo, lcw, refined);
            if (pass) {
                if (part == 0) reinterpret_cast<val_t *>(p.consensus)[(size_t)r * W3_D + gc] = ctr;
                break;
            }
            refined1 = refined;
            if (part == 0) {
                c1_s[c] = ctr;
                if (p.essence1) reinterpret_cast<val_t *>(p.essence1)[(size_t)r * W3_D + gc] = ctr;
            }
            __syncthreads();

            // ---- phase 2: quadratic risk.  Row-wise sweep of the tile: 16 lanes share a row (a quarter-warp reads
            //      one 128-byte line), a thread visits rows RT*t + rg, t = 0 .. NT-1, and folds (x - e1)^2 over its four
            //      columns; the NT partials of a thread and the 16 lanes of a row are then folded together by a
            //      transposing shuffle tree (xor 8, 4, 2, 1: fixed order), which leaves every thread with the total
            //      of ONE row over the CTA's 64 columns (NT = 8: pairs of lanes hold the same row; one of them owns it).
            acc_t own;
            {
                val_t e4[4];
#pragma unroll
                for (int k = 0; k < 4; ++k) e4[k] = c1_s[4 * rh + k];
                acc_t v[NT];
                bool bad = false;
#pragma unroll
                for (int t = 0; t < NT; ++t) {
                    const int4 raw = *reinterpret_cast<const int4 *>(rsrc + t * RT * 32);
                    in_t xr[4];
                    memcpy(xr, &raw, 16);
                    acc_t q = 0;
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        if constexpr (Tr::fixed) {
                            const bool out = p.constrained ? ((xr[k] < 0) | ((i64)xr[k] > WSAD))
                                                           : (((i64)xr[k] > ENGINE_RANGE) | ((i64)xr[k] < -ENGINE_RANGE));
                            bad |= out & (FULL || RT * t + rg < N);
                        }
                        q += Tr::dev(xr[k], e4[k]);
                    }
                    v[t] = q;
                }
                // halving steps: lane bit 3 keeps the upper / lower half of the trips, bit 2 of what is left, ...
                int lanebit = 8;
#pragma unroll
                for (int n = NT / 2; n >= 1; n /= 2, lanebit >>= 1) {
                    const bool up = rh & lanebit;
#pragma unroll
                    for (int i = 0; i < n; ++i) {
                        const acc_t keep = up ? v[i + n] : v[i], send = up ? v[i] : v[i + n];
                        v[i] = keep + __shfl_xor_sync(0xffffffffu, send, lanebit);
                    }
                }
                own = v[0];
                if (NT == 8) own += __shfl_xor_sync(0xffffffffu, own, 1);     // both lanes of the pair: the same sum
                // to the peer: this row's partial (and, fixed form, the warp's input-range flag)
                if (row_owner)
                    st_async_peer(theirs_remote + (uint32_t)((par * W3_ROWS + my_row) * sizeof(acc_t)), own, xbar_remote + par * 8u);
                if constexpr (Tr::fixed) {
                    const uint32_t anybad = __any_sync(0xffffffffu, bad) ? 1u : 0u;
                    if (lane == 0) {
                        lflag_s[par * G::WARPS + warp] = anybad;
                        st_async_peer(xflag_remote + (par * G::WARPS + warp) * 4u, anybad, xbar_remote + par * 8u);
                    }
                }
            }
            mbar_wait_sleepy(xbar + par, (it >> 1) & 1u);
            {
                const acc_t other = theirs_s[par * W3_ROWS + my_row];
                const acc_t q = rank == 0 ? own + other : other + own;        // partial(rank 0) + partial(rank 1) on both CTAs
                if (row_owner) {
                    qr_s[my_row] = q;
                    if (rank == 0 && p.qr && my_row < N) reinterpret_cast<acc_t *>(p.qr)[(size_t)r * p.N + my_row] = q;
                }
            }
            __syncthreads();

            // ---- phase 3: which rows are kept — warp 0 (w2_select_rows) while warp 1 does the first-pass reliability;
            //      exact all-pairs ranking when the selection gives up on this round's risks.  Both CTAs of the pair
            //      compute the same thing from the same risks; rank 0 writes the per-round outputs.
            if (warp == 0) {
                unsigned words[NS];
                acc_t qrow[NS];
#pragma unroll
                for (int j = 0; j < NS; ++j) qrow[j] = qr_s[j * 32 + lane];
                const bool fast = w2_select_rows<acc_t, NS>(qrow, N, R, high, wq, wi, ww, words);
                if (fast) {
                    unsigned mw = 0;
#pragma unroll
                    for (int j = 0; j < NS; ++j) mw = lane == j ? words[j] : mw;
                    if (lane < NS) maskw[lane] = mw;
                }
                if (lane == 0) *sel_s = fast ? 1 : 0;
            } else if (warp == 1) {
                acc_t s1 = 0;
#pragma unroll
                for (int j = 0; j < NS; ++j)
                    if (j * 32 + lane < N) s1 += qr_s[j * 32 + lane];
                for (int o = 16; o; o >>= 1) s1 += __shfl_xor_sync(0xffffffffu, s1, o);
                if (lane == 0) {
                    int stw = STW_OK;
                    if constexpr (Tr::fixed) {
                        uint32_t anybad = 0;
#pragma unroll
                        for (int k = 0; k < G::WARPS; ++k) anybad |= lflag_s[par * G::WARPS + k] | xflag_s[par * G::WARPS + k];
                        if (anybad) stw = p.constrained ? STW_INTERVAL : STW_ENGINE_RANGE;
                        const i64 rel1 = reliability_fixed(s1 / (u64)N, p, STW_SQRT_REL1, stw);
                        if (rel1 < 0 || rel1 > WSAD) stw = min(stw, STW_REL1_RANGE);
                        if (rank == 0) reinterpret_cast<i64 *>(p.rel1)[r] = rel1;
                    } else {
                        const val_t rel1 = reliability_float<val_t>(s1 / (val_t)N, p);
                        if (!(rel1 >= 0 && rel1 <= 1)) stw = min(stw, STW_REL1_RANGE);
                        if (rank == 0) reinterpret_cast<val_t *>(p.rel1)[r] = rel1;
                    }
                    if (stw != STW_OK) atomicMin(stw_s + par, stw);
                }
            }
            __syncthreads();
            if (!*sel_s) {
                // exact ranking of row `tid` under (risk asc, index desc | asc): rare
                if (tid < W3_ROWS) {
                    const acc_t mineq = qr_s[tid];
                    int rk = 0;
                    for (int j = 0; j < N; ++j) {
                        const acc_t q = qr_s[j];
                        rk += (q < mineq) | ((q == mineq) & (high ? j > tid : j < tid));
                    }
                    const unsigned bal = __ballot_sync(0xffffffffu, tid < N && rk < R);
                    if (lane == 0) maskw[warp] = bal;
                }
                __syncthreads();
            }
            // every thread turns the mask bit of row `tid` into the byte forms the column units read: keep_s[row]
            // (0 / 1, exact fallbacks) and patch_s[part][value] (0x7f where dropped: OR-ed into the keys)
            if (tid < W3_ROWS) {
                const bool kept = (maskw[tid >> 5] >> (tid & 31)) & 1u;
                keep_s[tid] = kept ? 1 : 0;
                patch_s[((tid & 7) / E) * NV + E * (tid >> 3) + (tid % E)] = kept ? 0 : 0x7f;
            }
            __syncthreads();                 // keep_s / patch_s are complete
            // ---- phase 4 set-up: keys of the kept rows only; the loop then runs the second-pass centre
            if (shape2) {
                if (part == 0) reinterpret_cast<val_t *>(p.consensus)[(size_t)r * W3_D + gc] = 0;
                break;
            }
            if (!col_const && !col_fb) {
                if (refined1) {
                    // the first pass re-keyed the column: back to the first-pass scale
#pragma unroll
                    for (int q = 0; q < NKW; ++q) kw[q] = 0u;
                    sel_build_keys<Tr, G, true, FULL, false, false>(nullptr, val, kw, N, lo, hi, sc);
                }
                // dropped rows -> sentinel key (keys are <= 127, so OR-ing 0x7f gives 127)
#pragma unroll
                for (int q4 = 0; q4 < NKW / 4; ++q4) {
                    const uint4 d = *reinterpret_cast<const uint4 *>(patch_s + part * NV + 16 * q4);
                    kw[4 * q4] |= d.x; kw[4 * q4 + 1] |= d.y; kw[4 * q4 + 2] |= d.z; kw[4 * q4 + 3] |= d.w;
                }
            }
        }
        // ---- this warp has made its last read of the tile: the LAST warp to get here copies the next round in
        //      (an arrival counter; nobody waits — an earlier version had warp 0 wait on an "empty" mbarrier, and with
        //      lane 0 alone waiting while its siblings span on the full barrier that only lane 0's copy completes,
        //      the kernel hung whenever another warp was last: divergent spin loops inside one warp)
        __syncwarp();
        if (lane == 0 && has_next) {
            if (atomicAdd(done_s, 1u) == G::WARPS - 1) {
                *done_s = 0u;
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                mbar_expect_tx(full_bar, tile_bytes);
                const int row0 = (int)((r + n_clusters) * p.N);
                tma_load_box(tile_s, &tmap, (int)rank * W3_COLS, row0, full_bar);
                tma_load_box(tile_s + W3_BOX_WORDS, &tmap, (int)rank * W3_COLS + 32, row0, full_bar);
            }
        }
        __syncwarp();
        // ---- second-pass reliability, mask and status: warp 0, at the end of the round so that it delays no barrier
        if (warp == 0) {
            acc_t s2 = 0;
#pragma unroll
            for (int j = 0; j < NS; ++j)
                if (j * 32 + lane < N && ((maskw[j] >> lane) & 1u)) s2 += qr_s[j * 32 + lane];
            if (rank == 0 && p.reliable && lane < p.words) p.reliable[(size_t)r * p.words + lane] = maskw[lane];
            for (int o = 16; o; o >>= 1) s2 += __shfl_xor_sync(0xffffffffu, s2, o);
            if (lane == 0) {
                int stw = STW_OK;
                if constexpr (Tr::fixed) {
                    i64 rel2 = 0;
                    if (shape2) stw = min(stw, STW_SECOND_SHAPE);
                    else {
                        rel2 = reliability_fixed(s2 / (u64)R, p, STW_SQRT_REL2, stw);
                        if (rel2 < 0 || rel2 > WSAD) stw = min(stw, STW_REL2_RANGE);
                    }
                    if (rank == 0) reinterpret_cast<i64 *>(p.rel2)[r] = rel2;
                } else {
                    val_t rel2 = 0;
                    if (shape2) stw = min(stw, STW_SECOND_SHAPE);
                    else {
                        rel2 = reliability_float<val_t>(s2 / (val_t)R, p);
                        if (!(rel2 >= 0 && rel2 <= 1)) stw = min(stw, STW_REL2_RANGE);
                    }
                    if (rank == 0) reinterpret_cast<val_t *>(p.rel2)[r] = rel2;
                }
                if (stw != STW_OK) atomicMin(stw_s + par, stw);
                if (rank == 0) {         // nothing later contributes to the status
                    const int sw = stw_s[par];
                    p.status[r] = (sw == STW_OK) ? SVOC_ST_OK : (sw & 0xff);
                }
            }
        }
    }
}

// host side: the tensor map over the whole batch, [T*N rows][128 columns] of 32-bit words
typedef CUresult (*W3EncodeFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                               const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                               CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
W3EncodeFn w3_encode_fn();

template <class Tr, int P>
int launch_wide3_p(const Params &p, const Plan &pl, cudaStream_t st) {
    W3EncodeFn enc = w3_encode_fn();
    if (!enc) return fail("cuTensorMapEncodeTiled is not available from this driver");
    CUtensorMap tmap;
    const cuuint64_t dims[2] = {(cuuint64_t)W3_D, (cuuint64_t)p.T * (cuuint64_t)p.N};
    const cuuint64_t strides[1] = {(cuuint64_t)W3_D * 4};
    const cuuint32_t box[2] = {32, (cuuint32_t)W3_ROWS};
    const cuuint32_t estr[2] = {1, 1};
    const CUresult cr = enc(&tmap, CU_TENSOR_MAP_DATA_TYPE_INT32, 2, const_cast<void *>(p.pred), dims, strides, box, estr,
                            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (cr != CUDA_SUCCESS) return fail("cuTensorMapEncodeTiled failed with code " + std::to_string((int)cr));
    void (*kern)(const Params, const CUtensorMap) = (p.N == W3_ROWS) ? consensus_wide3_kernel<Tr, P, true> : consensus_wide3_kernel<Tr, P, false>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl.smem);
    if (e != cudaSuccess) return fail(std::string("cudaFuncSetAttribute: ") + cudaGetErrorString(e));
    // Three CTAs of 75 KB only fit an SM with the largest shared-memory carveout; left to its default the driver
    // sometimes configured less (two resident CTAs, a second wave for the rest of the grid: 14.2 instead of 15.0 M
    // rounds/s, box by box).  Ask for the carveout, then size the persistent grid by what is really resident.
    e = cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, (int)cudaSharedmemCarveoutMaxShared);
    if (e != cudaSuccess) return fail(std::string("cudaFuncSetAttribute(carveout): ") + cudaGetErrorString(e));
    int resident = 0;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&resident, kern, W3Geo<P>::THREADS, pl.smem);
    if (e != cudaSuccess || resident < 1) return fail(std::string("wide3 occupancy query: ") + cudaGetErrorString(e));
    int sms = 0, dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    long long grid = pl.grid;
    const long long fit = 2 * (((long long)sms * resident) / 2);          // whole clusters
    if (fit >= 2 && grid > fit) grid = fit;
    kern<<<(unsigned)grid, W3Geo<P>::THREADS, pl.smem, st>>>(p, tmap);
    e = cudaGetLastError();
    if (e != cudaSuccess) return fail(std::string("kernel launch: ") + cudaGetErrorString(e));
    count_launch();
    return 0;
}

template <class Tr>
int launch_wide3(const Params &p, const Plan &pl, cudaStream_t st) {
    return pl.wP == 8 ? launch_wide3_p<Tr, 8>(p, pl, st) : launch_wide3_p<Tr, 4>(p, pl, st);
}

}  // namespace svoc
