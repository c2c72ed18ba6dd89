// consensus_wide2.cuh — the `wide2` variant of the fused consensus round: 64 < N <= 512 oracles, 32-bit
// elements (fp32 / int32), rounds aligned to 16 bytes.  This is the kernel config 4 (N=256, D=128, fp32)
// runs.  It keeps the selection method of `wide` (monotone 7-bit keys, SWAR bisection, exact ranking of a
// small window) and changes the work decomposition after profiling it:
//   * the P row-parts of a column sit in the SAME warp (lane = part * (32/P) + column-in-warp), so partial
//     counts, candidate offsets and rank results are combined with a few shuffles: the bisection has no
//     shared-memory traffic, no atomics and no barriers;
//   * the round is not staged in shared memory.  Its sweeps read global memory (a warp request covers
//     P rows x 32 contiguous bytes = whole sectors): the first touch comes from HBM, the later ones from
//     L2, and the CTA's next round is prefetched into L2 (cp.async.bulk.prefetch.L2).  Without the
//     128 KB tile two CTAs — two rounds — share an SM and fill each other's stalls;
//   * the key scale comes from a 64-row sample of the column (values outside it clamp to the end
//     buckets, which keeps keys monotone), saving one sweep.
//   * which rows are kept is found by ONE warp from the shared risk array (7-bit keys of the
//     risks in registers, SWAR bisection with warp reductions, exact ranking of the cut bucket) while the
//     other warps wait at the next barrier; all-pairs counting over all warps is the fallback.
// A CTA-wide barrier is needed only where rows meet columns (3 per round; 5 when the all-pairs ranking runs).
// Reference semantics: see consensus_kernel.cuh (contract.cairo:442-503 etc.).
#pragma once
#include <type_traits>

#include "consensus_wide.cuh"

namespace svoc {

#ifndef W2_THREADS_N
#define W2_THREADS_N 512
#endif
constexpr int W2_THREADS = W2_THREADS_N;
constexpr int W2_CTAS_PER_SM = 1024 / W2_THREADS;
// 64-bit elements: one CTA per SM (128 registers) — at 64 registers the double / int64 arithmetic spills
template <class Tr>
constexpr int w2_ctas_per_sm() { return sizeof(typename Tr::in_t) == 8 ? 1 : W2_CTAS_PER_SM; }
#ifndef W2_CAP_N
#define W2_CAP_N 32
#endif
#ifndef W2_SAMPLE
#define W2_SAMPLE 64
#endif
constexpr int W2_CAP = W2_CAP_N;
constexpr int W2_RCAP = 32;          // rows of the cut bucket one warp ranks exactly (one per lane)
#ifndef W2_CHUNK
#define W2_CHUNK 2
#endif
#ifndef W2_RANKW
#define W2_RANKW 4
#endif
#ifndef W2_SELECT_FAST
#define W2_SELECT_FAST 1
#endif

struct W2Ctx {
    int N, D, RP;          // rows, dimensions, rows per part (multiple of 32)
    int c, part;           // this thread's column and row part
    unsigned cmask;        // the lanes that share this thread's column
    int sh1, sh2, sh3;     // xor distances that walk over those lanes (0 = unused)
    uint8_t *mykeys;       // keysT + c*CS + part*RP
};

template <class T>
__device__ __forceinline__ T col_sum(const W2Ctx &w, T v) {
    if (w.sh1) v += __shfl_xor_sync(w.cmask, v, w.sh1);
    if (w.sh2) v += __shfl_xor_sync(w.cmask, v, w.sh2);
    if (w.sh3) v += __shfl_xor_sync(w.cmask, v, w.sh3);
    return v;
}
template <class T, class Op>
__device__ __forceinline__ T col_reduce(const W2Ctx &w, T v, Op op) {
    if (w.sh1) v = op(v, __shfl_xor_sync(w.cmask, v, w.sh1));
    if (w.sh2) v = op(v, __shfl_xor_sync(w.cmask, v, w.sh2));
    if (w.sh3) v = op(v, __shfl_xor_sync(w.cmask, v, w.sh3));
    return v;
}

template <class Tr, bool FIRST, int DT, int RPT>
__device__ __forceinline__ void w2_build_keys(const W2Ctx &w, const typename Tr::in_t *tile, const uint8_t *keep,
                                              typename Tr::in_t lo, typename Tr::in_t hi,
                                              const typename WideKey<Tr>::Scale &sc) {
    const int WD = DT ? DT : w.D;
    const int WRP = RPT ? RPT : w.RP;
    const int row_lo = w.part * WRP;
    const typename Tr::in_t *ptr = tile + row_lo * WD + w.c;
#pragma unroll 2
    for (int g = 0; g < WRP; g += 16) {
        unsigned f[16];
        if (!keep && row_lo + g + 16 <= w.N) {
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
        *reinterpret_cast<uint4 *>(w.mykeys + g) =
            make_uint4(pack4(f[0], f[1], f[2], f[3]), pack4(f[4], f[5], f[6], f[7]), pack4(f[8], f[9], f[10], f[11]),
                       pack4(f[12], f[13], f[14], f[15]));
        ptr += 16 * WD;
    }
}

// r += (u < v) as a compare and a predicated add (the compiler's own choice is an add and a predicated move
// on top of the compare)
__device__ __forceinline__ void count_lt(int &r, float u, float v) {
    asm("{\n\t.reg .pred p;\n\tsetp.lt.f32 p, %1, %2;\n\t@p add.s32 %0, %0, 1;\n\t}" : "+r"(r) : "f"(u), "f"(v));
}
__device__ __forceinline__ void count_lt(int &r, double u, double v) {
    asm("{\n\t.reg .pred p;\n\tsetp.lt.f64 p, %1, %2;\n\t@p add.s32 %0, %0, 1;\n\t}" : "+r"(r) : "d"(u), "d"(v));
}
__device__ __forceinline__ void count_lt(int &r, i64 u, i64 v) {
    asm("{\n\t.reg .pred p;\n\tsetp.lt.s64 p, %1, %2;\n\t@p add.s32 %0, %0, 1;\n\t}" : "+r"(r) : "l"(u), "l"(v));
}
__device__ __forceinline__ void count_lt(int &r, int32_t u, int32_t v) {
    asm("{\n\t.reg .pred p;\n\tsetp.lt.s32 p, %1, %2;\n\t@p add.s32 %0, %0, 1;\n\t}" : "+r"(r) : "r"(u), "r"(v));
}

// Centre of the flagged rows of this thread's column (all rows when keep == nullptr).  Executed by the P
// lanes of the column together (shuffles under w.cmask); every one of them returns the result.
template <class Tr, int DT, int PT, int RPT>
__device__ __forceinline__ typename Tr::val_t w2_centre(const W2Ctx &w, const typename Tr::in_t *tile,
                                                        const uint8_t *keep, int n_valid, int kind, bool col_const,
                                                        bool col_fallback, typename Tr::in_t const_value,
                                                        typename Tr::in_t *list, int P_rt, bool &refined) {
    using in_t = typename Tr::in_t;
    using val_t = typename Tr::val_t;
    using Key = WideKey<Tr>;
    using key_t = decltype(okey(in_t()));          // order-preserving unsigned image of an element
    const int WD = DT ? DT : w.D;
    const int WP = PT ? PT : P_rt;
    const int WRP = RPT ? RPT : w.RP;
    refined = false;
    const int row_lo = w.part * WRP, row_hi = min(w.N, row_lo + WRP);
    if (kind == SVOC_CENTRE_MEAN) {
        val_t s = 0;
        for (int i = row_lo; i < row_hi; ++i)
            if (!keep || keep[i]) s += (val_t)tile[i * WD + w.c];
        return Tr::mean(col_sum(w, s), n_valid);     // integer sums are exact; float sums differ only in order
    }
    if (col_const) return Tr::half_sum((val_t)const_value, (val_t)const_value);
    if (col_fallback) {
        val_t sk, ku;
        int stw = STW_OK;
        return column_centre_generic<Tr>(tile + w.c, WD, w.N, keep, n_valid, kind, false, sk, ku, stw);
    }
    int klo, khi;
    centre_ranks(kind, n_valid, klo, khi);
    const int slots = WP * WRP;
    in_t *lc = list + w.c;                        // this column's candidate values: slot k at lc[k * WD]
    in_t v_lo = const_value, v_hi = const_value;

    for (;;) {
        // ---- bisection over the 7 key bits for rank klo: own part, then a shuffle sum over the column's lanes
        unsigned prefix = 0;
        int L = 0, LE = n_valid;
        const int sum_keys = key_byte_sum<RPT>(w.mykeys, WRP);
#pragma unroll
        for (int bit = 6; bit >= 0; --bit) {
            const unsigned cand = prefix | (1u << bit);
            const int ge = col_sum(w, (int)count_ge<RPT>(w.mykeys, WRP, cand, sum_keys));
            const int lt = slots - ge;            // sentinels (127) are never below a pivot
            if (lt <= klo) { prefix = cand; L = lt; }
            else LE = lt;                         // the last rejected pivot is prefix + 1: LE = #(key <= prefix)
        }
        // ---- the window: the bucket of rank klo alone when rank khi is inside it, else that and the next two
        const unsigned span = (khi < LE) ? 1u : 3u;
        const unsigned up = prefix + span > 127u ? 127u : prefix + span;
        const unsigned c1b = (128u - prefix) * 0x01010101u, c2b = (128u - up) * 0x01010101u;
        // flags of this part's rows, RP/32 words (compile-time count when RPT is given)
        unsigned fl[(RPT ? RPT : 512) / 32];
        int mine = 0;
#pragma unroll
        for (int g = 0; g < (RPT ? RPT : 512); g += 32) {
            fl[g / 32] = 0;
            if (g < WRP) {
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
                fl[g / 32] = acc;
                mine += __popc(acc);
            }
        }
        // slot offset of this part = candidates of the lower parts; m = all of them
        int off = 0, m = 0;
        {
            const int lane_part = w.part;
#pragma unroll
            for (int q = 0; q < (PT ? PT : 8); ++q) {
                if (q < WP) {
                    const int cq = __shfl_sync(w.cmask, mine, (threadIdx.x & 31 & ~(w.sh1 | w.sh2 | w.sh3)) | (q * (32 / WP)));
                    m += cq;
                    if (q < lane_part) off += cq;
                }
            }
        }
        const int j1 = klo - L, j2 = khi - L;
        bool rescue = false;
        in_t thr = const_value;
        key_t kmin = ~key_t(0), kmax = 0;
        if (m <= W2_CAP) {
            // values of this part's window rows -> slots off..  W2_CHUNK loads are issued before the first value is
            // stored (W2_CHUNK = 2 measured best: 1 serialises the round trips, 4 spends too many idle slots)
            int k = off;
#pragma unroll
            for (int g = 0; g < (RPT ? RPT : 512); g += 32) {
                unsigned acc = fl[g / 32];
                while (acc) {
                    in_t v[W2_CHUNK];
                    bool on[W2_CHUNK];
                    const int pos0 = __ffs(acc) - 1;
#pragma unroll
                    for (int u = 0; u < W2_CHUNK; ++u) {
                        on[u] = acc != 0u;
                        const int pos = on[u] ? __ffs(acc) - 1 : pos0;   // idle slots re-read a valid row
                        acc &= acc - 1;
                        v[u] = tile[(row_lo + g + 4 * (7 - (pos & 7)) + (pos >> 3)) * WD + w.c];
                    }
#pragma unroll
                    for (int u = 0; u < W2_CHUNK; ++u)
                        if (on[u]) lc[(k++) * WD] = v[u];
                }
            }
            __syncwarp(w.cmask);
            // rank this part's share of the slots: lt(v) = #{u < v}; rank j's value has the largest lt <= j
            int b1 = -1, b2 = -1;
            for (int base = w.part; base < m; base += W2_RANKW * WP) {
                in_t v[W2_RANKW];
                int rr[W2_RANKW];
#pragma unroll
                for (int t = 0; t < W2_RANKW; ++t) {
                    const int a = base + t * WP;
                    v[t] = a < m ? lc[a * WD] : lc[base * WD];
                    rr[t] = 0;
                }
#pragma unroll 1
                for (int b = 0; b < m; ++b) {
                    const in_t u = lc[b * WD];
#pragma unroll
                    for (int t = 0; t < W2_RANKW; ++t) count_lt(rr[t], u, v[t]);
                }
#pragma unroll
                for (int t = 0; t < W2_RANKW; ++t) {
                    const int a = base + t * WP;
                    if (a < m) {                      // largest (rank, slot) code with rank <= j1 / j2
                        const int code = (rr[t] << 8) | a;
                        if (rr[t] <= j1 && code > b1) b1 = code;
                        if (rr[t] <= j2 && code > b2) b2 = code;
                    }
                }
            }
            auto imax = [](int a, int b) { return a > b ? a : b; };
            b1 = col_reduce(w, b1, imax);
            b2 = col_reduce(w, b2, imax);
            v_lo = lc[(b1 & 0xff) * WD];
            v_hi = lc[(b2 & 0xff) * WD];
            rescue = j2 >= m;             // rank khi above the window: every rank is <= j2 then, so v_hi is the
            thr = v_hi;                   // window's largest value, the threshold of the rescue scan
            __syncwarp(w.cmask);                  // the list may be overwritten by the next selection
        } else {
            // crowded window (rare): its value range decides between "all one value" and a finer re-keying
#pragma unroll
            for (int g = 0; g < (RPT ? RPT : 512); g += 32) {
                unsigned acc = fl[g / 32];
                while (acc) {
                    const int pos = __ffs(acc) - 1;
                    acc &= acc - 1;
                    const key_t kx = okey(tile[(row_lo + g + 4 * (7 - (pos & 7)) + (pos >> 3)) * WD + w.c]);
                    kmin = kx < kmin ? kx : kmin;
                    kmax = kx > kmax ? kx : kmax;
                }
            }
            kmin = col_reduce(w, kmin, [](key_t a, key_t b) { return a < b ? a : b; });
            kmax = col_reduce(w, kmax, [](key_t a, key_t b) { return a > b ? a : b; });
            in_t a, b;
            unkey(kmin, a);
            unkey(kmax, b);
            typename Key::Scale sc;
            if (!(b > a)) {                       // the whole window is one value
                v_lo = a;
                v_hi = a;
                thr = a;
                rescue = j2 >= m;
            } else if (!Key::make(a, b, sc)) {
                val_t sk, ku;
                int stw = STW_OK;
                return column_centre_generic<Tr>(tile + w.c, WD, w.N, keep, n_valid, kind, false, sk, ku, stw);
            } else {
                w2_build_keys<Tr, false, DT, RPT>(w, tile, keep, a, b, sc);
                refined = true;
                continue;
            }
        }
        if (rescue) {                             // rank khi lies above the window: smallest flagged value above it
            // sixteen independent loads per trip (the flags are applied afterwards): a dependent chain of 64
            // L2 round trips here would hold the whole CTA at its next barrier
            key_t best = ~key_t(0);
            const in_t *ptr = tile + row_lo * WD + w.c;
            for (int row = row_lo; row < row_hi; row += 16, ptr += 16 * WD) {
                in_t x[16];
#pragma unroll
                for (int k = 0; k < 16; ++k) x[k] = (row + k < row_hi) ? ptr[k * WD] : thr;
                uint4 kf = make_uint4(0x01010101u, 0x01010101u, 0x01010101u, 0x01010101u);
                if (keep) kf = *reinterpret_cast<const uint4 *>(keep + row);      // row is a multiple of 16
                const unsigned kw4[4] = {kf.x, kf.y, kf.z, kf.w};
#pragma unroll
                for (int k = 0; k < 16; ++k) {
                    const bool on = (kw4[k >> 2] >> (8 * (k & 3))) & 0xffu;
                    if (on && x[k] > thr) { const key_t kx = okey(x[k]); best = kx < best ? kx : best; }
                }
            }
            best = col_reduce(w, best, [](key_t a, key_t b) { return a < b ? a : b; });
            unkey(best, v_hi);
        }
        break;
    }
    return Tr::half_sum((val_t)v_lo, (val_t)v_hi);
}

// Which rows are kept: the R rows that come first under (risk asc, index desc | asc) — contract.cairo:470-486
// drops the tail of that order.  Run by ONE warp (the others wait at the next barrier and leave the SM to
// the co-resident CTA); all-pairs counting over 16 warps costs about 25 times the instructions.
// Lane l holds the rows j*32 + l (j < NS) in q[].  A level works on the still undecided rows: 7-bit monotone
// keys over their exact range — linear in the BIT PATTERN of the float image of the risk, i.e. piecewise
// logarithmic in its value, because the kept rows usually form a narrow cluster far below the dropped
// ones — packed four per register; bisection over the key bits for the bucket of the cut (SWAR counting as
// in count_ge, one warp reduction per bit).  Rows in lower buckets are kept, rows in higher ones dropped,
// the cut bucket stays undecided: if it holds at most W2_RCAP rows they are ranked exactly, one per lane;
// if all its rows are equal the index order decides; else the next level re-keys it over its own range.
// Returns false when the risks are negative / non-finite or the levels run out: the caller then runs the
// all-pairs ranking.  words[j] = kept rows j*32 .. j*32+31 (same value in every lane).
template <class acc_t, int NS>
__device__ __forceinline__ bool w2_select_rows(const acc_t (&q)[NS], int N, int R, bool high, acc_t *wq, int *wi,
                                               unsigned *ww, unsigned (&words)[NS]) {
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    const unsigned below_me = (1u << lane) - 1u;
    constexpr int NW = (NS + 3) / 4;
    constexpr int NB = 32 * 4 * NW;                  // key bytes in the warp, sentinels included
    unsigned bits[NS];
    unsigned cand = 0, kept = 0;                     // slot masks of this lane: undecided / kept
#pragma unroll
    for (int j = 0; j < NS; ++j) {
        bits[j] = __float_as_uint((float)q[j]);      // monotone image (rounding keeps order); ties are settled on q
        if (j * 32 + lane < N) cand |= 1u << j;
    }
    int m = N, need = R;                             // undecided rows, and how many of them are kept
    bool done = false;
#pragma unroll 1
    for (int level = 0; level < 8 && !done; ++level) {
        unsigned bmin = 0xffffffffu, bmax = 0u;
#pragma unroll
        for (int j = 0; j < NS; ++j)
            if ((cand >> j) & 1u) {
                bmin = bits[j] < bmin ? bits[j] : bmin;
                bmax = bits[j] > bmax ? bits[j] : bmax;
            }
        bmin = __reduce_min_sync(full, bmin);
        bmax = __reduce_max_sync(full, bmax);
        if (bmax >= 0x7f800000u) return false;       // sign bit, inf, nan: non-negative finite floats order like their bits
        if (bmax == bmin) {
            if constexpr (!std::is_same<acc_t, float>::value) {
                // equal fp32 images need not be equal 64-bit risks (integers or doubles): check
                const unsigned first = __ballot_sync(full, cand != 0);
                acc_t ref = 0;
#pragma unroll
                for (int j = NS - 1; j >= 0; --j)
                    if ((cand >> j) & 1u) ref = q[j];
                ref = __shfl_sync(full, ref, __ffs(first) - 1);
                bool same = true;
#pragma unroll
                for (int j = 0; j < NS; ++j)
                    if ((cand >> j) & 1u) same &= q[j] == ref;
                if (!__all_sync(full, same)) return false;
            }
            // equal risks: the index order decides (rows are held in index order: slot-major, then lane)
            int base = 0;
#pragma unroll
            for (int j = 0; j < NS; ++j) {
                const unsigned c = (cand >> j) & 1u;
                const unsigned bal = __ballot_sync(full, c != 0u);
                const int pos = base + __popc(bal & below_me);
                kept |= (c & (unsigned)(high ? pos >= m - need : pos < need)) << j;
                base += __popc(bal);
            }
            done = true;
            break;
        }
        const float inv = 1.0f / (float)(bmax - bmin);
        unsigned key[NS], kw[NW];
#pragma unroll
        for (int j = 0; j < NS; ++j) {
            const float t = __saturatef((float)(bits[j] - bmin) * inv);
            key[j] = ((cand >> j) & 1u) ? (__float_as_uint(__fmaf_rn(t, 126.0f, 8388608.0f)) & 0xffu) : 127u;
        }
        int sk = 0;
#pragma unroll
        for (int g = 0; g < NW; ++g) {
            kw[g] = pack4(key[4 * g], 4 * g + 1 < NS ? key[4 * g + 1] : 127u, 4 * g + 2 < NS ? key[4 * g + 2] : 127u,
                          4 * g + 3 < NS ? key[4 * g + 3] : 127u);
            sk = (int)__dp4a(kw[g], 0x01010101u, (unsigned)sk);
        }
        sk = __reduce_add_sync(full, sk);
        const int k = need - 1;
        unsigned prefix = 0;
        int L = 0, LE = m;
#pragma unroll
        for (int bit = 6; bit >= 0; --bit) {
            const unsigned pv = prefix | (1u << bit);
            const unsigned cb = (128u - pv) * 0x01010101u;
            int acc = 0;
#pragma unroll
            for (int g = 0; g < NW; ++g) acc = __dp4a((int)(kw[g] + cb), 0x01010101, acc);
            acc = __reduce_add_sync(full, acc);
            const int lt = NB - ((sk + NB * (int)(128u - pv) - acc) >> 8);
            if (lt <= k) { prefix = pv; L = lt; }
            else LE = lt;
        }
        unsigned next = 0;
#pragma unroll
        for (int j = 0; j < NS; ++j) {
            if (key[j] < prefix) kept |= 1u << j;
            if (key[j] == prefix) next |= 1u << j;
        }
        cand = next;
        m = LE - L;
        need -= L;
        if (need == m) {                              // the cut falls on the bucket's end
            kept |= cand;
            done = true;
        } else if (m <= W2_RCAP) {
            // compact the cut bucket's rows (index order) into the list; lane t ranks entry t among them
            int base = 0;
#pragma unroll
            for (int j = 0; j < NS; ++j) {
                const bool c = (cand >> j) & 1u;
                const unsigned bal = __ballot_sync(full, c);
                if (c) {
                    const int pos = base + __popc(bal & below_me);
                    wq[pos] = q[j];
                    wi[pos] = j * 32 + lane;
                }
                base += __popc(bal);
            }
            if (lane < NS) ww[lane] = 0u;
            __syncwarp();
            const int t = lane < m ? lane : 0;
            const acc_t qv = wq[t];
            const int iv = wi[t];
            int before = 0;
            for (int b = 0; b < m; ++b) {
                const acc_t qb = wq[b];
                const int ib = wi[b];
                before += (qb < qv) | ((qb == qv) & (high ? ib > iv : ib < iv));
            }
            if (lane < m && before < need) atomicOr(&ww[iv >> 5], 1u << (iv & 31));
            __syncwarp();
#pragma unroll
            for (int j = 0; j < NS; ++j) kept |= ((ww[j] >> lane) & 1u) << j;
            __syncwarp();                             // the list is reused by the next round
            done = true;
        }
    }
    if (!done) return false;
#pragma unroll
    for (int j = 0; j < NS; ++j) words[j] = __ballot_sync(full, (kept >> j) & 1u);
    return true;
}

template <class Tr, int DT, int PT, int RPT>
__global__ void __launch_bounds__(W2_THREADS, w2_ctas_per_sm<Tr>()) consensus_wide2_kernel(const Params p) {
    using in_t = typename Tr::in_t;
    using val_t = typename Tr::val_t;
    using acc_t = typename Tr::acc_t;
    using Key = WideKey<Tr>;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int N = p.N, D = DT ? DT : p.D, R = p.R, P = PT ? PT : p.wP, RP = RPT ? RPT : p.wRP, CS = p.wCS;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    auto al = [](size_t x) { return (x + 15) & ~(size_t)15; };

    size_t off = 0;
    uint8_t *keysT = smem_raw + off;                            off += al((size_t)D * CS);
    in_t *list = reinterpret_cast<in_t *>(smem_raw + off);      off += al((size_t)W2_CAP * D * sizeof(in_t));
    val_t *c1 = reinterpret_cast<val_t *>(smem_raw + off);      off += al((size_t)D * sizeof(val_t));
    acc_t *qr_s = reinterpret_cast<acc_t *>(smem_raw + off);    off += al((size_t)P * RP * sizeof(acc_t));
    acc_t *wq = reinterpret_cast<acc_t *>(smem_raw + off);      off += al((size_t)W2_RCAP * sizeof(acc_t));
    int *wi = reinterpret_cast<int *>(smem_raw + off);          off += al((size_t)W2_RCAP * sizeof(int));
    unsigned *ww = reinterpret_cast<unsigned *>(smem_raw + off); off += 64;
    uint32_t *maskw = reinterpret_cast<uint32_t *>(smem_raw + off); off += al((size_t)(P * RP / 32 + 1) * 4);
    int *stw_s = reinterpret_cast<int *>(smem_raw + off);       off += 16;
    int *sel_s = stw_s + 1;
    uint8_t *keep_s = smem_raw + off;

    // column units: lane = part * CW + column-in-warp, CW = 32 / P columns per warp
    const int CW = 32 / P;
    W2Ctx w;
    w.N = N; w.D = D; w.RP = RP;
    w.part = lane / CW;
    w.c = warp * CW + (lane - w.part * CW);
    const bool has_unit = w.c < D;
    w.cmask = 0;
    for (int q = 0; q < P; ++q) w.cmask |= 1u << (q * CW + (lane % CW));
    w.sh1 = P >= 2 ? CW : 0;
    w.sh2 = P >= 4 ? 2 * CW : 0;
    w.sh3 = P >= 8 ? 4 * CW : 0;
    w.mykeys = keysT + (size_t)(has_unit ? w.c : 0) * CS + (size_t)w.part * RP;

    // row units for the row phases: HR lanes per row inside one warp (lane = h * (32/HR) + row-in-warp)
    const int HR = p.wH;                       // 1, 2 or 4
    const int RW = 32 / HR;                    // rows per warp
    const int rh = lane / RW;

    const size_t round_elems = (size_t)N * D;
    const uint32_t round_bytes = (uint32_t)(round_elems * sizeof(in_t));
    const int row_lo = w.part * RP;
    const bool shape2 = second_pass_shape_panic(p);

    for (i64 r = blockIdx.x; r < p.T; r += gridDim.x) {
        const in_t *tile = reinterpret_cast<const in_t *>(p.pred) + (size_t)r * round_elems;
        if (tid == 0) *stw_s = STW_OK;
        // ---- phase 1: first-pass centre of every column
        bool col_const = false, col_fb = false, refined1 = false, refined2 = false;
        in_t lo = 0, hi = 0;
        typename Key::Scale sc{};
        if (has_unit) {
            // key scale from a sample: the first 64/P rows of every part (values outside clamp to the end buckets)
            lo = std::numeric_limits<in_t>::max();
            hi = std::numeric_limits<in_t>::lowest();
            const int ns = min(W2_SAMPLE / P, max(0, min(N, row_lo + RP) - row_lo));
            const in_t *ptr = tile + row_lo * D + w.c;
            if (ns == W2_SAMPLE / 4) {           // the usual case, fully unrolled: loads at constant offsets
#pragma unroll
                for (int k = 0; k < W2_SAMPLE / 4; ++k) {
                    const in_t x = ptr[k * D];
                    lo = vmin(lo, x);
                    hi = vmax(hi, x);
                }
            } else {
                for (int k = 0; k < ns; ++k, ptr += D) {
                    const in_t x = *ptr;
                    lo = vmin(lo, x);
                    hi = vmax(hi, x);
                }
            }
            lo = col_reduce(w, lo, [](in_t a, in_t b) { return vmin(a, b); });
            hi = col_reduce(w, hi, [](in_t a, in_t b) { return vmax(a, b); });
            if (!(hi > lo)) {                    // degenerate sample: exact sweep, which also detects a constant column
                const in_t *q = tile + row_lo * D + w.c;
                for (int row = row_lo; row < min(N, row_lo + RP); ++row, q += D) {
                    const in_t x = *q;
                    lo = vmin(lo, x);
                    hi = vmax(hi, x);
                }
                lo = col_reduce(w, lo, [](in_t a, in_t b) { return vmin(a, b); });
                hi = col_reduce(w, hi, [](in_t a, in_t b) { return vmax(a, b); });
            }
            col_const = !(hi > lo);
            col_fb = !col_const && !Key::make(lo, hi, sc);
            if (!col_const && !col_fb) w2_build_keys<Tr, true, DT, RPT>(w, tile, nullptr, lo, hi, sc);
            const val_t e1 = w2_centre<Tr, DT, PT, RPT>(w, tile, nullptr, N, p.centre, col_const, col_fb, lo, list, P, refined1);
            if (w.part == 0) {
                c1[w.c] = e1;
                if (p.essence1) reinterpret_cast<val_t *>(p.essence1)[(size_t)r * D + w.c] = e1;
            }
        }
        __syncthreads();
        // the first sweep of this round is done: pull the CTA's next round towards L2 now (two rounds per CTA, 76 MB
        // in all, fit the L2; issued at the start of the round it evicted lines still in use, issued after the
        // risk sweep it came 0.5 % too late)
        if (tid == 0 && r + gridDim.x < p.T)
            asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(tile + (size_t)gridDim.x * round_elems), "r"(round_bytes) : "memory");

        // ---- phase 2: quadratic risk of every row.  A warp reads one whole row per request (lane <-> 4
        //      consecutive dimensions, 512 contiguous bytes for D = 128), so the loads are fully coalesced;
        //      the lanes' partial sums meet in a shuffle tree (fixed order: deterministic).
        if ((D & 3) == 0) {
            // four rows per trip: their lane partials are folded with a transposing shuffle tree
            // (2 + 1 + 3 shuffles for four row sums instead of 4 x 5)
            for (int i0 = 4 * warp; i0 < N; i0 += 4 * (W2_THREADS / 32)) {
                acc_t q[4] = {0, 0, 0, 0};
                bool bad = false;
                // one row pointer per trip, opaque to the compiler: it would otherwise rebuild the 64-bit round
                // offset in front of every load
                const in_t *rp = tile + (size_t)i0 * D + 4 * lane;
                asm volatile("" : "+l"(rp));
                for (int d = 4 * lane; d < D; d += 128, rp += 128) {
                    val_t cc[4];
#pragma unroll
                    for (int k = 0; k < 4; ++k) cc[k] = c1[d + k];
                    constexpr int NQ = sizeof(in_t) / 4;          // 16-byte loads per 4 elements
                    int4 xv[4][NQ];
#pragma unroll
                    for (int u = 0; u < 4; ++u)
#pragma unroll
                        for (int h = 0; h < NQ; ++h)
                            xv[u][h] = (i0 + u < N) ? reinterpret_cast<const int4 *>(rp + u * D)[h] : make_int4(0, 0, 0, 0);
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        in_t x[4];
                        memcpy(x, &xv[u][0], 4 * sizeof(in_t));
                        if (i0 + u < N) {
#pragma unroll
                            for (int k = 0; k < 4; ++k) {
                                if constexpr (Tr::fixed) {
                                    if (p.constrained) bad |= (x[k] < 0) | ((i64)x[k] > WSAD);
                                    else bad |= ((i64)x[k] > ENGINE_RANGE) | ((i64)x[k] < -ENGINE_RANGE);
                                }
                                q[u] += Tr::dev(x[k], cc[k]);
                            }
                        }
                    }
                }
                // lanes with bit 4 set end up owning rows 2,3; bit 3 picks the odd row of the pair
                const bool up16 = lane & 16, up8 = lane & 8;
                acc_t k0 = up16 ? q[2] : q[0], s0 = up16 ? q[0] : q[2];
                acc_t k1 = up16 ? q[3] : q[1], s1 = up16 ? q[1] : q[3];
                k0 += __shfl_xor_sync(0xffffffffu, s0, 16);
                k1 += __shfl_xor_sync(0xffffffffu, s1, 16);
                acc_t kk = up8 ? k1 : k0, ss = up8 ? k0 : k1;
                kk += __shfl_xor_sync(0xffffffffu, ss, 8);
                kk += __shfl_xor_sync(0xffffffffu, kk, 4);
                kk += __shfl_xor_sync(0xffffffffu, kk, 2);
                kk += __shfl_xor_sync(0xffffffffu, kk, 1);
                const int mine_row = i0 + (up16 ? 2 : 0) + (up8 ? 1 : 0);
                if ((lane & 7) == 0 && mine_row < N) {
                    qr_s[mine_row] = kk;
                    if (p.qr) reinterpret_cast<acc_t *>(p.qr)[(size_t)r * N + mine_row] = kk;
                }
                if (bad) atomicMin(stw_s, p.constrained ? STW_INTERVAL : STW_ENGINE_RANGE);
            }
        } else {
            for (int i = tid; i < N; i += W2_THREADS) {
                const in_t *row = tile + (size_t)i * D;
                acc_t q = 0;
                bool bad = false;
                for (int d = 0; d < D; ++d) {
                    const in_t x = row[d];
                    if constexpr (Tr::fixed) {
                        if (p.constrained) bad |= (x < 0) | ((i64)x > WSAD);
                        else bad |= ((i64)x > ENGINE_RANGE) | ((i64)x < -ENGINE_RANGE);
                    }
                    q += Tr::dev(x, c1[d]);
                }
                qr_s[i] = q;
                if (p.qr) reinterpret_cast<acc_t *>(p.qr)[(size_t)r * N + i] = q;
                if (bad) atomicMin(stw_s, p.constrained ? STW_INTERVAL : STW_ENGINE_RANGE);
            }
        }
        __syncthreads();

        // ---- phase 3: which rows are kept.  Warp 0 works it out alone (w2_select_rows) while the others wait;
        //      the all-pairs ranking — rank of every row under (qr asc, index desc | asc), rows spread over all
        //      warps — runs when ranks are an output or the selection gives up on this round's risks.
        {
            constexpr int NS = (PT && RPT) ? (PT * RPT) / 32 : 16;
            const bool high = p.tie == SVOC_TIE_HIGH_INDEX_KEPT;
            unsigned words[NS];
            acc_t qrow[NS];
            if (warp == 0) {
#pragma unroll
                for (int j = 0; j < NS; ++j)          // qr_s has P*RP entries; rows >= N are masked out by the selection
                    qrow[j] = (j * 32 < P * RP) ? qr_s[j * 32 + lane] : acc_t(0);
                bool fast = W2_SELECT_FAST && p.rank == nullptr;
                if (fast) fast = w2_select_rows<acc_t, NS>(qrow, N, R, high, wq, wi, ww, words);
                if (fast) {
                    unsigned mine = 0;
#pragma unroll
                    for (int j = 0; j < NS; ++j) mine = lane == j ? words[j] : mine;
                    if (lane * 32 < P * RP) maskw[lane] = mine;
                    __syncwarp();
                    // flag bytes, eight rows per lane and trip: bit b of a mask byte -> byte b (0 / 1)
                    for (int g = lane; g * 8 < P * RP; g += 32) {
                        const unsigned bits8 = (maskw[g >> 2] >> (8 * (g & 3))) & 0xffu;
                        const unsigned lo4 = ((bits8 & 0xfu) * 0x00204081u) & 0x01010101u;
                        const unsigned hi4 = ((bits8 >> 4) * 0x00204081u) & 0x01010101u;
                        *reinterpret_cast<uint2 *>(keep_s + 8 * g) = make_uint2(lo4, hi4);
                    }
                }
                if (lane == 0) *sel_s = fast ? 1 : 0;
            } else if (warp == 1) {
                // meanwhile an idle warp does the first-pass reliability, which needs all risks but no selection
                acc_t s1 = 0;
#pragma unroll
                for (int j = 0; j < NS; ++j)
                    if (j * 32 + lane < N) s1 += qr_s[j * 32 + lane];
                for (int o = 16; o; o >>= 1) s1 += __shfl_xor_sync(0xffffffffu, s1, o);
                if (lane == 0) {
                    int stw = STW_OK;
                    if constexpr (Tr::fixed) {
                        const i64 rel1 = reliability_fixed(s1 / (u64)N, p, STW_SQRT_REL1, stw);
                        if (rel1 < 0 || rel1 > WSAD) stw = min(stw, STW_REL1_RANGE);
                        reinterpret_cast<i64 *>(p.rel1)[r] = rel1;
                    } else {
                        const val_t rel1 = reliability_float<val_t>(s1 / (val_t)N, p);
                        if (!(rel1 >= 0 && rel1 <= 1)) stw = min(stw, STW_REL1_RANGE);
                        reinterpret_cast<val_t *>(p.rel1)[r] = rel1;
                    }
                    if (stw != STW_OK) atomicMin(stw_s, stw);
                }
            }
            __syncthreads();
            if (!*sel_s) {
                for (int ib = warp * RW; ib < N; ib += (W2_THREADS / 32) * RW) {
                    const int i = ib + (lane - rh * RW);
                    const bool ok = i < N;
                    const acc_t mine = qr_s[ok ? i : 0];
                    const int i1 = min(N, ib + RW);             // the warp's rows are [ib, i1)
                    const int jh = (((N + HR - 1) / HR) + 3) & ~3, jlo = rh * jh, jhi = min(N, jlo + jh);
                    int tie = 0;
                    // [jlo, jhi) minus the warp's own rows: one compare per pair, strict on one side of the warp's
                    // rows and non-strict on the other
                    auto count = [&](auto strict, int a, int b) {
                        int c = 0, j = a;
                        auto hit = [&](acc_t q) {
                            if constexpr (sizeof(acc_t) == 4) {      // set.{lt,le}.s32.f32: 0 or -1 in one instruction
                                int d;
                                float qf, mf;
                                memcpy(&qf, &q, 4);
                                memcpy(&mf, &mine, 4);
                                if constexpr (decltype(strict)::value) asm("set.lt.s32.f32 %0, %1, %2;" : "=r"(d) : "f"(qf), "f"(mf));
                                else asm("set.le.s32.f32 %0, %1, %2;" : "=r"(d) : "f"(qf), "f"(mf));
                                return -d;
                            } else {
                                if constexpr (decltype(strict)::value) return q < mine ? 1 : 0;
                                else return q <= mine ? 1 : 0;
                            }
                        };
                        if constexpr (sizeof(acc_t) == 4) {
                            for (; j < b && (j & 3); ++j) c += hit(qr_s[j]);
                            for (; j + 4 <= b; j += 4) {
                                const float4 v = *reinterpret_cast<const float4 *>(reinterpret_cast<const float *>(qr_s) + j);
                                acc_t q0, q1, q2, q3;
                                memcpy(&q0, &v.x, 4); memcpy(&q1, &v.y, 4); memcpy(&q2, &v.z, 4); memcpy(&q3, &v.w, 4);
                                c += (hit(q0) + hit(q1)) + (hit(q2) + hit(q3));
                            }
                        }
                        for (; j < b; ++j) c += hit(qr_s[j]);
                        return c;
                    };
                    const int a_end = min(jhi, max(jlo, ib)), b_beg = max(jlo, min(jhi, i1));
                    int rank = high ? count(std::true_type{}, jlo, a_end) + count(std::false_type{}, b_beg, jhi)
                                    : count(std::false_type{}, jlo, a_end) + count(std::true_type{}, b_beg, jhi);
                    for (int j = a_end; j < b_beg; ++j) {
                        const acc_t q = qr_s[j];
                        tie += (q < mine ? 1 : 0) + (((q == mine) & (high ? j > i : j < i)) ? 1 : 0);
                    }
                    rank += tie;
                    if (HR >= 2) rank += __shfl_xor_sync(0xffffffffu, rank, RW);
                    if (HR >= 4) rank += __shfl_xor_sync(0xffffffffu, rank, 2 * RW);
                    if (ok && rh == 0) {
                        keep_s[i] = rank < R ? 1 : 0;
                        if (p.rank) p.rank[(size_t)r * N + i] = rank;
                    }
                }
                __syncthreads();
                if (warp == 0) {
#pragma unroll
                    for (int j = 0; j < NS; ++j) {
                        words[j] = __ballot_sync(0xffffffffu, j * 32 + lane < N && keep_s[j * 32 + lane]);
                        if (lane == j && j * 32 < P * RP) maskw[j] = words[j];
                    }
                }
                __syncthreads();
            }
            // ---- the second-pass reliability and the status: warp 0, while the others start on their columns
            if (warp == 0) {
                // risks and mask words are read back from shared memory: nothing stays live across the barrier
                acc_t s2 = 0;
#pragma unroll
                for (int j = 0; j < NS; ++j)
                    if (j * 32 + lane < N && ((maskw[j] >> lane) & 1u)) s2 += qr_s[j * 32 + lane];
                if (p.reliable && lane < p.words) p.reliable[(size_t)r * p.words + lane] = maskw[lane];
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
                        reinterpret_cast<i64 *>(p.rel2)[r] = rel2;
                    } else {
                        val_t rel2 = 0;
                        if (shape2) stw = min(stw, STW_SECOND_SHAPE);
                        else {
                            rel2 = reliability_float<val_t>(s2 / (val_t)R, p);
                            if (!(rel2 >= 0 && rel2 <= 1)) stw = min(stw, STW_REL2_RANGE);
                        }
                        reinterpret_cast<val_t *>(p.rel2)[r] = rel2;
                    }
                    if (stw != STW_OK) atomicMin(stw_s, stw);
                    if (!p.want_moments) {   // nothing later contributes to the status: publish it now
                        const int sw = *stw_s;
                        p.status[r] = (sw == STW_OK) ? SVOC_ST_OK : (sw & 0xff);
                    }
                }
            }
        }
        // ---- phase 4: second-pass centre over the reliable rows (+ moments)
        if (has_unit) {
            val_t cons = 0;
            if (!shape2) {
                if (!col_const && !col_fb && p.centre2 != SVOC_CENTRE_MEAN) {
                    if (refined1) {
                        w2_build_keys<Tr, true, DT, RPT>(w, tile, keep_s, lo, hi, sc);
                    } else {
                        // dropped rows of this part -> sentinel key.  Branch-free: four drop bits are spread to
                        // the low bits of four bytes (one multiply), scaled to 0x7f and OR-ed into the key word
                        // (keys are <= 127, and rows >= N already hold the sentinel)
#pragma unroll
                        for (int g = 0; g < (RPT ? RPT : 512); g += 32) {
                            if (g < RP) {
                                const unsigned drop = ~maskw[(row_lo + g) >> 5];
                                uint4 a = *reinterpret_cast<const uint4 *>(w.mykeys + g);
                                uint4 b = *reinterpret_cast<const uint4 *>(w.mykeys + g + 16);
                                auto sent = [&](int j) { return ((((drop >> (4 * j)) & 0xfu) * 0x00204081u) & 0x01010101u) * 0x7fu; };
                                a.x |= sent(0); a.y |= sent(1); a.z |= sent(2); a.w |= sent(3);
                                b.x |= sent(4); b.y |= sent(5); b.z |= sent(6); b.w |= sent(7);
                                *reinterpret_cast<uint4 *>(w.mykeys + g) = a;
                                *reinterpret_cast<uint4 *>(w.mykeys + g + 16) = b;
                            }
                        }
                    }
                }
                cons = w2_centre<Tr, DT, PT, RPT>(w, tile, keep_s, R, p.centre2, col_const, col_fb, lo, list, P, refined2);
            }
            if (w.part == 0) {
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
        }
        // Without moments nothing below touches shared state that the next round writes before its own first
        // barrier (keys and candidate lists are private to a column's lanes), so warps run ahead into the next
        // round's column phase instead of waiting for the slowest column here.
        if (p.want_moments) {
            __syncthreads();
            if (tid == 0) {
                const int sw = *stw_s;
                p.status[r] = (sw == STW_OK) ? SVOC_ST_OK : (sw & 0xff);
            }
        }
    }
}

inline size_t wide2_smem_bytes(int N, int D, int P, int RP, int CS, size_t es, size_t vs) {
    auto al = [](size_t x) { return (x + 15) & ~(size_t)15; };
    size_t s = 0;
    s += al((size_t)D * CS);
    s += al((size_t)W2_CAP * D * es);
    s += al((size_t)D * vs);
    s += al((size_t)P * RP * vs);
    s += al((size_t)W2_RCAP * vs);
    s += al((size_t)W2_RCAP * 4);
    s += 64;
    s += al((size_t)(P * RP / 32 + 1) * 4);
    s += 16;
    s += al((size_t)P * RP);
    return s;
}

template <class Tr>
int launch_wide2(const Params &p, const Plan &pl, cudaStream_t st) {
    void (*kern)(const Params) = consensus_wide2_kernel<Tr, 0, 0, 0>;
#if W2_THREADS_N == 512
    if (p.D == 128 && p.wP == 4 && p.wRP == 64) kern = consensus_wide2_kernel<Tr, 128, 4, 64>;
    else if (sizeof(typename Tr::in_t) == 4 && p.D == 128 && p.wP == 4) kern = consensus_wide2_kernel<Tr, 128, 4, 0>;
#else
    if (p.D == 128 && p.wP == 2 && p.wRP == 128) kern = consensus_wide2_kernel<Tr, 128, 2, 128>;
#endif
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl.smem);
    if (e != cudaSuccess) return fail(std::string("cudaFuncSetAttribute: ") + cudaGetErrorString(e));
    // several CTAs per SM are planned: ask for the full shared-memory carveout (the default is not always enough)
    cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, (int)cudaSharedmemCarveoutMaxShared);
    kern<<<(unsigned)pl.grid, W2_THREADS, pl.smem, st>>>(p);
    e = cudaGetLastError();
    if (e != cudaSuccess) return fail(std::string("kernel launch: ") + cudaGetErrorString(e));
    count_launch();
    return 0;
}

}  // namespace svoc
