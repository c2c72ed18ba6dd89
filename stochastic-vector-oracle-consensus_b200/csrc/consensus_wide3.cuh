// consensus_wide3.cuh — the `wide3` variant of the fused consensus round: 64 < N <= 256 oracles, D = 128
// dimensions, 32-bit elements (fp32 / int32).  This is the kernel BASELINE config 4 runs.
//
// It keeps the selection method of `wide2` (monotone 7-bit keys, SWAR bisection, exact ranking of a small
// window; risks selected by one warp) and changes where the round lives, after the round-1 profiles showed
// `wide2` waiting on L2 round trips (every sweep after the first re-read the round from L2) and spending a
// fifth of its instructions on loads, address arithmetic and shared-memory key traffic:
//
//   * A round is split by COLUMNS over a cluster of two CTAs (64 columns each, 256 threads, 128 registers):
//     two CTAs — halves of two different rounds — share an SM and fill each other's barrier waits.
//   * Each CTA's half round (256 rows x 64 columns, 64 KB) is fetched by two 2-D TMA tensor-map copies
//     (cp.async.bulk.tensor, 128-byte swizzle, completing on an mbarrier) into a shared-memory landing buffer,
//     and moved from there into REGISTERS: a thread owns 64 values of one column (rows 8k + 2*part + {0,1}; with
//     the swizzle the 32 lanes of a warp read 32 different banks).  Keys (16 packed words) live in registers
//     too.  As soon as every thread holds its values the buffer is free, so the NEXT round's copy is issued
//     right then and lands while this round is being computed: HBM latency is never exposed, and after the
//     one LDS per value nothing in the round touches global memory or L2 again except the outputs.
//   * The per-oracle risks need all 128 columns: each CTA sums its 64 columns (a transposing shuffle tree over
//     the 8 columns of a warp, then 8 warp partials per row through shared memory, fixed order), writes its 256
//     partial sums into the PEER's shared memory (DSMEM) and one cluster barrier later both CTAs hold the same
//     full risks — partial(rank 0) + partial(rank 1) on both sides, bit-identical — and select the same rows
//     redundantly.  Partial buffers are double-buffered by round parity, so one cluster barrier per round is
//     enough.
//
// Anything this kernel does not cover (moments, rank output, mean centres, other D, misaligned rounds) is
// dispatched to `wide2`.  Reference semantics: see consensus_kernel.cuh (contract.cairo:442-503 etc.).
#pragma once
#include <cuda.h>

#include "consensus_wide2.cuh"

namespace svoc {

constexpr int W3_THREADS = 256;
constexpr int W3_COLS = 64;          // columns per CTA
constexpr int W3_D = 128;            // columns per round (two CTAs)
constexpr int W3_ROWS = 256;         // rows of a landing box (rows >= N are ignored)
constexpr int W3_NV = 64;            // values per thread
constexpr int W3_NKW = 16;           // packed key words per thread
constexpr int W3_CAP = 32;           // candidate slots per column
constexpr int W3_BOX_BYTES = 32 * W3_ROWS * 4;     // one 32-column box

struct W3Ctx {
    int N;
    int part;            // 0..3: rows 8k + 2*part + {0, 1}
    unsigned cmask;      // the four lanes of this thread's column
    int lane_base;       // lane of part 0 of this column
};

__device__ __forceinline__ int w3_row(int j, int part) { return 8 * (j >> 1) + 2 * part + (j & 1); }

template <class T>
__device__ __forceinline__ T w3_col_sum(const W3Ctx &w, T v) {
    v += __shfl_xor_sync(w.cmask, v, 8);
    v += __shfl_xor_sync(w.cmask, v, 16);
    return v;
}
template <class T, class Op>
__device__ __forceinline__ T w3_col_reduce(const W3Ctx &w, T v, Op op) {
    v = op(v, __shfl_xor_sync(w.cmask, v, 8));
    v = op(v, __shfl_xor_sync(w.cmask, v, 16));
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
__device__ __forceinline__ void st_peer(uint32_t addr, float v) {
    asm volatile("st.shared::cluster.f32 [%0], %1;" ::"r"(addr), "f"(v) : "memory");
}
__device__ __forceinline__ void st_peer(uint32_t addr, u64 v) {
    asm volatile("st.shared::cluster.u64 [%0], %1;" ::"r"(addr), "l"(v) : "memory");
}
__device__ __forceinline__ void st_peer(uint32_t addr, int v) {
    asm volatile("st.shared::cluster.s32 [%0], %1;" ::"r"(addr), "r"(v) : "memory");
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// one box of the round: 32 columns x 256 rows starting at (col0, row0) of the [T*N][128] tensor
__device__ __forceinline__ void tma_load_box(void *dst, const CUtensorMap *map, int col0, int row0, u64 *bar) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(
            smem_u32(dst)),
        "l"(map), "r"(col0), "r"(row0), "r"(smem_u32(bar))
        : "memory");
}

// if (flag) { [slot] = v; slot += one candidate slot; } without a branch
__device__ __forceinline__ void store_if(uint32_t &slot, unsigned flag, float v) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.u32 p, %1, 0;\n\t@p st.shared.f32 [%0], %2;\n\t@p add.u32 %0, %0, 256;\n\t}"
                 : "+r"(slot) : "r"(flag), "f"(v) : "memory");
}
__device__ __forceinline__ void store_if(uint32_t &slot, unsigned flag, int32_t v) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.u32 p, %1, 0;\n\t@p st.shared.s32 [%0], %2;\n\t@p add.u32 %0, %0, 256;\n\t}"
                 : "+r"(slot) : "r"(flag), "r"(v) : "memory");
}
static_assert(W3_COLS * 4 == 256, "store_if advances by one candidate slot");

// the exact, slow column centre as a real call: it is reached only by degenerate columns, and inlining its four
// call sites would add a sixth to the kernel's code
template <class Tr>
__device__ __noinline__ typename Tr::val_t w3_exact_centre(const typename Tr::in_t *gcol, int N, const uint8_t *keep,
                                                           int n_valid, int kind) {
    typename Tr::val_t sk, ku;
    int stw = STW_OK;
    return column_centre_generic<Tr>(gcol, W3_D, N, keep, n_valid, kind, false, sk, ku, stw);
}

// ---- keys in registers
template <class Tr, bool FIRST, bool FULL, bool ONLY_VALID>
__device__ __forceinline__ void w3_build_keys(const typename Tr::in_t (&x)[W3_NV], unsigned (&kw)[W3_NKW], int N, int part,
                                              typename Tr::in_t lo, typename Tr::in_t hi,
                                              const typename WideKey<Tr>::Scale &sc) {
    // ONLY_VALID: rows whose current key is the sentinel (dropped / padded) keep it
#pragma unroll
    for (int q = 0; q < W3_NKW; ++q) {
        unsigned f[4];
        const unsigned old = kw[q];
#pragma unroll
        for (int b = 0; b < 4; ++b) {
            const int j = 4 * q + b;
            f[b] = wide_key<Tr, FIRST>(x[j], lo, hi, sc);        // pack4 takes the low byte
            if (!FULL && w3_row(j, part) >= N) f[b] = 127u;
            if (ONLY_VALID && ((old >> (8 * b)) & 0xffu) == 127u) f[b] = 127u;
        }
        kw[q] = pack4(f[0], f[1], f[2], f[3]);
    }
}

__device__ __forceinline__ int w3_key_sum(const unsigned (&kw)[W3_NKW]) {
    unsigned a0 = 0, a1 = 0;
#pragma unroll
    for (int q = 0; q < W3_NKW; q += 2) {
        a0 = __dp4a(kw[q], 0x01010101u, a0);
        a1 = __dp4a(kw[q + 1], 0x01010101u, a1);
    }
    return (int)(a0 + a1);
}

// #(key >= cand) among this thread's 64 keys — see count_ge in consensus_wide.cuh for the SWAR identity
__device__ __forceinline__ int w3_count_ge(const unsigned (&kw)[W3_NKW], unsigned cand, int sum_keys) {
    const unsigned cb = (128u - cand) * 0x01010101u;
    int a0 = 0, a1 = 0, a2 = 0, a3 = 0;
#pragma unroll
    for (int q = 0; q < W3_NKW; q += 4) {
        a0 = __dp4a((int)__viaddmax_u32(kw[q], cb, 0u), 0x01010101, a0);
        a1 = __dp4a((int)__viaddmax_u32(kw[q + 1], cb, 0u), 0x01010101, a1);
        a2 = __dp4a((int)__viaddmax_u32(kw[q + 2], cb, 0u), 0x01010101, a2);
        a3 = __dp4a((int)__viaddmax_u32(kw[q + 3], cb, 0u), 0x01010101, a3);
    }
    return (int)((unsigned)(sum_keys + W3_NV * (int)(128u - cand) - ((a0 + a1) + (a2 + a3))) >> 8);
}

// Centre of the valid rows of this thread's column (valid = key != sentinel: all rows < N in the first pass, the
// kept rows in the second).  Executed by the four lanes of the column together; every one returns the result.
// gcol / keep serve only the rare exact fallback, which walks the column in global memory.
template <class Tr, bool FULL>
__device__ __forceinline__ typename Tr::val_t w3_centre(const W3Ctx &w, const typename Tr::in_t (&x)[W3_NV],
                                                        unsigned (&kw)[W3_NKW], const typename Tr::in_t *gcol,
                                                        const uint8_t *keep, int n_valid, int kind, bool col_const,
                                                        bool col_fallback, typename Tr::in_t const_value,
                                                        typename Tr::in_t *lc, bool &refined) {
    using in_t = typename Tr::in_t;
    using val_t = typename Tr::val_t;
    using Key = WideKey<Tr>;
    refined = false;
    if (col_const) return Tr::half_sum((val_t)const_value, (val_t)const_value);
    if (col_fallback) return w3_exact_centre<Tr>(gcol, w.N, keep, n_valid, kind);
    int klo, khi;
    centre_ranks(kind, n_valid, klo, khi);
    in_t v_lo = const_value, v_hi = const_value;

    for (;;) {
        // ---- bisection over the 7 key bits for rank klo: own 64 keys, then a shuffle sum over the column's lanes
        unsigned prefix = 0;
        int L = 0, LE = n_valid;
        const int sum_keys = w3_key_sum(kw);
#pragma unroll
        for (int bit = 6; bit >= 0; --bit) {
            const unsigned cand = prefix | (1u << bit);
            const int ge = w3_col_sum(w, w3_count_ge(kw, cand, sum_keys));
            const int lt = 4 * W3_NV - ge;        // sentinels (127) are never below a pivot
            if (lt <= klo) { prefix = cand; L = lt; }
            else LE = lt;                         // the last rejected pivot is prefix + 1: LE = #(key <= prefix)
        }
        // ---- the window: the bucket of rank klo alone when rank khi is inside it, else that and the next two
        const unsigned span = (khi < LE) ? 1u : 3u;
        const unsigned up = prefix + span > 127u ? 127u : prefix + span;
        const unsigned c1b = (128u - prefix) * 0x01010101u, c2b = (128u - up) * 0x01010101u;
        unsigned fl[W3_NKW];                      // byte MSB set <=> that key is inside the window
        int mine = 0;
#pragma unroll
        for (int q = 0; q < W3_NKW; ++q) {
            fl[q] = (kw[q] + c1b) & ~(kw[q] + c2b) & 0x80808080u;
            mine += __popc(fl[q]);
        }
        // slot offset of this part = candidates of the lower parts; m = all of them
        int off = 0, m = 0;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const int cq = __shfl_sync(w.cmask, mine, w.lane_base + 8 * q);
            m += cq;
            if (q < w.part) off += cq;
        }
        const int j1 = klo - L, j2 = khi - L;
        bool rescue = false;
        in_t thr = const_value;
        if (m <= W3_CAP) {
            // values of this part's window rows -> slots off.. (straight from the registers).  Written as a test, a
            // predicated store and a predicated address bump per value: left to itself the compiler branches
            // around every store (a dozen instructions per element)
            uint32_t slot = smem_u32(lc) + (uint32_t)off * (W3_COLS * 4);
#pragma unroll
            for (int q = 0; q < W3_NKW; ++q) {
#pragma unroll
                for (int b = 0; b < 4; ++b) store_if(slot, fl[q] & (0x80u << (8 * b)), x[4 * q + b]);
            }
            __syncwarp(w.cmask);
            // rank this part's share of the slots: lt(v) = #{u < v}; rank j's value has the largest lt <= j
            int b1 = -1, b2 = -1;
            for (int base = w.part; base < m; base += W2_RANKW * 4) {
                in_t v[W2_RANKW];
                int rr[W2_RANKW];
#pragma unroll
                for (int t = 0; t < W2_RANKW; ++t) {
                    const int a = base + t * 4;
                    v[t] = a < m ? lc[a * W3_COLS] : lc[base * W3_COLS];
                    rr[t] = 0;
                }
#pragma unroll 1
                for (int b = 0; b < m; ++b) {
                    const in_t u = lc[b * W3_COLS];
#pragma unroll
                    for (int t = 0; t < W2_RANKW; ++t) count_lt(rr[t], u, v[t]);
                }
#pragma unroll
                for (int t = 0; t < W2_RANKW; ++t) {
                    const int a = base + t * 4;
                    if (a < m) {                      // largest (rank, slot) code with rank <= j1 / j2
                        const int code = (rr[t] << 8) | a;
                        if (rr[t] <= j1 && code > b1) b1 = code;
                        if (rr[t] <= j2 && code > b2) b2 = code;
                    }
                }
            }
            auto imax = [](int a, int b) { return a > b ? a : b; };
            b1 = w3_col_reduce(w, b1, imax);
            b2 = w3_col_reduce(w, b2, imax);
            v_lo = lc[(b1 & 0xff) * W3_COLS];
            v_hi = lc[(b2 & 0xff) * W3_COLS];
            rescue = j2 >= m;             // rank khi above the window: every rank is <= j2 then, so v_hi is the
            thr = v_hi;                   // window's largest value, the threshold of the rescue scan
            __syncwarp(w.cmask);          // the list may be overwritten by the next selection
        } else {
            // crowded window (rare): its value range decides between "all one value" and a finer re-keying
            uint32_t kmin = 0xffffffffu, kmax = 0u;
#pragma unroll
            for (int q = 0; q < W3_NKW; ++q) {
#pragma unroll
                for (int b = 0; b < 4; ++b) {
                    if (fl[q] & (0x80u << (8 * b))) {
                        const uint32_t kx = okey(x[4 * q + b]);
                        kmin = kx < kmin ? kx : kmin;
                        kmax = kx > kmax ? kx : kmax;
                    }
                }
            }
            kmin = w3_col_reduce(w, kmin, [](uint32_t a, uint32_t b) { return a < b ? a : b; });
            kmax = w3_col_reduce(w, kmax, [](uint32_t a, uint32_t b) { return a > b ? a : b; });
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
                return w3_exact_centre<Tr>(gcol, w.N, keep, n_valid, kind);
            } else {
                w3_build_keys<Tr, false, FULL, true>(x, kw, w.N, w.part, a, b, sc);
                refined = true;
                continue;
            }
        }
        if (rescue) {                             // rank khi lies above the window: smallest valid value above it
            uint32_t best = 0xffffffffu;
#pragma unroll
            for (int q = 0; q < W3_NKW; ++q) {
#pragma unroll
                for (int b = 0; b < 4; ++b) {
                    const bool valid = ((kw[q] >> (8 * b)) & 0xffu) != 127u;
                    const in_t xv = x[4 * q + b];
                    if (valid && xv > thr) { const uint32_t kx = okey(xv); best = kx < best ? kx : best; }
                }
            }
            best = w3_col_reduce(w, best, [](uint32_t a, uint32_t b) { return a < b ? a : b; });
            unkey(best, v_hi);
        }
        break;
    }
    return Tr::half_sum((val_t)v_lo, (val_t)v_hi);
}

struct W3Smem {
    // offsets in bytes from the 1024-aligned base
    size_t tile, list, part, mine, theirs, xst, qr, wq, wi, ww, maskw, misc, keep, bar, total;
};
__host__ __device__ inline W3Smem w3_smem_layout(size_t vs) {
    W3Smem s{};
    size_t off = 0;
    auto take = [&](size_t bytes, size_t align) { off = (off + align - 1) & ~(align - 1); const size_t at = off; off += bytes; return at; };
    s.tile = take(2 * W3_BOX_BYTES, 1024);
    s.list = take((size_t)W3_CAP * W3_COLS * 4, 16);
    s.part = take((size_t)8 * W3_ROWS * vs, 16);          // [8 warps][256 rows] partial risks
    s.mine = take((size_t)2 * W3_ROWS * vs, 16);          // [parity][row]: this CTA's 64-column partial
    s.theirs = take((size_t)2 * W3_ROWS * vs, 16);        // written by the peer CTA
    s.xst = take(4 * sizeof(int), 16);                    // [parity] peer's status word, [2 + parity] unused
    s.qr = take((size_t)W3_ROWS * vs, 16);
    s.wq = take((size_t)W2_RCAP * vs, 16);
    s.wi = take((size_t)W2_RCAP * 4, 16);
    s.ww = take(64, 16);
    s.maskw = take(16 * 4, 16);
    s.misc = take(16, 16);                                // stw, sel
    s.keep = take(W3_ROWS, 16);
    s.bar = take(16, 16);
    s.total = off;
    return s;
}

template <class Tr, bool FULL>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(W3_THREADS, 2)
consensus_wide3_kernel(const Params p, const __grid_constant__ CUtensorMap tmap) {
    using in_t = typename Tr::in_t;
    using val_t = typename Tr::val_t;
    using acc_t = typename Tr::acc_t;
    using Key = WideKey<Tr>;
    extern __shared__ __align__(1024) unsigned char smem_w3[];     // the 128-byte swizzle wants the boxes 1024-aligned
    unsigned char *const smem_raw = smem_w3;
    const W3Smem S = w3_smem_layout(sizeof(acc_t));
    const int N = FULL ? W3_ROWS : p.N, R = p.R;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t rank = cluster_ctarank(), peer = rank ^ 1u;
    const i64 cluster_id = blockIdx.x >> 1, n_clusters = gridDim.x >> 1;

    in_t *tile_s = reinterpret_cast<in_t *>(smem_raw + S.tile);
    in_t *list = reinterpret_cast<in_t *>(smem_raw + S.list);
    acc_t *part_s = reinterpret_cast<acc_t *>(smem_raw + S.part);
    acc_t *mine_s = reinterpret_cast<acc_t *>(smem_raw + S.mine);
    acc_t *theirs_s = reinterpret_cast<acc_t *>(smem_raw + S.theirs);
    int *xst_s = reinterpret_cast<int *>(smem_raw + S.xst);
    acc_t *qr_s = reinterpret_cast<acc_t *>(smem_raw + S.qr);
    acc_t *wq = reinterpret_cast<acc_t *>(smem_raw + S.wq);
    int *wi = reinterpret_cast<int *>(smem_raw + S.wi);
    unsigned *ww = reinterpret_cast<unsigned *>(smem_raw + S.ww);
    uint32_t *maskw = reinterpret_cast<uint32_t *>(smem_raw + S.maskw);
    int *stw_s = reinterpret_cast<int *>(smem_raw + S.misc);
    int *sel_s = stw_s + 1;
    uint8_t *keep_s = smem_raw + S.keep;
    u64 *bar = reinterpret_cast<u64 *>(smem_raw + S.bar);

    // this thread's unit: column c of the CTA's 64, rows 8k + 2*part + {0, 1}
    const int cc = lane & 7;
    W3Ctx w;
    w.N = N;
    w.part = lane >> 3;
    w.lane_base = cc;
    w.cmask = 0x01010101u << cc;
    const int part = w.part;
    const int c = warp * 8 + cc;                  // column within the CTA
    const int gc = (int)rank * W3_COLS + c;       // column within the round
    in_t *lc = list + c;
    // landing-buffer addresses of this thread's two row phases (128-byte swizzle: the 16-byte chunk index of a row
    // is XOR-ed with the row number mod 8, which is 2*part / 2*part + 1 for every row this thread reads)
    const int c32 = c & 31;
    const in_t *src0 = tile_s + (c >> 5) * (W3_BOX_BYTES / 4) + (2 * part) * 32 + ((((c32 >> 2) ^ (2 * part)) << 2) | (c32 & 3));
    const in_t *src1 = tile_s + (c >> 5) * (W3_BOX_BYTES / 4) + (2 * part + 1) * 32 + ((((c32 >> 2) ^ (2 * part + 1)) << 2) | (c32 & 3));

    const uint32_t tile_bytes = 2 * W3_BOX_BYTES;
    if (tid == 0) {
        if (smem_u32(tile_s) & 1023u) __trap();      // the swizzle pattern is a function of the address bits
        mbar_init(bar, 1);
        if (cluster_id < p.T) {
            mbar_expect_tx(bar, tile_bytes);
            const int row0 = (int)(cluster_id * p.N);
            tma_load_box(tile_s, &tmap, (int)rank * W3_COLS, row0, bar);
            tma_load_box(tile_s + W3_BOX_BYTES / 4, &tmap, (int)rank * W3_COLS + 32, row0, bar);
        }
    }
    __syncthreads();
    // the peer must be running before anything is written into its shared memory
    cluster_sync_all();

    const bool shape2 = second_pass_shape_panic(p);
    const bool high = p.tie == SVOC_TIE_HIGH_INDEX_KEPT;
    const uint32_t theirs_remote = map_to_peer(smem_u32(theirs_s), peer);
    const uint32_t xst_remote = map_to_peer(smem_u32(xst_s), peer);

    uint32_t it = 0;
    for (i64 r = cluster_id; r < p.T; r += n_clusters, ++it) {
        const uint32_t par = it & 1u;
        const in_t *gcol = reinterpret_cast<const in_t *>(p.pred) + (size_t)r * p.N * W3_D + gc;
        if (tid == 0) *stw_s = STW_OK;
        // ---- phase 0: the round's values into registers
        mbar_wait(bar, par);
        in_t x[W3_NV];
#pragma unroll
        for (int k = 0; k < W3_NV / 2; ++k) {
            x[2 * k] = src0[k * 256];             // 8 rows of 32 words
            x[2 * k + 1] = src1[k * 256];
        }
        // ---- phase 1a: key scale from a 64-row sample (rows 0..63: the first 16 values of every part) and the keys
        bool col_const = false, col_fb = false, refined1 = false;
        in_t lo = std::numeric_limits<in_t>::max(), hi = std::numeric_limits<in_t>::lowest();
        typename Key::Scale sc{};
#pragma unroll
        for (int j = 0; j < 16; ++j) {
            lo = vmin(lo, x[j]);
            hi = vmax(hi, x[j]);
        }
        lo = w3_col_reduce(w, lo, [](in_t a, in_t b) { return vmin(a, b); });
        hi = w3_col_reduce(w, hi, [](in_t a, in_t b) { return vmax(a, b); });
        if (!(hi > lo)) {                         // degenerate sample: exact sweep, which also detects a constant column
#pragma unroll
            for (int j = 0; j < W3_NV; ++j) {
                if (FULL || w3_row(j, part) < N) {
                    lo = vmin(lo, x[j]);
                    hi = vmax(hi, x[j]);
                }
            }
            lo = w3_col_reduce(w, lo, [](in_t a, in_t b) { return vmin(a, b); });
            hi = w3_col_reduce(w, hi, [](in_t a, in_t b) { return vmax(a, b); });
        }
        col_const = !(hi > lo);
        col_fb = !col_const && !Key::make(lo, hi, sc);
        if (col_fb) {
            // No keys will be built, so nothing below has read values 16..63 yet.  They must have ARRIVED in their
            // registers before the barrier that hands the landing buffer to the next copy: fold them into a word
            // the compiler has to produce (the store never happens).
            uint32_t fold = 0;
#pragma unroll
            for (int j = 16; j < W3_NV; ++j) fold ^= okey(x[j]);
            if (fold == 0x9e3779b9u && p.T < 0) *sel_s = (int)fold;
        }
        unsigned kw[W3_NKW];
#pragma unroll
        for (int q = 0; q < W3_NKW; ++q) kw[q] = 0u;
        if (!col_const && !col_fb) w3_build_keys<Tr, true, FULL, false>(x, kw, N, part, lo, hi, sc);
        // every value has been consumed: the landing buffer is free for the next round, whose copy now has the
        // whole of this round's computation to arrive
        __syncthreads();
        if (tid == 0 && r + n_clusters < p.T) {
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            mbar_expect_tx(bar, tile_bytes);
            const int row0 = (int)((r + n_clusters) * p.N);
            tma_load_box(tile_s, &tmap, (int)rank * W3_COLS, row0, bar);
            tma_load_box(tile_s + W3_BOX_BYTES / 4, &tmap, (int)rank * W3_COLS + 32, row0, bar);
        }
        // ---- the two passes share ONE inlined copy of the centre code (the kernel is 250 KB of SASS otherwise):
        //      pass 0 = first-pass centre, risks, row selection, key patching; pass 1 = second-pass centre
        val_t e1 = 0;
#pragma unroll 1
        for (int pass = 0;; ++pass) {
            bool refined = false;
            const val_t ctr = w3_centre<Tr, FULL>(w, x, kw, gcol, pass ? keep_s : nullptr, pass ? R : N,
                                                  pass ? p.centre2 : p.centre, col_const, col_fb, lo, lc, refined);
            if (pass) {
                if (part == 0) reinterpret_cast<val_t *>(p.consensus)[(size_t)r * W3_D + gc] = ctr;
                break;
            }
            e1 = ctr;
            refined1 = refined;
            if (part == 0 && p.essence1) reinterpret_cast<val_t *>(p.essence1)[(size_t)r * W3_D + gc] = e1;

        // ---- phase 2: quadratic risk.  A thread holds (x - e1)^2 of 64 rows of ONE column; the 8 columns of the
        //      warp are folded with a transposing shuffle tree (lane bit 0 keeps the rows of its own parity, bits 1
        //      and 2 the next two row-index bits), 8 values at a time, which leaves every lane with 8 row sums over
        //      the warp's columns; warp partials meet in shared memory, summed in a fixed order.
        {
            const bool b0 = cc & 1, b1 = cc & 2, b2 = cc & 4;
            bool bad = false;
#pragma unroll
            for (int n = 0; n < 8; ++n) {          // values 8n .. 8n+7: rows 32n + 8*(0..3) + 2*part + {0,1}
                acc_t u[4];
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const in_t xa = x[8 * n + 2 * i], xb = x[8 * n + 2 * i + 1];
                    if constexpr (Tr::fixed) {
                        bool ba, bb;
                        if (p.constrained) {
                            ba = (xa < 0) | ((i64)xa > WSAD);
                            bb = (xb < 0) | ((i64)xb > WSAD);
                        } else {
                            ba = ((i64)xa > ENGINE_RANGE) | ((i64)xa < -ENGINE_RANGE);
                            bb = ((i64)xb > ENGINE_RANGE) | ((i64)xb < -ENGINE_RANGE);
                        }
                        if (!FULL) {
                            ba &= w3_row(8 * n + 2 * i, part) < N;
                            bb &= w3_row(8 * n + 2 * i + 1, part) < N;
                        }
                        bad |= ba | bb;
                    }
                    const acc_t da = Tr::dev(xa, e1), db = Tr::dev(xb, e1);
                    const acc_t keep = b0 ? db : da, send = b0 ? da : db;
                    u[i] = keep + __shfl_xor_sync(0xffffffffu, send, 1);
                }
                acc_t t[2];
#pragma unroll
                for (int i = 0; i < 2; ++i) {
                    const acc_t keep = b1 ? u[2 * i + 1] : u[2 * i], send = b1 ? u[2 * i] : u[2 * i + 1];
                    t[i] = keep + __shfl_xor_sync(0xffffffffu, send, 2);
                }
                const acc_t keep = b2 ? t[1] : t[0], send = b2 ? t[0] : t[1];
                const acc_t s = keep + __shfl_xor_sync(0xffffffffu, send, 4);
                // this lane's row: value index j = 8n + 4*b2 + 2*b1 + b0
                const int row = 32 * n + 8 * ((b2 ? 2 : 0) + (b1 ? 1 : 0)) + 2 * part + (b0 ? 1 : 0);
                part_s[warp * W3_ROWS + row] = s;
            }
            if (bad) atomicMin(stw_s, p.constrained ? STW_INTERVAL : STW_ENGINE_RANGE);
        }
        __syncthreads();
        {
            // this CTA's 64-column partial of row `tid`, to both CTAs of the pair
            acc_t q = part_s[tid];
#pragma unroll
            for (int k = 1; k < 8; ++k) q += part_s[k * W3_ROWS + tid];
            mine_s[par * W3_ROWS + tid] = q;
            st_peer(theirs_remote + (uint32_t)((par * W3_ROWS + tid) * sizeof(acc_t)), q);
            if (tid == 0) st_peer(xst_remote + par * 4u, *stw_s);
        }
        cluster_sync_all();
        {
            const acc_t a = mine_s[par * W3_ROWS + tid], b = theirs_s[par * W3_ROWS + tid];
            const acc_t q = rank == 0 ? a + b : b + a;        // partial(rank 0) + partial(rank 1) on both CTAs
            qr_s[tid] = q;
            if (rank == 0 && p.qr && tid < N) reinterpret_cast<acc_t *>(p.qr)[(size_t)r * p.N + tid] = q;
            if (tid == 0) {
                const int other = xst_s[par];
                if (other != STW_OK) atomicMin(stw_s, other);
            }
        }
        __syncthreads();

        // ---- phase 3: which rows are kept — warp 0 (w2_select_rows) while warp 1 does the first-pass reliability;
        //      exact all-pairs ranking when the selection gives up on this round's risks.  Both CTAs of the pair
        //      compute the same thing from the same risks; rank 0 writes the per-round outputs.
        {
            constexpr int NS = W3_ROWS / 32;
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
                    __syncwarp();
                    // keep flags as bytes for the exact fallbacks: bit b of a mask byte -> byte b (0 / 1)
                    const unsigned bits8 = (maskw[lane >> 2] >> (8 * (lane & 3))) & 0xffu;
                    const unsigned lo4 = ((bits8 & 0xfu) * 0x00204081u) & 0x01010101u;
                    const unsigned hi4 = ((bits8 >> 4) * 0x00204081u) & 0x01010101u;
                    *reinterpret_cast<uint2 *>(keep_s + 8 * lane) = make_uint2(lo4, hi4);
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
                        const i64 rel1 = reliability_fixed(s1 / (u64)N, p, STW_SQRT_REL1, stw);
                        if (rel1 < 0 || rel1 > WSAD) stw = min(stw, STW_REL1_RANGE);
                        if (rank == 0) reinterpret_cast<i64 *>(p.rel1)[r] = rel1;
                    } else {
                        const val_t rel1 = reliability_float<val_t>(s1 / (val_t)N, p);
                        if (!(rel1 >= 0 && rel1 <= 1)) stw = min(stw, STW_REL1_RANGE);
                        if (rank == 0) reinterpret_cast<val_t *>(p.rel1)[r] = rel1;
                    }
                    if (stw != STW_OK) atomicMin(stw_s, stw);
                }
            }
            __syncthreads();
            if (!*sel_s) {
                // exact ranking of row `tid` under (risk asc, index desc | asc): rare
                const acc_t mineq = qr_s[tid];
                int rk = 0;
                for (int j = 0; j < N; ++j) {
                    const acc_t q = qr_s[j];
                    rk += (q < mineq) | ((q == mineq) & (high ? j > tid : j < tid));
                }
                const bool kept = tid < N && rk < R;
                keep_s[tid] = kept ? 1 : 0;
                const unsigned bal = __ballot_sync(0xffffffffu, kept);
                if (lane == 0) maskw[warp] = bal;
                __syncthreads();
            }
            // second-pass reliability, mask and status
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
                    if (stw != STW_OK) atomicMin(stw_s, stw);
                    if (rank == 0) {         // nothing later contributes to the status
                        const int sw = *stw_s;
                        p.status[r] = (sw == STW_OK) ? SVOC_ST_OK : (sw & 0xff);
                    }
                }
            }
        }
        // ---- phase 4 set-up: keys of the kept rows only; the loop then runs the second-pass centre
            if (shape2) {
                if (part == 0) reinterpret_cast<val_t *>(p.consensus)[(size_t)r * W3_D + gc] = 0;
                break;
            }
            if (!col_const && !col_fb) {
                if (refined1) {
                    // the first pass re-keyed the column: back to the first-pass scale, dropped rows -> sentinel
#pragma unroll
                    for (int q = 0; q < W3_NKW; ++q) kw[q] = 0u;
                    w3_build_keys<Tr, true, FULL, false>(x, kw, N, part, lo, hi, sc);
                }
                // dropped rows of this thread -> sentinel key.  Key word q holds rows 32*(q>>1) + 16*(q&1) +
                // {0, 1, 8, 9} + 2*part: two bit pairs of mask word q>>1, spread to four bytes by one multiply
#pragma unroll
                for (int g = 0; g < W3_NKW / 2; ++g) {
                    const unsigned drop = ~maskw[g] >> (2 * part);
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        const unsigned t = drop >> (16 * h);
                        const unsigned nib = (t & 3u) | ((t >> 6) & 0xcu);
                        kw[2 * g + h] |= ((nib * 0x00204081u) & 0x01010101u) * 0x7fu;
                    }
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

template <class Tr>
int launch_wide3(const Params &p, const Plan &pl, cudaStream_t st) {
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
    void (*kern)(const Params, const CUtensorMap) = (p.N == W3_ROWS) ? consensus_wide3_kernel<Tr, true> : consensus_wide3_kernel<Tr, false>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl.smem);
    if (e != cudaSuccess) return fail(std::string("cudaFuncSetAttribute: ") + cudaGetErrorString(e));
    kern<<<(unsigned)pl.grid, W3_THREADS, pl.smem, st>>>(p, tmap);
    e = cudaGetLastError();
    if (e != cudaSuccess) return fail(std::string("kernel launch: ") + cudaGetErrorString(e));
    count_launch();
    return 0;
}

}  // namespace svoc
