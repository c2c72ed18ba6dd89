// consensus_mid.cuh — the `mid` variant of the fused consensus round: 32 < N <= 64 oracles, D <= 32
// dimensions, 32-bit elements (int32 fixed point / fp32).  This is the kernel BASELINE config 3 (N = 64,
// D = 16, fixed point, bit-exact) runs.
//
// The `small<64>` kernel it replaces gave every (round, dimension) column to one thread, which sorted the
// 64 values with a 543-comparator network — twice per round — and ranked the 64 risks by all-pairs counting:
// 5.7 k warp instructions per round, 177 lane operations per element, 6.6 % of the HBM roofline
// (profiles/r01n_small64_c3_full.txt).  Here ONE WARP owns a round:
//   * columns: lane = part * (32 / P) + column, P lanes per column (P = 2 for D = 16: 32 rows each); the column
//     centre is found like in `wide3` — monotone 7-bit keys packed in registers, SWAR bisection with the counts
//     combined by shuffles, the bucket resolved exactly from a handful of values (sel_centre, consensus_wide3.cuh);
//   * risks: a lane owns two rows and walks their D elements with a per-lane rotation (conflict-free);
//   * the kept rows come from the single-warp bisection on the risks (w2_select_rows), all-pairs as its fallback;
//   * the two Newton square roots of a round (math.cairo:271-292, data-dependent trip counts) are NOT run inside the
//     round — one busy lane would hold 31 idle ones for ~1000 instructions — but collected per CTA (128 rounds) and
//     done at the end, one task per lane.
// A CTA of 8 warps takes 128 consecutive rounds; every warp stages its own rounds with its own TMA bulk copies.
// Moments, rank output and mean centres stay with `small<64>`.
// Reference semantics: see consensus_kernel.cuh (contract.cairo:442-503 etc.).
#pragma once
#include "consensus_wide3.cuh"

namespace svoc {

constexpr int MID_WARPS = 8;         // warps (= rounds in flight) per CTA
constexpr int MID_ITERS = 16;        // rounds per warp: 128 rounds per CTA, so the square-root tail has one task per thread
constexpr int MID_ROUNDS = MID_WARPS * MID_ITERS;
constexpr int MID_ROWS = 64;

template <int P>
struct MidGeo {
    static constexpr int LANES = P;
    static constexpr int CW = 32 / P;              // column slots per warp = lane distance between the parts of a column
    static constexpr int NV = MID_ROWS / P;        // values per lane
    static constexpr int NKW = NV / 4;
    static constexpr int MW = (NKW + 7) / 8;
    static constexpr int SLOTS = MID_ROWS;
    // 64 values per column: 6-bit keys (about two values per bucket near the median) and the next FOUR buckets searched
    // for the value above the median.  With 7 bits most buckets are empty and the scan over all values that replaces an
    // empty window ran in nearly every warp; with 5 bits the buckets held 4-5 values and their exact ranking cost as
    // much as the sorting network this kernel replaces (profiles/r02h, r02j).
    static constexpr int KBITS = 6;
    static constexpr int NEXT = 4;
};

// a lane's values of its column in the staged round: rows P*j + part.  STRIDE = P * D when that is a compile-time
// constant (D = 32 / P: every load then has an immediate offset), 0 = run-time stride
template <int P, int STRIDE, class in_t>
struct MidVals {
    const in_t *base;        // tile + part * D + column
    int stride;              // P * D
    int part;
    __device__ __forceinline__ in_t operator()(int j) const { return base[j * (STRIDE ? STRIDE : stride)]; }
    __device__ __forceinline__ int row(int j) const { return P * j + part; }
};

struct MidSmem {
    size_t tile, c1, qr, list, patch, keep, wq, wi, ww, sum1, sum2, stw, bar, total;
};
__host__ __device__ inline MidSmem mid_smem_layout(int N, int D, size_t vs) {
    MidSmem s{};
    size_t off = 0;
    auto take = [&](size_t bytes, size_t align) { off = (off + align - 1) & ~(align - 1); const size_t at = off; off += bytes; return at; };
    s.tile = take((size_t)MID_WARPS * N * D * 4, 128);
    s.c1 = take((size_t)MID_WARPS * 32 * vs, 16);
    s.qr = take((size_t)MID_WARPS * MID_ROWS * vs, 16);
    s.list = take((size_t)MID_WARPS * 32 * W3_LSTRIDE * 4, 16);
    s.patch = take((size_t)MID_WARPS * MID_ROWS, 16);
    s.keep = take((size_t)MID_WARPS * MID_ROWS, 16);
    s.wq = take((size_t)MID_WARPS * W2_RCAP * vs, 16);
    s.wi = take((size_t)MID_WARPS * W2_RCAP * 4, 16);
    s.ww = take((size_t)MID_WARPS * 16, 16);
    s.sum1 = take((size_t)MID_ROUNDS * vs, 16);
    s.sum2 = take((size_t)MID_ROUNDS * vs, 16);
    s.stw = take((size_t)MID_ROUNDS * 4, 16);
    s.bar = take(8 * MID_WARPS, 16);
    s.total = off;
    return s;
}

// DEXACT: D == 32 / P exactly (config 3: D = 16, P = 2) — strides and the risk walk are compile-time then
template <class Tr, int P, bool FULL, bool DEXACT>
__global__ void __launch_bounds__(MID_WARPS * 32, 3) consensus_mid_kernel(const Params p) {
    using in_t = typename Tr::in_t;
    using val_t = typename Tr::val_t;
    using acc_t = typename Tr::acc_t;
    using Key = WideKey<Tr>;
    using G = MidGeo<P>;
    constexpr int NV = G::NV, NKW = G::NKW, CW = G::CW;
    extern __shared__ __align__(128) unsigned char smem_mid[];
    const int N = FULL ? MID_ROWS : p.N, D = DEXACT ? CW : p.D, R = p.R;
    const MidSmem S = mid_smem_layout(p.N, D, sizeof(acc_t));
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const size_t round_elems = (size_t)p.N * D;

    in_t *tile_s = reinterpret_cast<in_t *>(smem_mid + S.tile);
    val_t *c1_w = reinterpret_cast<val_t *>(smem_mid + S.c1) + warp * 32;
    acc_t *qr_w = reinterpret_cast<acc_t *>(smem_mid + S.qr) + warp * MID_ROWS;
    in_t *list_w = reinterpret_cast<in_t *>(smem_mid + S.list) + warp * 32 * W3_LSTRIDE;
    uint8_t *patch_w = smem_mid + S.patch + warp * MID_ROWS;
    uint8_t *keep_w = smem_mid + S.keep + warp * MID_ROWS;
    acc_t *wq = reinterpret_cast<acc_t *>(smem_mid + S.wq) + warp * W2_RCAP;
    int *wi = reinterpret_cast<int *>(smem_mid + S.wi) + warp * W2_RCAP;
    unsigned *ww = reinterpret_cast<unsigned *>(smem_mid + S.ww) + warp * 4;
    acc_t *sum1_s = reinterpret_cast<acc_t *>(smem_mid + S.sum1);
    acc_t *sum2_s = reinterpret_cast<acc_t *>(smem_mid + S.sum2);
    int *stw_s = reinterpret_cast<int *>(smem_mid + S.stw);
    u64 *bar = reinterpret_cast<u64 *>(smem_mid + S.bar);

    const i64 r_first = (i64)blockIdx.x * MID_ROUNDS;
    const int n_here = (int)min((i64)MID_ROUNDS, p.T - r_first);        // rounds of this CTA

    // column unit of this lane
    const int cq = lane % CW;
    W3Ctx w;
    w.N = N;
    w.part = lane / CW;
    w.lane_base = cq;
    w.cmask = 0;
#pragma unroll
    for (int k = 0; k < P; ++k) w.cmask |= 1u << (cq + CW * k);
    const int part = w.part;
    const bool has_unit = cq < D;
    in_t *lcw = list_w + cq * W3_LSTRIDE;
    const bool shape2 = second_pass_shape_panic(p);
    const bool high = p.tie == SVOC_TIE_HIGH_INDEX_KEPT;
    const int rot = (32 % D == 0) ? lane / (32 / D) : lane;            // per-lane rotation of the risk walk

    for (int k = tid; k < MID_ROUNDS; k += blockDim.x) stw_s[k] = STW_OK;
    if (tid < MID_WARPS) mbar_init(bar + tid, 1);
    __syncthreads();

    // Every warp stages its OWN rounds (one TMA bulk copy per round into the warp's slot of the tile, completing on the
    // warp's mbarrier) and issues the copy of its next round as soon as it has finished the current one: no CTA-wide
    // barrier inside the loop, and the copy's latency is covered by the other warps of the SM.
    const in_t *g_first = reinterpret_cast<const in_t *>(p.pred) + (size_t)r_first * round_elems;
    const uint32_t rbytes = (uint32_t)(round_elems * sizeof(in_t));
    const bool bulk = ((reinterpret_cast<uintptr_t>(g_first) & 15) == 0) && ((rbytes & 15) == 0);
    in_t *tile = tile_s + (size_t)warp * round_elems;
    auto stage = [&](int slot) {          // this warp's round `slot` -> its tile slot
        const in_t *src = g_first + (size_t)slot * round_elems;
        if (bulk) {
            if (lane == 0) {
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                mbar_expect_tx(bar + warp, rbytes);
                bulk_g2s(tile, src, rbytes, bar + warp);
            }
        } else {
            for (int e = lane; e < (int)round_elems; e += 32) tile[e] = src[e];
        }
    };
    if (warp < n_here) stage(warp);
    uint32_t phase = 0;
    for (int it = 0; it * MID_WARPS < n_here; ++it) {
        const int nr = min(MID_WARPS, n_here - it * MID_WARPS);         // rounds of this batch
        const in_t *gsrc = g_first + (size_t)(it * MID_WARPS) * round_elems;
        if (warp < nr) {
            const int slot = it * MID_WARPS + warp;
            const i64 r = r_first + slot;
            if (bulk) mbar_wait(bar + warp, phase & 1u);
            ++phase;
            __syncwarp();
            const in_t *gcol = gsrc + (size_t)warp * round_elems + cq;
            const MidVals<P, DEXACT ? 32 : 0, in_t> val{tile + part * D + cq, P * D, part};
            // ---- first-pass centre of the lane's column
            bool col_const = false, col_fb = false, refined1 = false, bad = false;
            in_t lo = std::numeric_limits<in_t>::max(), hi = std::numeric_limits<in_t>::lowest();
            typename Key::Scale sc{};
            unsigned kw[NKW];
#pragma unroll
            for (int q = 0; q < NKW; ++q) kw[q] = 0u;
            if (has_unit) {
                in_t x[NV];
#pragma unroll
                for (int j = 0; j < NV; ++j) {
                    const bool ok = FULL || val.row(j) < N;
                    x[j] = ok ? val(j) : in_t(0);
                    if (ok) {
                        lo = vmin(lo, x[j]);
                        hi = vmax(hi, x[j]);
                    }
                }
                lo = sel_col_reduce<G>(w, lo, [](in_t a, in_t b) { return vmin(a, b); });
                hi = sel_col_reduce<G>(w, hi, [](in_t a, in_t b) { return vmax(a, b); });
                col_const = !(hi > lo);
                col_fb = !col_const && !Key::make(lo, hi, sc);
                if constexpr (Tr::fixed) {          // the interval check (math.cairo:298-310) on the column's extremes
                    if (p.constrained) bad = (lo < 0) | ((i64)hi > WSAD);
                    else bad = ((i64)hi > ENGINE_RANGE) | ((i64)lo < -ENGINE_RANGE);
                }
                if (!col_const && !col_fb) sel_build_keys<Tr, G, true, FULL, false, true>(x, val, kw, N, lo, hi, sc);
                const val_t e1 = sel_centre<Tr, G, FULL>(w, val, kw, gcol, D, nullptr, N, p.centre, col_const, col_fb, lo, lcw, refined1);
                if (part == 0) {
                    c1_w[cq] = e1;
                    if (p.essence1) reinterpret_cast<val_t *>(p.essence1)[(size_t)r * D + cq] = e1;
                }
            }
            __syncwarp();
            // ---- quadratic risk.  Fixed form: a lane owns rows lane and lane + 32 and walks their D elements with a
            //      per-lane rotation (conflict-free; integer sums do not depend on the order).  Float form: identical
            //      rows must get bit-identical risks (colluders tie exactly, and the tie rule decides), so every row is
            //      summed in the SAME order: column-wise, (x - e1)^2 of the lane's own column folded over the column
            //      lanes by a transposing shuffle tree, which leaves lane cq with values 2cq, 2cq + 1 of its part.
            if constexpr (Tr::fixed) {
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    const int row = lane + 32 * h;
                    if (row < N) {
                        const in_t *xr = tile + (size_t)row * D;
                        acc_t q = 0;
                        if (p.constrained) {
                            // |x - centre| <= 1e6: 32-bit squares (wmul_sq_small), 32-bit sum (at most 32 * 1 000 001)
                            uint32_t q32 = 0;
                            if (DEXACT) {
#pragma unroll
                                for (int k = 0; k < CW; ++k) {
                                    const int dd = (rot + k) & (CW - 1);
                                    const int diff = (int)xr[dd] - (int)c1_w[dd];
                                    q32 += wmul_sq_small((uint32_t)(diff < 0 ? -diff : diff));
                                }
                            } else {
                                int dd = rot % D;
                                for (int k = 0; k < D; ++k) {
                                    const int diff = (int)xr[dd] - (int)c1_w[dd];
                                    q32 += wmul_sq_small((uint32_t)(diff < 0 ? -diff : diff));
                                    dd = (dd + 1 == D) ? 0 : dd + 1;
                                }
                            }
                            q = q32;
                        } else {
                            int dd = rot % D;
                            for (int k = 0; k < D; ++k) {
                                q += Tr::dev(xr[dd], c1_w[dd]);
                                dd = (dd + 1 == D) ? 0 : dd + 1;
                            }
                        }
                        qr_w[row] = q;
                        if (p.qr) reinterpret_cast<acc_t *>(p.qr)[(size_t)r * p.N + row] = q;
                    }
                }
            } else {
                acc_t v[NV];
                const val_t e1 = has_unit ? c1_w[cq] : val_t(0);
#pragma unroll
                for (int j = 0; j < NV; ++j) v[j] = (has_unit && (FULL || val.row(j) < N)) ? Tr::dev(val(j), e1) : acc_t(0);
                int lanebit = CW / 2;
#pragma unroll
                for (int n = NV / 2; n >= 2; n /= 2, lanebit >>= 1) {
                    const bool up = cq & lanebit;
#pragma unroll
                    for (int i = 0; i < n; ++i) {
                        const acc_t keep = up ? v[i + n] : v[i], send = up ? v[i] : v[i + n];
                        v[i] = keep + __shfl_xor_sync(0xffffffffu, send, lanebit);
                    }
                }
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    const int row = P * (2 * cq + h) + part;
                    if (row < N) {
                        qr_w[row] = v[h];
                        if (p.qr) reinterpret_cast<acc_t *>(p.qr)[(size_t)r * p.N + row] = v[h];
                    }
                }
            }
            if constexpr (Tr::fixed) {
                if (__any_sync(0xffffffffu, bad) && lane == 0) atomicMin(stw_s + slot, p.constrained ? STW_INTERVAL : STW_ENGINE_RANGE);
            }
            __syncwarp();
            // ---- which rows are kept
            unsigned words[2];
            acc_t qrow[2];
            qrow[0] = qr_w[lane];
            qrow[1] = lane + 32 < N ? qr_w[lane + 32] : acc_t(0);
            bool fast = w2_select_rows<acc_t, 2>(qrow, N, R, high, wq, wi, ww, words);
            if (!fast) {          // exact ranking under (risk asc, index desc | asc)
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    const int row = lane + 32 * h;
                    int rk = 0;
                    for (int j = 0; j < N; ++j) {
                        const acc_t q = qr_w[j];
                        rk += (q < qrow[h]) | ((q == qrow[h]) & (high ? j > row : j < row));
                    }
                    words[h] = __ballot_sync(0xffffffffu, row < N && rk < R);
                }
            }
            {
                acc_t s1 = (lane < N ? qrow[0] : acc_t(0)) + (lane + 32 < N ? qrow[1] : acc_t(0));
                acc_t s2 = (((words[0] >> lane) & 1u) ? qrow[0] : acc_t(0)) + (((words[1] >> lane) & 1u) ? qrow[1] : acc_t(0));
                for (int o = 16; o; o >>= 1) {
                    s1 += __shfl_xor_sync(0xffffffffu, s1, o);
                    s2 += __shfl_xor_sync(0xffffffffu, s2, o);
                }
                if (lane == 0) {
                    sum1_s[slot] = s1;
                    sum2_s[slot] = s2;
                }
                if (p.reliable && lane < p.words) p.reliable[(size_t)r * p.words + lane] = words[lane];
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    const int row = lane + 32 * h;
                    const bool kept = (words[h] >> lane) & 1u;
                    keep_w[row] = kept ? 1 : 0;
                    patch_w[(row % P) * NV + row / P] = kept ? 0 : 0x7f;
                }
            }
            __syncwarp();
            // ---- second-pass centre over the kept rows
            if (has_unit) {
                val_t cons = 0;
                if (!shape2) {
                    if (!col_const && !col_fb) {
                        if (refined1) {
#pragma unroll
                            for (int q = 0; q < NKW; ++q) kw[q] = 0u;
                            sel_build_keys<Tr, G, true, FULL, false, false>(nullptr, val, kw, N, lo, hi, sc);
                        }
#pragma unroll
                        for (int q = 0; q < NKW; ++q) kw[q] |= reinterpret_cast<const unsigned *>(patch_w + part * NV)[q];
                    }
                    bool refined2 = false;
                    cons = sel_centre<Tr, G, FULL>(w, val, kw, gcol, D, keep_w, R, p.centre2, col_const, col_fb, lo, lcw, refined2);
                }
                if (part == 0) reinterpret_cast<val_t *>(p.consensus)[(size_t)r * D + cq] = cons;
            }
            __syncwarp();             // every lane has made its last read of the slot
            if (slot + MID_WARPS < n_here) stage(slot + MID_WARPS);
        }
    }
    __syncthreads();

    // ---- the reliabilities of the CTA's rounds, one task (round, pass) per thread: two Newton square roots per round
    for (int task = tid; task < 2 * n_here; task += blockDim.x) {
        const int slot = task >> 1, second = task & 1;
        const acc_t sum = second ? sum2_s[slot] : sum1_s[slot];
        const int cnt = second ? R : N;
        int stw = STW_OK;
        if (second && shape2) stw = STW_SECOND_SHAPE;
        if constexpr (Tr::fixed) {
            i64 rel = 0;
            if (!(second && shape2)) {
                rel = reliability_fixed(sum / (u64)cnt, p, second ? STW_SQRT_REL2 : STW_SQRT_REL1, stw);
                if (rel < 0 || rel > WSAD) stw = min(stw, second ? STW_REL2_RANGE : STW_REL1_RANGE);
            }
            reinterpret_cast<i64 *>(second ? p.rel2 : p.rel1)[r_first + slot] = rel;
        } else {
            val_t rel = 0;
            if (!(second && shape2)) {
                rel = reliability_float<val_t>(sum / (val_t)cnt, p);
                if (!(rel >= 0 && rel <= 1)) stw = min(stw, second ? STW_REL2_RANGE : STW_REL1_RANGE);
            }
            reinterpret_cast<val_t *>(second ? p.rel2 : p.rel1)[r_first + slot] = rel;
        }
        if (stw != STW_OK) atomicMin(stw_s + slot, stw);
    }
    __syncthreads();
    for (int k = tid; k < n_here; k += blockDim.x) {
        const int sw = stw_s[k];
        p.status[r_first + k] = (sw == STW_OK) ? SVOC_ST_OK : (sw & 0xff);
    }
}

template <class Tr, int P>
int launch_mid_p(const Params &p, const Plan &pl, cudaStream_t st) {
    const bool dexact = p.D == 32 / P;
    void (*kern)(const Params) = (p.N == MID_ROWS) ? (dexact ? consensus_mid_kernel<Tr, P, true, true> : consensus_mid_kernel<Tr, P, true, false>)
                                                   : (dexact ? consensus_mid_kernel<Tr, P, false, true> : consensus_mid_kernel<Tr, P, false, false>);
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl.smem);
    if (e != cudaSuccess) return fail(std::string("cudaFuncSetAttribute: ") + cudaGetErrorString(e));
    // several CTAs per SM are planned: ask for the full shared-memory carveout (the default is not always enough)
    cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, (int)cudaSharedmemCarveoutMaxShared);
    kern<<<(unsigned)pl.grid, MID_WARPS * 32, pl.smem, st>>>(p);
    e = cudaGetLastError();
    if (e != cudaSuccess) return fail(std::string("kernel launch: ") + cudaGetErrorString(e));
    count_launch();
    return 0;
}

template <class Tr>
int launch_mid(const Params &p, const Plan &pl, cudaStream_t st) {
    switch (pl.wP) {
        case 1: return launch_mid_p<Tr, 1>(p, pl, st);
        case 2: return launch_mid_p<Tr, 2>(p, pl, st);
        default: return launch_mid_p<Tr, 4>(p, pl, st);
    }
}

}  // namespace svoc
