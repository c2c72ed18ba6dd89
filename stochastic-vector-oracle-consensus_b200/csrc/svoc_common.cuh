// svoc_common.cuh — shared device helpers: wsad fixed-point arithmetic, order-preserving
// keys, sorting networks, TMA bulk-copy wrappers.  sm_100a only.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <utility>
#include "svoc.h"

namespace svoc {

typedef unsigned long long u64;
typedef long long i64;
typedef unsigned __int128 u128;
typedef __int128 i128;

constexpr i64 WSAD = 1000000;       // contract/src/signed_decimal.cairo:82
constexpr i64 HALF_WSAD = 500000;   // :83
constexpr i64 ENGINE_RANGE = (i64)1 << 30;

// ---------------------------------------------------------------- status words
// A round's status is the FIRST panic in the contract's evaluation order
// (contract.cairo:442-503).  Threads of a round publish (priority << 8 | code) with
// atomicMin; the smallest priority wins.
__host__ __device__ constexpr int st_word(int prio, int code) { return (prio << 8) | code; }
constexpr int STW_OK = 0x7fffffff;
constexpr int STW_INTERVAL = st_word(1, SVOC_ST_INTERVAL_INPUT);
constexpr int STW_ENGINE_RANGE = st_word(1, SVOC_ST_ENGINE_RANGE);
constexpr int STW_FIRST_SHAPE = st_word(2, SVOC_ST_FIRST_PASS_SHAPE);
constexpr int STW_SQRT_REL1 = st_word(3, SVOC_ST_SQRT_REL1);
constexpr int STW_SPREAD_ZERO = st_word(4, SVOC_ST_MAX_SPREAD_ZERO);
constexpr int STW_REL1_RANGE = st_word(5, SVOC_ST_REL1_RANGE);
constexpr int STW_SECOND_SHAPE = st_word(6, SVOC_ST_SECOND_PASS_SHAPE);
constexpr int STW_SQRT_REL2 = st_word(7, SVOC_ST_SQRT_REL2);
constexpr int STW_REL2_RANGE = st_word(9, SVOC_ST_REL2_RANGE);
constexpr int STW_SKEW = st_word(10, SVOC_ST_SKEW_PANIC);
constexpr int STW_KURT = st_word(11, SVOC_ST_KURT_PANIC);

// ---------------------------------------------------------------- wsad arithmetic (device)
// signed_decimal.cairo:52-63: I128Div divides magnitudes and re-applies the sign, i.e.
// truncation toward zero — exactly C++'s `/` on signed integers.

// wsad_mul(x, x) for |x| <= 2^31: the product is non-negative, so the rounding add and
// the division are unsigned (signed_decimal.cairo:110-112, math.cairo:170-173).
__device__ __forceinline__ u64 wmul_sq(i64 diff) {
    u64 a = (u64)(diff < 0 ? -diff : diff);
    return (a * a + (u64)HALF_WSAD) / (u64)WSAD;
}

// The same for |diff| <= 1 048 999 (constrained inputs: both the value and the centre lie in [0, 1e6]) in 32-bit
// arithmetic only: with a = 1000 a1 + a0,  a^2 = 1e6 a1^2 + 2000 a1 a0 + a0^2, so
//     (a^2 + 500000) / 1e6 = a1^2 + (2000 a1 a0 + a0^2 + 500000) / 1e6        (truncating divisions)
// and the second numerator stays below 2^31 (2000 * 1048 * 999 + 999^2 + 5e5 = 2 095 402 001).  Ten 32-bit
// instructions instead of a 64-bit multiply and a 64-bit division by a constant; exhaustively checked against the
// 64-bit form over the whole range in tests/test_host_logic.py.
__host__ __device__ __forceinline__ uint32_t wmul_sq_small(uint32_t a) {
    const uint32_t a1 = a / 1000u, a0 = a - a1 * 1000u;
    return a1 * a1 + (2000u * a1 * a0 + a0 * a0 + 500000u) / 1000000u;
}

// wsad_div(a, b) = (a*W + b/2) / b with i128 intermediates (signed_decimal.cairo:114-116).
__device__ __forceinline__ i64 wdiv_wide(i64 a, i64 b) {
    i128 num = (i128)a * (i128)WSAD + (i128)(b / 2);
    return (i64)(num / (i128)b);
}

// Non-negative fast path: a >= 0, b > 0, a*W + b/2 < 2^64.
__device__ __forceinline__ u64 wdiv_u64(u64 a, u64 b) { return (a * (u64)WSAD + (b >> 1)) / b; }

// math.cairo:271-292 — Newton iteration in wsad with the 50-iteration cap.
// value >= 0.  Returns false when the contract would divide by zero (value == 1).
__device__ inline bool wsqrt(u64 value, u64 &out) {
    if (value == 0) { out = 0; return true; }
    u64 g = value >> 1;
    u64 g2 = g + (u64)WSAD;
    const bool narrow = value <= 9000000000000ull;  // value*W + g/2 stays below 2^64
    // value*W + g/2 below 2^53 (g never exceeds max(value/2, W) + a little): numerator and divisor are exact
    // doubles, and a quotient rounded toward zero truncates to the exact integer quotient — a dozen FP64
    // instructions instead of the ~100 of the 64-bit integer division
    const bool tiny = value < 8000000000ull;
    for (int i = 0; i < 50; ++i) {
        if (g == g2) break;
        if (g == 0) return false;
        u64 n;
#ifdef __CUDA_ARCH__
        if (tiny) n = (u64)__ddiv_rz((double)(value * (u64)WSAD + (g >> 1)), (double)g);
        else
#endif
        if (narrow) n = wdiv_u64(value, g);
        else n = (u64)(((u128)value * (u128)WSAD + (u128)(g >> 1)) / (u128)g);
        g2 = g;
        g = (g + n) >> 1;
    }
    out = g;
    return true;
}

// ---------------------------------------------------------------- order-preserving keys
__device__ __forceinline__ uint32_t okey(int32_t v) { return (uint32_t)v ^ 0x80000000u; }
__device__ __forceinline__ u64 okey(i64 v) { return (u64)v ^ 0x8000000000000000ull; }
__device__ __forceinline__ uint32_t okey(float v) {
    uint32_t b = __float_as_uint(v);
    return b ^ ((b >> 31) ? 0xffffffffu : 0x80000000u);
}
__device__ __forceinline__ u64 okey(double v) {
    u64 b = (u64)__double_as_longlong(v);
    return b ^ ((b >> 63) ? 0xffffffffffffffffull : 0x8000000000000000ull);
}
__device__ __forceinline__ void unkey(uint32_t k, int32_t &v) { v = (int32_t)(k ^ 0x80000000u); }
__device__ __forceinline__ void unkey(u64 k, i64 &v) { v = (i64)(k ^ 0x8000000000000000ull); }
__device__ __forceinline__ void unkey(uint32_t k, float &v) {
    v = __uint_as_float(k ^ ((k >> 31) ? 0x80000000u : 0xffffffffu));
}
__device__ __forceinline__ void unkey(u64 k, double &v) {
    v = __longlong_as_double((i64)(k ^ ((k >> 63) ? 0x8000000000000000ull : 0xffffffffffffffffull)));
}

// ---------------------------------------------------------------- sorting network
// Batcher's merge-exchange network on a register array of power-of-two length; every
// index is a compile-time constant after unrolling, so the array never leaves registers.
template <class T>
__host__ __device__ __forceinline__ void cswap(T &a, T &b) {
    T lo = a < b ? a : b;
    T hi = a < b ? b : a;
    a = lo;
    b = hi;
}
// fp32: two FMNMX instead of a compare and two selects
__host__ __device__ __forceinline__ void cswap(float &a, float &b) {
    const float lo = fminf(a, b), hi = fmaxf(a, b);
    a = lo;
    b = hi;
}

// The comparator list is built at compile time and applied through an index sequence so
// that every array index is a constant expression (plain unrolled loops left the array in
// local memory for N >= 16).
template <int N>
struct NetPairs {
    static constexpr int cap = N * 10;
    unsigned char lo[cap] = {}, hi[cap] = {};
    int n = 0;
    constexpr NetPairs() {
        static_assert((N & (N - 1)) == 0 && N <= 128, "power of two, at most 128");
        for (int p = 1; p < N; p <<= 1)
            for (int k = p; k >= 1; k >>= 1)
                for (int j = k % p; j <= N - 1 - k; j += 2 * k)
                    for (int i = 0; i < k; ++i)
                        if (i + j + k < N && ((i + j) / (2 * p) == (i + j + k) / (2 * p))) {
                            lo[n] = (unsigned char)(i + j);
                            hi[n] = (unsigned char)(i + j + k);
                            ++n;
                        }
    }
};

template <class T, int N, size_t... I>
__host__ __device__ __forceinline__ void network_apply(T (&a)[N], std::index_sequence<I...>) {
    constexpr NetPairs<N> P{};
    (cswap(a[P.lo[I]], a[P.hi[I]]), ...);
}

template <class T, int N>
__host__ __device__ __forceinline__ void network_sort(T (&a)[N]) {
    constexpr NetPairs<N> P{};
    network_apply<T, N>(a, std::make_index_sequence<(size_t)P.n>{});
}

template <class T, int N, size_t... I>
__host__ __device__ __forceinline__ T pick_impl(const T (&a)[N], int k, std::index_sequence<I...>) {
    T v = a[0];
    ((v = ((int)I == k) ? a[I] : v), ...);
    return v;
}

// a[k] for a run-time k without dynamic register indexing
template <class T, int N>
__host__ __device__ __forceinline__ T pick(const T (&a)[N], int k) {
    return pick_impl<T, N>(a, k, std::make_index_sequence<N>{});
}

// ---------------------------------------------------------------- TMA bulk copy (1-D)
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

// A load through the 32-bit shared-memory window.  A pointer into a staged tile reaches the kernels' helper
// functions as a GENERIC pointer: every element then costs a 64-bit address (multiply, carry, two moves) before
// its LD; with the window address it is one integer multiply-add and an LDS.
__device__ __forceinline__ void lds_at(float &v, uint32_t addr) { asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(addr)); }
__device__ __forceinline__ void lds_at(double &v, uint32_t addr) { asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(addr)); }
__device__ __forceinline__ void lds_at(int32_t &v, uint32_t addr) { asm volatile("ld.shared.s32 %0, [%1];" : "=r"(v) : "r"(addr)); }
__device__ __forceinline__ void lds_at(i64 &v, uint32_t addr) { asm volatile("ld.shared.s64 %0, [%1];" : "=l"(v) : "r"(addr)); }

__device__ __forceinline__ void mbar_init(u64 *bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(u64 *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, uint32_t bytes, u64 *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
// The same wait for a barrier that usually completes MUCH later (a tile copy, a peer CTA's data): poll with a
// growing nanosleep in between.  try_wait with a suspend-time hint came back after ~30 ns each time here (ncu opcode
// histogram of wide3: ~90 SYNCS.TRYWAIT + NANOSLEEP.SYNCS + SYNCS.PHASECHK + BRA per warp and round, a tenth of
// the issue slots of an issue-bound kernel); a sleeping warp costs no issue slot.
#ifndef SVOC_WAIT_CAP_NS
#define SVOC_WAIT_CAP_NS 256u
#endif
__device__ __forceinline__ bool mbar_test(u64 *bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0u;
}
__device__ __forceinline__ void mbar_wait_sleepy(u64 *bar, uint32_t parity) {
    if (mbar_test(bar, parity)) return;
    unsigned ns = 32u;
    do {
        __nanosleep(ns);
        ns = ns < SVOC_WAIT_CAP_NS ? 2u * ns : SVOC_WAIT_CAP_NS;
    } while (!mbar_test(bar, parity));
}
__device__ __forceinline__ void mbar_wait(u64 *bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}

}  // namespace svoc
