// common.cuh -- shared device helpers for libnrw (sm_100a only).
//
// Philox2x32-10 (Salmon et al., "Parallel random numbers: as easy as 1, 2, 3", SC'11) is the
// counter-based generator behind the on-device walk, key = hash(seed).  It replaces the reference's
// single sequential MT19937 stream (fluence_mini_app.py:243, :523) in the statistical ("philox")
// mode; the bit-exact mode consumes the reference's own stream instead (walk_replay.cuh).
//
// Draw of walker `id` for model step t (1-based), by table layout:
//   * WIDE alias tables and the CDF search: one call per step, counter (id, t); the alias slot comes
//     from word 0, the coin from word 1; the CDF search uses both as one 53-bit uniform.
//   * COMPACT alias tables ("paired draws"): one call per PAIR of steps, counter (id, (t-1)|1); step
//     t uses word 0 when t is odd, word 1 when even.  The word w and the row length n give
//     v = w * n (64 bit): slot = v >> 32, coin = v & 0xffffffff, primary target iff coin <= threshold.
//     Rows have n <= 1023 entries, so slot and coin deviate from exact uniformity by < n / 2^32
//     (2.4e-7 relative at worst, 2e-9 on a 9-point stencil) -- far below any Monte-Carlo error reachable.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace nrw {

#define NRW_FULL_MASK 0xffffffffu

constexpr uint32_t PHILOX_M2x32 = 0xD256D193u;
constexpr uint32_t PHILOX_W32   = 0x9E3779B9u;

// 10 rounds; the key schedule depends only on the (kernel-uniform) key, so the compiler
// keeps the ten round keys in uniform registers: per round one IMAD.WIDE.U32 + one LOP3.
__device__ __forceinline__ uint2 philox2x32_10(uint32_t c0, uint32_t c1, uint32_t key) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint32_t hi = __umulhi(PHILOX_M2x32, c0);
        const uint32_t lo = PHILOX_M2x32 * c0;
        c0 = hi ^ key ^ c1;
        c1 = lo;
        key += PHILOX_W32;
    }
    return make_uint2(c0, c1);
}

// counter word 1 of the call that serves model step t (t >= 1, < 2^24) under the paired scheme, and the
// word of its result that step t uses
__host__ __device__ __forceinline__ uint32_t pair_counter(uint32_t t) { return (t - 1u) | 1u; }
__host__ __device__ __forceinline__ uint32_t pair_word(uint2 r, uint32_t t) { return (t & 1u) ? r.x : r.y; }

// 64-bit seed -> 32-bit Philox key (host mirror: nrw/philox.py::seed_to_key).
__host__ __device__ __forceinline__ uint32_t seed_to_key(uint64_t seed) {
    uint64_t z = seed + 0x9E3779B97F4A7C15ull;          // splitmix64 finaliser
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    z = z ^ (z >> 31);
    return (uint32_t)(z ^ (z >> 32));
}

// ---- COMPACT row descriptor (include/nrw.h).
//   w0 = off[0:15) | n[15:25) | detabs[25] | term[26] | cls[27:29) | inc[29:32)
//   w1 = base[0:29) | detlast[29] | sinc[30:32)
// detabs marks a row whose single successor is absorbing; term = (n == 0) | detabs: a walker landing here ends in
// this iteration; cls = source class of the row; detlast: the absorbing successor is the last row (the fluence
// sink).  inc and sinc are what ARRIVING on the row adds to a walker's score and step count when rows with an
// absorbing successor are stepped without a draw: inc = cls + (detabs ? class of the successor : 0),
// sinc = 1 + detabs.  They sit in the top bits so that one shift-add each (LEA.HI) applies them, and so that a
// base carrying its flag bits still addresses the right descriptor: (base + offset) << 3 drops bits 29..31
// (tables.py::build_alias).
constexpr uint32_t ROW_OFF_MASK = 0x00007fffu;
constexpr uint32_t ROW_N_MASK   = 0x01ff8000u;
constexpr uint32_t ROW_DETABS   = 0x02000000u;
constexpr uint32_t ROW_TERM     = 0x04000000u;
constexpr uint32_t ROW_BASE_MASK = 0x1fffffffu;
constexpr uint32_t ROW_DETLAST  = 0x20000000u;              // in w1
__host__ __device__ __forceinline__ uint32_t row_off(uint32_t w0) { return w0 & ROW_OFF_MASK; }
__host__ __device__ __forceinline__ uint32_t row_len(uint32_t w0) { return (w0 >> 15) & 0x3ffu; }
__host__ __device__ __forceinline__ uint32_t row_class(uint32_t w0) { return (w0 >> 27) & 3u; }
__host__ __device__ __forceinline__ uint32_t row_inc(uint32_t w0) { return w0 >> 29; }
__host__ __device__ __forceinline__ uint32_t row_detcls(uint32_t w0) { return (w0 >> 29) - ((w0 >> 27) & 3u); }
__host__ __device__ __forceinline__ uint32_t row_sinc(uint32_t w1) { return w1 >> 30; }
__host__ __device__ __forceinline__ uint32_t row_base(uint32_t w1) { return w1 & ROW_BASE_MASK; }

// numpy legacy random_sample(): (a >> 5, b >> 6) -> 53-bit double in [0, 1).
__device__ __forceinline__ double u53(uint32_t a, uint32_t b) {
    return ((double)(a >> 5) * 67108864.0 + (double)(b >> 6)) * (1.0 / 9007199254740992.0);
}

// searchsorted(side='left') on an ascending fp64 slice: first i with cum[i] >= u.
// Returns n when u is above every entry (the reference then raises IndexError,
// fluence_mini_app.py:244-245; callers clamp to n-1 -- documented deviation).
__device__ __forceinline__ int lower_bound_f64(const double* __restrict__ cum, int n, double u) {
    int lo = 0, hi = n;
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (__ldg(cum + mid) < u) lo = mid + 1; else hi = mid;
    }
    return lo;
}

// One CSR draw with the reference's semantics (Q6 of SURVEY.md): u <= 0 -> dense column 0.
__device__ __forceinline__ int sample_cdf(const int32_t* __restrict__ indptr,
                                          const int32_t* __restrict__ cols,
                                          const double* __restrict__ cum, int row, double u) {
    const int lo = __ldg(indptr + row), hi = __ldg(indptr + row + 1);
    int i = lower_bound_f64(cum + lo, hi - lo, u);
    if (i >= hi - lo) i = hi - lo - 1;
    int c = __ldg(cols + lo + i);
    if (u <= 0.0) c = 0;
    return c;
}

__device__ __forceinline__ uint32_t lanemask_lt() {
    uint32_t m;
    asm volatile("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
    return m;
}

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return (uint32_t)__cvta_generic_to_shared(p);
}
// 8-byte shared-memory load from a 32-bit shared-window address (lets callers do the address arithmetic in 32
// bits: an index whose top bits carry flags addresses the right entry once it is shifted left by 3)
__device__ __forceinline__ uint2 lds64(uint32_t addr) {
    uint2 v;
    asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(addr));
    return v;
}

// fire-and-forget increment of a 32-bit shared-memory word at a 32-bit shared-window address
__device__ __forceinline__ void red_shared_inc(uint32_t addr) {
    asm volatile("red.shared.add.u32 [%0], 1;" ::"r"(addr) : "memory");
}

// ---- mbarrier + 1-D TMA bulk copy (global -> shared), SASS: UBLKCP ---------------------------
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
                 "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t}" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
// bytes must be a multiple of 16; src/dst 16-byte aligned.
__device__ __forceinline__ void tma_bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes,
                                             uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::
            "r"(smem_u32(dst_smem)),
        "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
        : "memory");
}

}  // namespace nrw
