// walk_tail.cuh -- what the last CTAs of a walk launch do after their own tallies: hand the launch's ticket
// slot back, zero the tally buffer of the next launch, and -- on N > 1 GPUs -- run the path's one collective
// (the sum of the per-start scores and counters over ranks, the all-reduce a multi-process run(args) of
// fluence_mini_app.py:397-438 needs once per solve) inside the same kernel, over NVLink peer memory.
//
// Every CTA ends with "fence, count myself done".  The CTA that counts LAST knows that all tallies of this rank
// are in the rank's buffer; the TAIL_HELPERS - 1 CTAs that counted just before it stay as helpers (they spin on a
// word of the ticket slot; every other CTA has exited, so there is always a free SM for a CTA still to come).
//   (1) last CTA, world > 1: publishes "epoch E complete" into every peer's flag array (st.release.sys) and waits
//       for the peers' flags (ld.acquire.sys; %globaltimer time-out instead of a hang), then releases the helpers;
//   (2) last CTA + helpers: every thread sums ONE 16-byte pair of the vector over all peers, the loads of all
//       peers issued before any is consumed -- the whole 80 KB vector in one NVLink round trip, where one CTA
//       alone needs five -- writes the sum to this rank's output region and zeroes the pair of the NEXT launch's
//       buffer (two buffers alternate by epoch parity: a buffer is rewritten only after the wait of the following
//       exchange, which proves that every peer has finished reading it -- so callers never launch a memset);
//   (3) the CTA that finishes last of these resets the ticket slot (no memset per launch) and writes the status.
// A separate reduction kernel cost a launch boundary, its stream gap and the rank skew it exposes: 28 us of a
// 0.55 ms step at N = 8 (round 1); here the exchange starts the moment the last warp retires.
#pragma once
#include "common.cuh"

namespace nrw {

constexpr int TAIL_HELPERS = 8;        // CTAs that run the exchange (the last one to finish its walk + 7 before it)
constexpr int TICKET_WORDS = 4;        // ticket slot: {work_next, done, go, helpers_done}

struct TailParams {
    unsigned long long* slot;                      // this launch's ticket slot (TICKET_WORDS words, all zero when free)
    int world, rank;                               // world <= 1: no exchange
    long long* const* bufs;                        // [world] every rank's tally buffer as mapped into this process
    unsigned long long* const* flags;              // [world] every rank's flag array (uint64[world])
    long long* own;                                // this rank's tally buffer (== bufs[rank]; NULL: resolved on the device)
    long long off, zero_off, out_off, n;           // tallies of this launch: own[off, off + n); zero own[zero_off, +n) (< 0: none);
                                                   // sums over ranks land in own[out_off, out_off + n)
    unsigned long long epoch;
    long long* status;                             // 0 ok, 1 a peer timed out (may be NULL)
    unsigned long long timeout_ns;
};

__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v) {
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_gpu(unsigned long long* p, unsigned long long v) {
    asm volatile("st.release.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_gpu(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ longlong2 ld_relaxed_sys_v2(const long long* p) {      // p 16-byte aligned
    longlong2 v;
    asm volatile("ld.relaxed.sys.global.v2.s64 {%0, %1}, [%2];" : "=l"(v.x), "=l"(v.y) : "l"(p));
    return v;
}
__device__ __forceinline__ unsigned long long globaltimer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

// thread `tid` < world of ONE CTA: publish `epoch` into word `rank` of every peer's flag array, then wait for peer
// `tid`'s word in this rank's array.  Returns after a __syncthreads(); false when some peer timed out.
__device__ __forceinline__ bool signal_and_wait(const TailParams& r, int tid, int* s_bad) {
    if (tid == 0) *s_bad = 0;
    __syncthreads();
    if (tid < r.world) {
        __threadfence_system();
        st_release_sys(r.flags[tid] + r.rank, r.epoch);
        const unsigned long long* mine = r.flags[r.rank] + tid;
        const unsigned long long t0 = globaltimer_ns();
        while (ld_acquire_sys(mine) < r.epoch) {
            if (globaltimer_ns() - t0 > r.timeout_ns) { *s_bad = 1; break; }
            __nanosleep(20);
        }
    }
    __syncthreads();
    return *s_bad == 0;
}

// Pairs part, part + parts, ... of the vector: sum over all peers (world > 1) into this rank's output region, and
// zero the same pairs of the next launch's buffer.  n is taken up to the next even count (the buffers are allocated
// that way and the pad word stays zero) so that everything moves as 16-byte pairs.
__device__ __forceinline__ void sum_and_zero(const TailParams& r, long long* own, long long part, long long parts) {
    long long* mine_next = (r.zero_off >= 0 && own) ? own + r.zero_off : nullptr;
    const long long pairs = (r.n + 1) / 2;
    for (long long i = part; i < pairs; i += parts) {
        if (r.world > 1) {
            longlong2 acc = make_longlong2(0, 0);
            for (int q0 = 0; q0 < r.world; q0 += 8) {
                longlong2 v[8];
#pragma unroll
                for (int k = 0; k < 8; ++k)
                    v[k] = (q0 + k < r.world) ? ld_relaxed_sys_v2(r.bufs[q0 + k] + r.off + 2 * i) : make_longlong2(0, 0);
#pragma unroll
                for (int k = 0; k < 8; ++k) { acc.x += v[k].x; acc.y += v[k].y; }
            }
            *reinterpret_cast<longlong2*>(own + r.out_off + 2 * i) = acc;
        }
        if (mine_next) *reinterpret_cast<longlong2*>(mine_next + 2 * i) = make_longlong2(0, 0);
    }
}

// Called by every thread of every CTA as the last statement of a walk kernel, after the CTA's tallies have
// been issued.  slot[0] is the launch's ticket counter (fetch_range).
__device__ __noinline__ void walk_tail(const TailParams& r) {
    __shared__ int s_idx, s_bad;
    __shared__ unsigned long long s_go;
    const int tid = threadIdx.x, nthr = (int)blockDim.x;
    unsigned long long* done = r.slot + 1;
    unsigned long long* go = r.slot + 2;
    unsigned long long* hdone = r.slot + 3;
    const bool work = r.world > 1 || r.zero_off >= 0;          // anything to do beyond handing the slot back?
    const int helpers = work ? min(TAIL_HELPERS, (int)gridDim.x) : 1;
    __threadfence();                               // this thread's tallies (RED / atomics) before the CTA counts itself done
    __syncthreads();
    if (tid == 0) s_idx = (int)atomicAdd(done, 1ull) - ((int)gridDim.x - helpers);
    __syncthreads();
    const int hidx = s_idx;                        // helpers - 1: this CTA counted last
    if (hidx < 0) return;
    bool ok = true;
    if (hidx == helpers - 1) {
        __threadfence();                           // every CTA's tallies are visible from here on
        if (r.world > 1) ok = signal_and_wait(r, tid, &s_bad);
        if (tid == 0 && helpers > 1) st_release_gpu(go, ok ? r.epoch : ~0ull);
    } else {
        if (tid == 0) {
            unsigned long long g;
            while ((g = ld_acquire_gpu(go)) == 0ull) __nanosleep(20);
            s_go = g;
        }
        __syncthreads();
        ok = s_go != ~0ull;
    }
    long long* own = r.own ? r.own : (r.world > 1 ? r.bufs[r.rank] : nullptr);
    // a peer never arrived: the sums would be wrong and the next buffer may still be read by a peer -- leave
    // every buffer as it is and poison the status word (callers test it per solve)
    if (ok && work) sum_and_zero(r, own, (long long)hidx * nthr + tid, (long long)helpers * nthr);
    __threadfence();
    __syncthreads();
    if (tid == 0) {
        const unsigned long long h = helpers > 1 ? atomicAdd(hdone, 1ull) : 0ull;
        if (h == (unsigned long long)helpers - 1ull) {          // the last of the tail CTAs: the slot is free again
            if (r.status) *r.status = ok ? 0 : 1;
            r.slot[0] = 0ull; *done = 0ull; *go = 0ull; *hdone = 0ull;
        }
    }
}

}  // namespace nrw
