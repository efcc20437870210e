// walk_tail.cuh -- what the LAST CTA of a walk launch does after its own tallies: hand the launch's ticket
// slot back, zero the tally buffer of the next launch, and -- on N > 1 GPUs -- run the path's one collective
// (the sum of the per-start scores and counters over ranks, the all-reduce a multi-process run(args) of
// fluence_mini_app.py:397-438 needs once per solve) inside the same kernel, over NVLink peer memory.
//
// Every CTA ends with "fence, count myself done"; the CTA that counts last knows that all tallies of this
// rank are in the rank's buffer.  It then
//   (1) resets the launch's ticket counter and done counter (the slot is reusable; no memset per launch),
//   (2) world > 1: publishes "epoch E complete" into every peer's flag array (st.release.sys), waits for the
//       peers' flags (ld.acquire.sys; %globaltimer time-out instead of a hang), issues the loads of all peers'
//       vectors before consuming any (one NVLink round trip per group of 8 peers) and writes the sums plus a
//       status word to `out`,
//   (3) zeroes this rank's buffer of the NEXT launch (two buffers alternate by epoch parity: a buffer is
//       rewritten only after the wait of the following reduction, which proves that every peer has finished
//       reading it -- so one signal / wait per step suffices and callers never launch a memset).
// A separate reduction kernel cost a launch boundary, its stream gap and the rank skew it exposes: 28 us of a
// 0.55 ms step at N = 8 (round 1); here the exchange starts the moment the last warp retires.
#pragma once
#include "common.cuh"

namespace nrw {

struct TailParams {
    unsigned long long* done;                      // CTAs-finished counter of this launch's ticket slot
    int world, rank;                               // world <= 1: no exchange
    const long long* const* bufs;                  // [world] every rank's tally buffer as mapped into this process
    unsigned long long* const* flags;              // [world] every rank's flag array (uint64[world])
    long long* own;                                // this rank's tally buffer (== bufs[rank])
    long long off, zero_off, n;                    // tallies of this launch: own[off, off + n); zero own[zero_off, +n) (< 0: none)
    unsigned long long epoch;
    long long* out;                                // [n] sums over ranks (world > 1)
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
__device__ __forceinline__ long long ld_relaxed_sys(const long long* p) {
    long long v;
    asm volatile("ld.relaxed.sys.global.s64 %0, [%1];" : "=l"(v) : "l"(p));
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

// Signal / wait / sum / zero, executed by ONE CTA (`nthr` threads, thread index `tid`); shared by the walk's
// tail and the stand-alone reduction kernel (p2p_reduce.cuh).  Returns false on this CTA when a peer timed out.
__device__ __forceinline__ bool exchange_and_sum(const TailParams& r, int tid, int nthr, int* s_bad) {
    if (tid == 0) *s_bad = 0;
    __syncthreads();
    if (r.world > 1) {
        if (tid < r.world) {
            __threadfence_system();
            st_release_sys(r.flags[tid] + r.rank, r.epoch);                   // tell peer `tid`: my epoch is complete
            const unsigned long long* mine = r.flags[r.rank] + tid;           // ... and wait for peer `tid`'s word
            const unsigned long long t0 = globaltimer_ns();
            while (ld_acquire_sys(mine) < r.epoch) {
                if (globaltimer_ns() - t0 > r.timeout_ns) { *s_bad = 1; break; }
                __nanosleep(32);
            }
        }
        __syncthreads();
        if (*s_bad) {
            // a peer never arrived: the sums would be wrong and the next buffer may still be read by a peer.
            // Poison the status word (callers test it per solve) and leave every buffer as it is.
            if (tid == 0 && r.status) *r.status = 1;
            return false;
        }
    }
    long long* own = r.own ? r.own : (r.world > 1 ? const_cast<long long*>(r.bufs[r.rank]) : nullptr);
    long long* mine_next = (r.zero_off >= 0 && own) ? own + r.zero_off : nullptr;
    if (r.world > 1) {
        // both halves 16-byte aligned: pairs of int64 per load
        const long long pairs = (((r.off | (r.zero_off < 0 ? 0 : r.zero_off)) & 1) == 0) ? r.n / 2 : 0;
        for (long long i = tid; i < pairs; i += nthr) {
            longlong2 acc = make_longlong2(0, 0);
            for (int q0 = 0; q0 < r.world; q0 += 8) {
                longlong2 v[8];
#pragma unroll
                for (int k = 0; k < 8; ++k)
                    v[k] = (q0 + k < r.world) ? ld_relaxed_sys_v2(r.bufs[q0 + k] + r.off + 2 * i) : make_longlong2(0, 0);
#pragma unroll
                for (int k = 0; k < 8; ++k) { acc.x += v[k].x; acc.y += v[k].y; }
            }
            *reinterpret_cast<longlong2*>(r.out + 2 * i) = acc;
            if (mine_next) *reinterpret_cast<longlong2*>(mine_next + 2 * i) = make_longlong2(0, 0);
        }
        for (long long i = 2 * pairs + tid; i < r.n; i += nthr) {
            long long acc = 0;
            for (int q = 0; q < r.world; ++q) acc += ld_relaxed_sys(r.bufs[q] + r.off + i);
            r.out[i] = acc;
            if (mine_next) mine_next[i] = 0;
        }
        if (tid == 0 && r.status) *r.status = 0;
    } else if (mine_next) {
        for (long long i = tid; i < r.n; i += nthr) mine_next[i] = 0;
    }
    return true;
}

// Called by every thread of every CTA as the last statement of a walk kernel, after the CTA's tallies have
// been issued.  `work_next` is the launch's ticket counter (fetch_range).
__device__ __forceinline__ void walk_tail(const TailParams& r, unsigned long long* work_next) {
    __shared__ int s_last, s_bad;
    const int tid = threadIdx.x;
    __threadfence();                               // this thread's tallies (RED / atomics) before the CTA counts itself done
    __syncthreads();
    if (tid == 0) {
        const unsigned long long d = atomicAdd(r.done, 1ull);
        s_last = (d == (unsigned long long)gridDim.x - 1ull) ? 1 : 0;
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();                               // the other CTAs' tallies are visible from here on
    if (tid == 0) { *work_next = 0ull; *r.done = 0ull; }      // the ticket slot is ready for a later launch
    if (r.world > 1 || r.zero_off >= 0) exchange_and_sum(r, tid, (int)blockDim.x, &s_bad);
}

}  // namespace nrw
