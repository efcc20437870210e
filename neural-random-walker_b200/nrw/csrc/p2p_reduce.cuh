// p2p_reduce.cuh -- the path's one collective without NCCL: sum of the per-start score vectors over the
// GPUs of one NVSwitch box, done by ONE kernel per rank over peer memory.
//
// Every rank owns a buffer (its walk kernel's scores) and a flag array in symmetric memory, mapped into
// all peers (torch.distributed._symmetric_memory does the allocation and the handle exchange:
// nrw/dist.py).  The kernel (1) publishes "my scores of epoch E are complete" into every peer's flag
// array, (2) waits until every peer has published epoch E, (3) reads all peers' vectors over NVLink and
// writes the sum to a local output.  80 KB per peer: the cost is one NVLink round trip plus the signal,
// where NCCL's all-reduce of the same vector costs 28 us at N = 8.
//
// Callers alternate between two score buffers (epoch parity): a rank re-zeroes buffer E only after the
// wait of its reduce E+1, which proves that every peer has finished reading buffer E -- so no second
// barrier is needed.  The kernel does that zeroing itself (zero_offset), which also saves the caller a
// memset launch per step.  A peer that never shows up trips a clock64() time-out that raises *err instead of hanging.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace nrw {

constexpr int P2P_THREADS = 512;

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

// bufs[q] / flags[q]: rank q's score buffer and flag array as mapped into THIS process.
__global__ void __launch_bounds__(P2P_THREADS)
p2p_allreduce_i64_kernel(long long* __restrict__ out, const long long* const* __restrict__ bufs,
                         unsigned long long* const* __restrict__ flags, long long offset, long long zero_offset,
                         long long n, int rank, int world, unsigned long long epoch, long long timeout_cycles,
                         int* __restrict__ err) {
    __shared__ int s_bad;
    const int tid = threadIdx.x;
    if (tid == 0) s_bad = 0;
    __syncthreads();
    if (blockIdx.x == 0 && tid < world) {
        __threadfence_system();
        st_release_sys(flags[tid] + rank, epoch);               // tell peer `tid`: my epoch is complete
    }
    if (tid < world) {
        const unsigned long long* mine = flags[rank] + tid;     // ... and wait for peer `tid`'s word
        const long long t0 = clock64();
        while (ld_acquire_sys(mine) < epoch) {
            if (clock64() - t0 > timeout_cycles) { s_bad = 1; break; }
            __nanosleep(64);
        }
    }
    __syncthreads();
    if (s_bad) {
        if (tid == 0) atomicExch(err, 1);
        return;
    }
    // all peers' loads of an element pair are issued before any is consumed: one NVLink round trip per
    // group of 8 peers instead of one per peer
    long long* mine_next = zero_offset >= 0 ? const_cast<long long*>(bufs[rank]) + zero_offset : nullptr;
    const long long pairs = (((offset | (zero_offset < 0 ? 0 : zero_offset)) & 1) == 0) ? n / 2 : 0;   // 16-byte aligned halves only
    for (long long i = (long long)blockIdx.x * P2P_THREADS + tid; i < pairs; i += (long long)gridDim.x * P2P_THREADS) {
        longlong2 acc = make_longlong2(0, 0);
        for (int q0 = 0; q0 < world; q0 += 8) {
            longlong2 v[8];
#pragma unroll
            for (int k = 0; k < 8; ++k)
                v[k] = (q0 + k < world) ? ld_relaxed_sys_v2(bufs[q0 + k] + offset + 2 * i) : make_longlong2(0, 0);
#pragma unroll
            for (int k = 0; k < 8; ++k) { acc.x += v[k].x; acc.y += v[k].y; }
        }
        *reinterpret_cast<longlong2*>(out + 2 * i) = acc;
        if (mine_next) *reinterpret_cast<longlong2*>(mine_next + 2 * i) = make_longlong2(0, 0);
    }
    for (long long i = 2 * pairs + (long long)blockIdx.x * P2P_THREADS + tid; i < n; i += (long long)gridDim.x * P2P_THREADS) {
        long long acc = 0;
        for (int q = 0; q < world; ++q) acc += ld_relaxed_sys(bufs[q] + offset + i);
        out[i] = acc;
        if (mine_next) mine_next[i] = 0;
    }
}

}  // namespace nrw
