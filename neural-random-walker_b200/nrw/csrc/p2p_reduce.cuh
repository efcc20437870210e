// p2p_reduce.cuh -- the path's collective as a stand-alone kernel: sum of int64 tally vectors over the GPUs of
// one NVSwitch box by ONE CTA per rank over peer memory (no NCCL).  The walk kernels run the same exchange in
// their own tail (walk_tail.cuh, nrw_walk.reduce); this entry serves tallies that no walk launch produced in
// place (tests, tools/allreduce_latency.py, tallies assembled by several launches).
//
// Every rank owns a buffer and a flag array in symmetric memory, mapped into all peers
// (torch.distributed._symmetric_memory does the allocation and the handle exchange: nrw/dist.py).  Protocol,
// buffer alternation by epoch parity and time-out behaviour: walk_tail.cuh (exchange_and_sum).  The sums land in
// the symmetric buffer itself (every rank pushes its slice into every peer's output region).
#pragma once
#include "walk_tail.cuh"

namespace nrw {

constexpr int P2P_THREADS = 1024;
constexpr int P2P_CTAS = 8;            // one 16-byte pair per thread: the whole vector in one NVLink round trip

// Every CTA signals / awaits on its own words of the flag arrays (word cta * world + peer), so the CTAs need no
// hand-off between them; the flag arrays hold P2P_CTAS * world words.
__global__ void __launch_bounds__(P2P_THREADS)
p2p_allreduce_i64_kernel(TailParams r, int* __restrict__ err) {
    __shared__ int s_bad;
    const int tid = (int)threadIdx.x;
    if (tid == 0) s_bad = 0;
    __syncthreads();
    if (tid < r.world) {
        const int w = (int)blockIdx.x * r.world;
        __threadfence_system();
        st_release_sys(r.flags[tid] + w + r.rank, r.epoch);
        const unsigned long long* mine = r.flags[r.rank] + w + tid;
        const unsigned long long t0 = globaltimer_ns();
        while (ld_acquire_sys(mine) < r.epoch) {
            if (globaltimer_ns() - t0 > r.timeout_ns) { s_bad = 1; break; }
            __nanosleep(20);
        }
    }
    __syncthreads();
    if (s_bad) {
        if (tid == 0 && err) atomicExch(err, 1);
        return;
    }
    sum_and_zero(r, r.bufs[r.rank], (long long)blockIdx.x * P2P_THREADS + tid, (long long)gridDim.x * P2P_THREADS);
}

}  // namespace nrw
