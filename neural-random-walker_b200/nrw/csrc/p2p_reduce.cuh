// p2p_reduce.cuh -- the path's collective as a stand-alone kernel: sum of int64 tally vectors over the GPUs of
// one NVSwitch box by ONE CTA per rank over peer memory (no NCCL).  The walk kernels run the same exchange in
// their own tail (walk_tail.cuh, nrw_walk.reduce); this entry serves tallies that no walk launch produced in
// place (tests, tools/allreduce_latency.py, tallies assembled by several launches).
//
// Every rank owns a buffer and a flag array in symmetric memory, mapped into all peers
// (torch.distributed._symmetric_memory does the allocation and the handle exchange: nrw/dist.py).  Protocol,
// buffer alternation by epoch parity and time-out behaviour: walk_tail.cuh (exchange_and_sum).
#pragma once
#include "walk_tail.cuh"

namespace nrw {

constexpr int P2P_THREADS = 1024;

__global__ void __launch_bounds__(P2P_THREADS)
p2p_allreduce_i64_kernel(const TailParams r, int* __restrict__ err) {
    __shared__ int s_bad;
    if (!exchange_and_sum(r, (int)threadIdx.x, P2P_THREADS, &s_bad) && threadIdx.x == 0 && err) atomicExch(err, 1);
}

}  // namespace nrw
