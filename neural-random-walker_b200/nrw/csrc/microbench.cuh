// microbench.cuh -- denominators of the memory/atomic roofline, measured on the device the walk runs
// on ("measure BW_L2 and R_atom with micro-benchmarks on the box first").
//
// Each kernel mimics ONE resource use of the walk at the walk's own geometry (148 CTAs x 1024
// threads, one CTA per SM) with everything else stripped away:
//   MB_SMEM_ATOMIC  random atomicAdd on a per-CTA shared-memory histogram of n bins
//                   (the occupancy tally of walk_philox.cuh, TL_HIST_S)
//   MB_SMEM_LOAD    dependent random 8-byte shared-memory loads (the alias-entry / row-descriptor
//                   look-ups: two per walker-step)
//   MB_L2_LOAD      dependent random 8-byte global loads from a table of n entries that is L2-resident
//                   (row descriptors in TAB_HYB / TAB_GLB)
//   MB_GLOBAL_ATOMIC random 64-bit atomicAdd (RED) on a global histogram of n bins (TL_HIST_G)
// The index stream is a 32-bit LCG per thread (2 instructions), so the measured rate is the
// resource's, not the generator's.
#pragma once
#include "common.cuh"

namespace nrw {

enum : int { MB_SMEM_ATOMIC = 0, MB_SMEM_LOAD = 1, MB_L2_LOAD = 2, MB_GLOBAL_ATOMIC = 3 };

template <int KIND>
__global__ void __launch_bounds__(1024, 1)
microbench_kernel(int n, int iters, const uint2* __restrict__ table, unsigned long long* __restrict__ hist,
                  unsigned long long* __restrict__ sink) {
    extern __shared__ __align__(16) unsigned char smem[];
    uint32_t* hist_s = reinterpret_cast<uint32_t*>(smem);
    uint2* tab_s = reinterpret_cast<uint2*>(smem);
    const int tid = threadIdx.x;
    uint32_t state = (blockIdx.x * 1024u + tid) * 2654435761u + 12345u;
    if constexpr (KIND == MB_SMEM_ATOMIC) {
        for (int i = tid; i < n; i += 1024) hist_s[i] = 0u;
    } else if constexpr (KIND == MB_SMEM_LOAD) {
        for (int i = tid; i < n; i += 1024) tab_s[i] = make_uint2((uint32_t)i * 2246822519u, (uint32_t)i);
    }
    __syncthreads();
    uint32_t acc = 0;
    for (int it = 0; it < iters; ++it) {
        state = state * 1664525u + 1013904223u;
        const uint32_t idx = __umulhi(state, (uint32_t)n);
        if constexpr (KIND == MB_SMEM_ATOMIC) {
            atomicAdd(hist_s + idx, 1u);
        } else if constexpr (KIND == MB_SMEM_LOAD) {
            const uint2 v = tab_s[idx];
            state ^= v.x;                      // dependent: the next index needs this load
            acc += v.y;
        } else if constexpr (KIND == MB_L2_LOAD) {
            const uint2 v = __ldg(table + idx);
            state ^= v.x;
            acc += v.y;
        } else {
            atomicAdd(hist + idx, 1ull);
        }
    }
    if constexpr (KIND == MB_SMEM_ATOMIC) {
        __syncthreads();
        for (int i = tid; i < n; i += 1024) acc += hist_s[i];
    }
    if (acc == 0xdeadbeefu) atomicAdd(sink, (unsigned long long)acc);   // keep the work observable
}

}  // namespace nrw
