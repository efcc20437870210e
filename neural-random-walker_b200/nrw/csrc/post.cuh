// post.cuh -- K3 (solution assembly), K4 (row prefix sums), K5 (tempdata expansion).
#pragma once
#include "common.cuh"

namespace nrw {

// K4: np.cumsum(myQ, axis=1) restricted to the non-zeros (fluence_mini_app.py:224).  Sequential
// fp64 adds per row, so every partial sum carries the reference's bits (adding the dense row's
// exact zeros never changes a partial sum).  One thread per row; rows are short (<= ND-1 entries).
__global__ void build_cum_kernel(int n_rows, const int32_t* __restrict__ indptr,
                                 const double* __restrict__ vals, double* __restrict__ cum) {
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n_rows) return;
    const int lo = indptr[r], hi = indptr[r + 1];
    double acc = 0.0;
    for (int i = lo; i < hi; ++i) {
        acc = __dadd_rn(acc, vals[i]);
        cum[i] = acc;
    }
}

// K3: fluence_mini_app.py:422-438.  The reference adds, in (t, k) order with zero counts skipped,
//   tempval = tempval + tempdata[k,j]*np.exp(-SigA*dt*j)*v*sourceTerm(...)*dt
// i.e. (((count*tw[t])*v)*src[k])*dt, each product rounded (no FMA contraction: explicit _rn
// intrinsics), then divides by M.  tw[] and src[] are computed by the Python host with the same
// numpy / Python expressions as the reference, so this reproduces rwSol bit for bit from the
// replay events.  One thread per start node: the sum is sequential by definition.
__global__ void assemble_events_kernel(int n_starts, const long long* __restrict__ ev_offsets,
                                       const int32_t* __restrict__ ev_t,
                                       const int32_t* __restrict__ ev_node,
                                       const long long* __restrict__ ev_count,
                                       const double* __restrict__ tw, const double* __restrict__ src,
                                       double v, double dt, const long long* __restrict__ start_counts,
                                       int nt, int n_live, double* __restrict__ out) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n_starts) return;
    double tempval = 0.0;
    const long long lo = ev_offsets[s], hi = ev_offsets[s + 1];
    for (long long i = lo; i < hi; ++i) {
        const int t = ev_t[i], k = ev_node[i];
        if (t >= 1 && t <= nt - 2 && k < n_live) {
            double term = __dmul_rn((double)ev_count[i], tw[t]);
            term = __dmul_rn(term, v);
            term = __dmul_rn(term, src[k]);
            term = __dmul_rn(term, dt);
            tempval = __dadd_rn(tempval, term);
        }
    }
    out[s] = __ddiv_rn(tempval, (double)start_counts[s]);
}

// K5a: events -> dense float64 tempdata[n_rows, nt] (out pre-zeroed by cudaMemsetAsync).
__global__ void expand_events_kernel(long long n_events, const int32_t* __restrict__ ev_t,
                                     const int32_t* __restrict__ ev_node,
                                     const long long* __restrict__ ev_count, int n_rows, int nt,
                                     double* __restrict__ out) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_events) return;
    const int t = ev_t[i], k = ev_node[i];
    if (t < nt && k < n_rows) out[(size_t)k * nt + t] = (double)ev_count[i];
}

// K5b: dense uint32 count_tn[t_cap, n_rows] -> float64 tempdata[n_rows, nt]: a tiled transpose
// (coalesced on both sides through a padded shared-memory tile), t >= t_cap padded with zeros,
// column 0 = the initial condition tempdata[place, 0] = amount (fluence_mini_app.py:231).
__global__ void expand_counts_kernel(const uint32_t* __restrict__ count_tn, int t_cap, int n_rows,
                                     int nt, int start_node, long long start_count,
                                     double* __restrict__ out) {
    __shared__ uint32_t tile[32][33];
    const int t0 = blockIdx.x * 32, k0 = blockIdx.y * 32;
    for (int i = threadIdx.y; i < 32; i += blockDim.y) {
        const int t = t0 + i, k = k0 + threadIdx.x;
        uint32_t v = 0;
        if (t >= 1 && t < t_cap && t < nt && k < n_rows) v = count_tn[(size_t)t * n_rows + k];
        tile[i][threadIdx.x] = v;
    }
    __syncthreads();
    for (int i = threadIdx.y; i < 32; i += blockDim.y) {
        const int k = k0 + i, t = t0 + threadIdx.x;
        if (k < n_rows && t < nt) {
            double v = (double)tile[threadIdx.x][i];
            if (t == 0) v = (k == start_node) ? (double)start_count : 0.0;
            out[(size_t)k * nt + t] = v;
        }
    }
}

}  // namespace nrw
