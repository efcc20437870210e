// walk_philox.cuh -- K1: the on-device random-walk engine (statistical mode).
//
// Replaces the triple loop of run_pc (fluence_mini_app.py:236-245, simple_flux_miniapp.py:189-201)
// for ALL start nodes of run() (fluence_mini_app.py:397-407) in one launch:
//   * one thread per walker; a lane whose walker was absorbed / left through the sink is refilled
//     at once from the warp's walker-id cursor (warp vote: __ballot_sync + popc ranks the dead
//     lanes; no atomics), so lanes stay busy although the mean walk is only ~2 steps long;
//   * warps pull contiguous walker-id ranges from one global counter (dynamic balance); the range sizes
//     are guided -- big first, halving over the tail of the id space (fetch_range below) -- so that the
//     warps finish together;
//   * Philox2x32-10 keyed by the seed with counter (walker id, step) -- (walker id, step pair) on
//     compact tables, see common.cuh: any walker's trajectory is
//     a pure function of (seed, id), independent of grid shape and of how ranks shard the ids;
//   * transition tables: de-duplicated alias patterns + per-row descriptors staged ONCE per CTA
//     into shared memory with 1-D TMA bulk copies (UBLKCP) when they fit (TAB_SMEM), descriptors
//     left L2-resident when only the patterns fit (TAB_HYB), everything L2-resident for arbitrary
//     graphs (TAB_GLB); TAB_CDF runs the reference's own searchsorted on the fp64 prefix sums;
//   * tallies: per-start score in registers (fused fluence_mini_app.py:428-434), occupancy in a
//     per-CTA shared-memory histogram merged once (sink exits aggregated by warp vote), and
//     time-resolved counts (tempdata) in shared memory with __match_any_sync aggregation or in
//     global memory.
#pragma once
#include <type_traits>

#include "common.cuh"
#include "walk_tail.cuh"

namespace nrw {

enum : int { TAB_SMEM = 0, TAB_HYB = 1, TAB_GLB = 2, TAB_CDF = 3, TAB_FAC = 4, TAB_TREE = 5 };
// compact alias tables (patterns in shared memory, paired draws); they differ in where a row's descriptor
// comes from: shared memory (SMEM), global memory / L2 (HYB), or derived from three small shared-memory
// tables on meshes with the (direction, cell) structure of the fluence matrix (FAC, fac_row below)
__host__ __device__ constexpr bool tab_compact(int tab) { return tab == TAB_SMEM || tab == TAB_HYB || tab == TAB_FAC; }
enum : int { TL_SCORE = 1, TL_TW = 2, TL_HIST_S = 4, TL_HIST_G = 8, TL_CNT_S = 16, TL_CNT_G = 32 };

constexpr int WALK_THREADS = 1024;   // NRW_MAX_PHASES (guided range sizes): include/nrw.h

struct WalkParams {
    int n_rows, n_live, n_alias;
    const int32_t* indptr; const int32_t* cols; const double* cum; const uint8_t* node_class;
    const void* rowinfo; const void* alias;
    const uint8_t* tree; const int32_t* tree_ptr;   // TAB_TREE: 8-bit Bernoulli thresholds in heap order per row
    int n_starts; const int32_t* start_nodes; const long long* start_prefix; const long long* start_id_base;
    long long lo, hi;
    uint32_t key; int nt;
    uint32_t rk[10];                             // Philox round keys key + r * W32
    int t_last;                                  // nt - 1
    uint32_t t_fast;                             // max(nt - 4, 0): a step pair that starts at t < t_fast ends before step nt - 2
    long long* score; double* score_sq;
    uint32_t cwpack; double cv[4]; const double* tw;
    unsigned long long* node_hist;
    uint32_t* count_tn; int t_cap; int per_start; int cnt_bins; int cnt_match;
    unsigned long long* counters;
    unsigned long long* work_next; long long range;   // work_next: the launch's ticket counter (a slot of the engine's ring)
    TailParams tail;                             // last-CTA work: slot reset, zeroing, the fused all-reduce (walk_tail.cuh)
    // guided range sizes (walk_score.cuh, fetch_range below): phase j hands out ids
    // [lo + ph_start[j], lo + ph_start[j+1]) in ranges of ph_range[j], tickets ph_ticket[j] .. ph_ticket[j+1]-1
    int n_phase;
    long long ph_start[NRW_MAX_PHASES + 1], ph_ticket[NRW_MAX_PHASES + 1], ph_range[NRW_MAX_PHASES];
    // shared-memory carve-up (byte offsets into dynamic smem)
    int sm_rowinfo, sm_alias, sm_hist, sm_cnt;
    // CTA-level cache of hot (t, node) bins in front of the global time-resolved counts: 2^ccache_bits
    // (tag, count) slots at sm_ccache; 0 bits = off
    int sm_ccache, ccache_bits;
    // most likely trajectory of every start (nrw_walk.count_path): walkers still on it are tallied through a per-CTA
    // histogram of the steps at which walkers LEAVE it (sm_left: n_starts x (t_cap + 1)); path staged at sm_path
    const int32_t* cnt_path; int sm_path, sm_left;
    uint32_t str_thr;                            // structured draw (walk_score.cuh, STR): the one alias threshold
    uint32_t str_n;                              // ... and the one length of the rows a walker draws from (DET)
    uint32_t zero;                               // 0 (walk_score.cuh adds it to its loop constants: they then stay in uniform registers)
    // TAB_FAC (tables.py::FactoredRows): live row k = i + j * fac_period; sources of the staged tables
    const void* fac_dir; const void* fac_tail; const void* fac_cls;
    int fac_period, fac_cells, fac_n_tail, fac_shift;
    uint32_t fac_magic, fac_exit_w0, fac_exit_base;
    int sm_fdir, sm_ftail, sm_fcls;
};

// Row descriptor (w0, base) of row k derived instead of loaded (TAB_FAC): the direction table gives
// the pattern and the cell shift, the 2-bit class table the source class; a row whose target cell lies
// outside the mesh is an exit row.  Bit-identical to the rowinfo table it replaces (checked on the
// host for every row before the mode is offered: tables.py::factorise_rows).
__device__ __forceinline__ uint2 fac_row(const WalkParams& p, int k) {
    extern __shared__ __align__(16) unsigned char smem_fac[];
    if (k >= p.n_live) return reinterpret_cast<const uint2*>(smem_fac + p.sm_ftail)[k - p.n_live];
    const unsigned long long prod = (unsigned long long)(uint32_t)k * p.fac_magic;
    const uint32_t j = __funnelshift_r((uint32_t)prod, (uint32_t)(prod >> 32), p.fac_shift);   // k / period
    const uint32_t i = (uint32_t)k - j * (uint32_t)p.fac_period;
    const uint2 d = reinterpret_cast<const uint2*>(smem_fac + p.sm_fdir)[i];
    const uint32_t jn = j + d.y;                                   // target cell (wraps when negative)
    const uint32_t cls = (reinterpret_cast<const uint32_t*>(smem_fac + p.sm_fcls)[k >> 4] >> ((k & 15) * 2)) & 3u;
    const bool leaves = jn >= (uint32_t)p.fac_cells;
    // the class enters the cls field [27:29) and the inc field [29:32) (inc = cls + class of an absorbing successor,
    // which the exit descriptor already carries): one multiply-add, no carry out of either field
    return make_uint2((leaves ? p.fac_exit_w0 : d.x) + cls * 0x28000000u,
                      leaves ? p.fac_exit_base : (jn * (uint32_t)p.fac_period) | 0x40000000u);
}

// Next id range [nb, nb + len) of a warp, or false when the launch's ids are exhausted (walk_score.cuh and the
// general kernel below).  Ranges are numbered by TICKETS drawn from one global counter and their sizes are GUIDED: phase 0 hands
// out big ranges (few fetches), and the tail of the id space goes out in phases of ranges half as long
// each, two ranges per warp and phase, down to `range_tail` ids -- so that the warps finish together.  With one range size the SM sub-partitions were
// active for 88 % of the kernel on average: a warp that draws a heavy range last finishes up to one range
// late (now 96 %).  Small ranges everywhere are no alternative: every fetch idles part of the warp for an
// iteration and the returning atomics on the one counter serialise in L2 (a launch with 128-id ranges for
// a quarter of its ids: 0.55 -> 0.96 ms); the guided sizes need ~13 tickets per warp.  (Drawing the next ticket ahead of time
// was measured and dropped: a warp that holds a range in advance keeps it from warps that run dry first.)
__device__ __forceinline__ bool fetch_range(const WalkParams& p, int lane, long long& nb, long long& len) {
    long long t = 0;
    if (lane == 0) t = (long long)atomicAdd(p.work_next, 1ull);
    t = __shfl_sync(NRW_FULL_MASK, t, 0);
    int j = 0;
#pragma unroll 1
    while (j < p.n_phase && t >= p.ph_ticket[j + 1]) ++j;
    if (j >= p.n_phase) return false;
    const long long a = p.ph_start[j] + (t - p.ph_ticket[j]) * p.ph_range[j];
    nb = p.lo + a;
    len = min(p.ph_range[j], p.ph_start[j + 1] - a);
    return true;
}

// Start index s with start_prefix[s] <= g < start_prefix[s + 1].  Interpolation guess first -- exact when
// every start holds the same number of walkers (the fluence solve: M per node), one round trip to L2
// instead of the log2(n_starts) dependent ones of the bisection that remains the fall-back.
__device__ __forceinline__ int locate_start(const WalkParams& p, long long g) {
    const long long* __restrict__ pre = p.start_prefix;
    const int n = p.n_starts;
    const long long total = __ldg(pre + n);
    int s0 = (int)((double)g * (double)n / (double)total);
    s0 = max(1, min(s0, n - 2));                             // window s0-1 .. s0+1 inside [0, n) when n >= 3
    if (n >= 3) {
        const long long a0 = __ldg(pre + s0 - 1), a1 = __ldg(pre + s0), a2 = __ldg(pre + s0 + 1), a3 = __ldg(pre + s0 + 2);
        if (a1 <= g && g < a2) return s0;
        if (a0 <= g && g < a1) return s0 - 1;
        if (a2 <= g && g < a3) return s0 + 1;
    }
    int lo_i = 0, hi_i = n;
    while (hi_i - lo_i > 1) {
        const int mid = (lo_i + hi_i) >> 1;
        if (__ldg(pre + mid) <= g) lo_i = mid; else hi_i = mid;
    }
    return lo_i;
}

template <int TAB>
struct RowRegs { uint32_t a, b, c, d; };   // COMPACT: a=w0 b=w1 | WIDE: off,n,base,cls | CDF: a=k b=n d=cls | TREE: a=k b=n c=tree offset d=cls

template <int TAB>
__device__ __forceinline__ RowRegs<TAB> load_row(const WalkParams& p, const uint2* rowinfo_s, int k) {
    RowRegs<TAB> r;
    if constexpr (TAB == TAB_SMEM) {               // compact: b = the column base without its flag bits
        const uint2 v = rowinfo_s[k];
        r.a = v.x; r.b = row_base(v.y); r.c = 0; r.d = 0;
    } else if constexpr (TAB == TAB_HYB) {
        const uint2 v = __ldg(reinterpret_cast<const uint2*>(p.rowinfo) + k);
        r.a = v.x; r.b = row_base(v.y); r.c = 0; r.d = 0;
    } else if constexpr (TAB == TAB_FAC) {
        const uint2 v = fac_row(p, k);
        r.a = v.x; r.b = row_base(v.y); r.c = 0; r.d = 0;
    } else if constexpr (TAB == TAB_GLB) {
        const uint4 v = __ldg(reinterpret_cast<const uint4*>(p.rowinfo) + k);
        r.a = v.x; r.b = v.y; r.c = v.z; r.d = v.w;
    } else {
        r.a = (uint32_t)k;
        r.b = (k < p.n_live) ? (uint32_t)(__ldg(p.indptr + k + 1) - __ldg(p.indptr + k)) : 0u;
        r.c = 0;
        if constexpr (TAB == TAB_TREE) r.c = (uint32_t)__ldg(p.tree_ptr + k);
        r.d = p.node_class ? (uint32_t)__ldg(p.node_class + k) : 0u;
    }
    return r;
}
template <int TAB>
__device__ __forceinline__ uint32_t row_n(const RowRegs<TAB>& r) {
    if constexpr (tab_compact(TAB)) return row_len(r.a);
    else return r.b;
}
template <int TAB>
__device__ __forceinline__ uint32_t row_cls(const RowRegs<TAB>& r) {
    if constexpr (tab_compact(TAB)) return row_class(r.a);
    else return r.d & 3u;
}

// Stage the de-duplicated alias patterns (and, for TAB_SMEM, the row descriptors) into shared
// memory: thread 0 arms one mbarrier with the byte count and issues 1-D TMA bulk copies (SASS
// UBLKCP) in <= 32 KB pieces; an odd entry count leaves one 8-byte tail outside the 16-byte
// granules, copied by thread 32.  Callers wait with mbar_wait(bar, 0) + __syncthreads().
template <int TAB>
__device__ __forceinline__ void stage_tables_issue(const WalkParams& p, uint64_t* bar, uint2* rowinfo_s,
                                                   uint2* alias_s, int tid) {
    if constexpr (tab_compact(TAB)) {
        if (tid == 0) mbar_init(bar, 1);
        __syncthreads();
        const uint32_t alias_bytes = ((uint32_t)p.n_alias * 8u) & ~15u;
        const uint32_t row_bytes = (TAB == TAB_SMEM) ? (((uint32_t)p.n_rows * 8u) & ~15u) : 0u;
        if (tid == 0) {
            mbar_expect_tx(bar, alias_bytes + row_bytes);
            for (uint32_t o = 0; o < alias_bytes; o += 32768u)
                tma_bulk_g2s(reinterpret_cast<unsigned char*>(alias_s) + o,
                             reinterpret_cast<const unsigned char*>(p.alias) + o,
                             min(32768u, alias_bytes - o), bar);
            if constexpr (TAB == TAB_SMEM)
                for (uint32_t o = 0; o < row_bytes; o += 32768u)
                    tma_bulk_g2s(reinterpret_cast<unsigned char*>(rowinfo_s) + o,
                                 reinterpret_cast<const unsigned char*>(p.rowinfo) + o,
                                 min(32768u, row_bytes - o), bar);
        }
        if (tid == 32) {
            if (p.n_alias & 1)
                alias_s[p.n_alias - 1] = __ldg(reinterpret_cast<const uint2*>(p.alias) + p.n_alias - 1);
            if (TAB == TAB_SMEM && (p.n_rows & 1))
                rowinfo_s[p.n_rows - 1] = __ldg(reinterpret_cast<const uint2*>(p.rowinfo) + p.n_rows - 1);
        }
        if constexpr (TAB == TAB_FAC) {          // three small tables (C4: 1 KB + 8 B + 40 KB): plain loads
            extern __shared__ __align__(16) unsigned char smem_stage[];
            const int nthr = (int)blockDim.x;
            uint32_t* d = reinterpret_cast<uint32_t*>(smem_stage + p.sm_fdir);
            for (int i = tid; i < 2 * p.fac_period; i += nthr) d[i] = __ldg(reinterpret_cast<const uint32_t*>(p.fac_dir) + i);
            d = reinterpret_cast<uint32_t*>(smem_stage + p.sm_ftail);
            for (int i = tid; i < 2 * p.fac_n_tail; i += nthr) d[i] = __ldg(reinterpret_cast<const uint32_t*>(p.fac_tail) + i);
            d = reinterpret_cast<uint32_t*>(smem_stage + p.sm_fcls);
            for (int i = tid; i < (p.n_live + 15) / 16; i += nthr) d[i] = __ldg(reinterpret_cast<const uint32_t*>(p.fac_cls) + i);
        }
    }
}

template <int TAB, int TL>
__global__ void __launch_bounds__(WALK_THREADS, 1)
walk_philox_kernel(const __grid_constant__ WalkParams p) {
    extern __shared__ __align__(16) unsigned char smem[];
    uint64_t* bar = reinterpret_cast<uint64_t*>(smem);
    uint2* rowinfo_s = reinterpret_cast<uint2*>(smem + p.sm_rowinfo);
    uint2* alias_s = reinterpret_cast<uint2*>(smem + p.sm_alias);
    uint32_t* hist_s = reinterpret_cast<uint32_t*>(smem + p.sm_hist);
    uint32_t* cnt_s = reinterpret_cast<uint32_t*>(smem + p.sm_cnt);
    const int tid = threadIdx.x, lane = tid & 31;

    // ---- stage the tables: TMA bulk copies complete on one mbarrier ---------------------------
    constexpr bool kStage = tab_compact(TAB);
    stage_tables_issue<TAB>(p, bar, rowinfo_s, alias_s, tid);
    if constexpr (TL & TL_HIST_S)
        for (int i = tid; i < p.n_rows; i += WALK_THREADS) hist_s[i] = 0u;
    if constexpr (TL & TL_CNT_S)
        for (int i = tid; i < p.cnt_bins; i += WALK_THREADS) cnt_s[i] = 0u;
    uint2* ccache_s = reinterpret_cast<uint2*>(smem + p.sm_ccache);     // .x = bin + 1 (0 = free), .y = count
    if constexpr (TL & TL_CNT_G)
        if (p.ccache_bits)
            for (int i = tid; i < (1 << p.ccache_bits); i += WALK_THREADS) ccache_s[i] = make_uint2(0u, 0u);
    int32_t* path_s = reinterpret_cast<int32_t*>(smem + p.sm_path);
    uint32_t* left_s = reinterpret_cast<uint32_t*>(smem + p.sm_left);
    const bool use_path = (TL & TL_CNT_G) && p.cnt_path != nullptr;
    if constexpr (TL & TL_CNT_G)
        if (use_path) {
            for (int i = tid; i < p.n_starts * p.t_cap; i += WALK_THREADS) path_s[i] = __ldg(p.cnt_path + i);
            for (int i = tid; i < p.n_starts * (p.t_cap + 1); i += WALK_THREADS) left_s[i] = 0u;
        }
    if constexpr (kStage) mbar_wait(bar, 0);
    __syncthreads();

    const uint32_t key = p.key;
    const int t_last = p.nt - 1;        // last model step (fluence_mini_app.py:236)
    const int t_score_max = p.nt - 2;   // solution sum skips the final step (:428, Q3)
    using ScoreT = typename std::conditional<(TL & TL_TW) != 0, long long, uint32_t>::type;

    // ---- lane state (32-bit wherever the range allows: the loop is issue-bound) ------------------
    int t = -1;                         // steps taken by the lane's walker; -1 = no walker
    uint32_t c0 = 0, chi = 0;           // Philox counter: stream id (low 32 | bits 32..39 << 24)
    RowRegs<TAB> row{0, 0, 0, 0};
    ScoreT x = 0;                       // this walker's score
    int s = -1;                         // start index the lane is accumulating for
    long long s_end = 0;                // global id at which start s ends
    uint32_t s_end_off = 0;             // the same, as an offset into the warp's current range
    long long sid_base = 0;             // stream id of offset 0 of the current range within start s
    int t_init = 0; uint32_t surv_init = 0;   // state of a fresh walker of start s (dead if absorbing)
    RowRegs<TAB> srow{0, 0, 0, 0};
    long long acc = 0; double acc2 = 0.0;
    uint32_t nsteps = 0, nsurv = 0, novf = 0, sinkc = 0;
    bool onpath = false;                // TL_CNT_G with count_path: the walker has followed its start's likely path so far
    // warp-uniform: the warp's current id range is [rbase + cur, rbase + end)
    long long rbase = 0;
    uint32_t cur = 0, end = 0;
    unsigned long long wwalk = 0;       // walkers taken by this warp (sum of range lengths)
    bool exhausted = false;
    const uint32_t lt_mask = lanemask_lt();

    auto flush_score = [&]() {
        if constexpr (TL & TL_SCORE) {
            if (s >= 0) {
                if (acc != 0) atomicAdd(reinterpret_cast<unsigned long long*>(p.score) + s,
                                        (unsigned long long)acc);
                if (p.score_sq && acc2 != 0.0) atomicAdd(p.score_sq + s, acc2);
            }
            acc = 0; acc2 = 0.0;
        }
    };
    // offset (clamped to 2^31) of the end of start s inside the current range
    auto rel_end = [&]() -> uint32_t {
        const long long d = s_end - rbase;
        return d <= 0 ? 0u : (d > 0x7fffffffll ? 0x7fffffffu : (uint32_t)d);
    };

    while (true) {
        // ---- refill dead lanes (warp vote ranks them; ids only ever grow for a lane) ---------
        const unsigned need = __ballot_sync(NRW_FULL_MASK, t < 0);
        if (need) {
            if (cur >= end) {
                if (!exhausted) {
                    long long nb = 0, len = 0;
                    cur = 0;
                    if (!fetch_range(p, lane, nb, len)) { exhausted = true; end = 0; }
                    else {
                        end = (uint32_t)len;
                        wwalk += end;
                        sid_base += nb - rbase;                  // same start, new range origin
                        rbase = nb;
                        s_end_off = rel_end();
                    }
                }
                if (exhausted && !__any_sync(NRW_FULL_MASK, t >= 0)) break;
            }
            const uint32_t off = cur + (uint32_t)__popc(need & lt_mask);
            const bool take = (t < 0) && (off < end);
            cur = min(cur + (uint32_t)__popc(need), end);
            if (take) {
                if (off >= s_end_off) {                           // next start (cold path)
                    const long long gnew = rbase + (long long)off;
                    flush_score();
                    long long s_lo = s_end;
                    bool found = false;
                    if (s >= 0 && s + 1 < p.n_starts) {
                        const long long e2 = __ldg(p.start_prefix + s + 2);
                        if (gnew < e2) { ++s; s_end = e2; found = true; }
                    }
                    if (!found) {
                        s = locate_start(p, gnew);
                        s_lo = __ldg(p.start_prefix + s);
                        s_end = __ldg(p.start_prefix + s + 1);
                    }
                    srow = load_row<TAB>(p, rowinfo_s, __ldg(p.start_nodes + s));
                    sid_base = rbase + (p.start_id_base ? (__ldg(p.start_id_base + s) - s_lo) : 0ll);
                    s_end_off = rel_end();
                    const bool absorbing = row_n<TAB>(srow) == 0u;   // started on the sink: an exit
                    t_init = (absorbing || t_last < 1) ? -1 : 0;
                    surv_init = (!absorbing && t_last < 1) ? 1u : 0u;
                }
                row = srow;
                const unsigned long long sid = (unsigned long long)(sid_base + (long long)off);
                c0 = (uint32_t)sid;
                chi = (uint32_t)(sid >> 32) << 24;
                t = t_init; x = 0;
                onpath = true;
                nsurv += surv_init;
            }
        }

        // ---- model steps for every live lane ------------------------------------------------------
        // compact tables: one Philox call serves a PAIR of steps (common.cuh, "paired draws"); lanes are
        // only refilled above, so every live lane is at an even t here and the warp shares the call
        constexpr bool PAIRED = tab_compact(TAB);
        uint2 rp = make_uint2(0u, 0u);
        if constexpr (PAIRED) {
            if (t >= 0) rp = philox2x32_10(c0, chi | (uint32_t)(t + 1), key);
        }
        auto model_step = [&](const bool first) {
            const bool stepped = t >= 0;
            int knew = 0;
            if (stepped) {
                ++t; ++nsteps;
                uint2 r = rp;
                if constexpr (!PAIRED) r = philox2x32_10(c0, chi | (uint32_t)t, key);
                if constexpr (TAB == TAB_CDF) {
                    knew = sample_cdf(p.indptr, p.cols, p.cum, (int)row.a, u53(r.x, r.y));
                } else if constexpr (TAB == TAB_TREE) {
                    // the probability circuit of loihirw/rw_core.py:11-156: one Bernoulli neuron per tree level,
                    // firing probability quantised to the 8-bit weight 128 p + 36 (here: threshold floor(128 p))
                    unsigned long long bits = (unsigned long long)r.x | ((unsigned long long)r.y << 32);
                    const int depth = row.b > 1u ? 32 - __clz((int)row.b - 1) : 0;
                    uint32_t node = 0;
                    for (int l = 0; l < depth; ++l) {
                        const uint32_t thr = __ldg(p.tree + row.c + node);
                        node = 2u * node + (((uint32_t)bits & 127u) < thr ? 1u : 2u);
                        bits >>= 7;
                    }
                    const uint32_t leaf = min(node - ((1u << depth) - 1u), row.b - 1u);
                    knew = __ldg(p.cols + __ldg(p.indptr + (int)row.a) + (int)leaf);
                } else if constexpr (TAB == TAB_GLB) {
                    const uint32_t slot = __umulhi(r.x, row.b);
                    const uint4 e = __ldg(reinterpret_cast<const uint4*>(p.alias) + row.a + slot);
                    knew = (int)(row.c + ((r.y <= e.x) ? e.y : e.z));
                } else {
                    const uint32_t n = row_len(row.a), off = row_off(row.a);
                    const unsigned long long v = (unsigned long long)(first ? r.x : r.y) * n;
                    const uint2 e = alias_s[off + (uint32_t)(v >> 32)];
                    knew = (int)(row.b + (((uint32_t)v <= e.x) ? (e.y & 0xffffu) : (e.y >> 16)));
                }
                row = load_row<TAB>(p, rowinfo_s, knew);
            }
            const bool in_sum = stepped && (t <= t_score_max);

            // ---- tallies --------------------------------------------------------------------------------
            if constexpr (TL & TL_SCORE) {
                if (in_sum) {
                    const uint32_t cls = row_cls<TAB>(row);
                    if constexpr (TL & TL_TW) {
                        const double cvv = cls == 0 ? p.cv[0] : cls == 1 ? p.cv[1] : cls == 2 ? p.cv[2] : p.cv[3];
                        x += __double2ll_rn(__ldg(p.tw + t) * cvv);
                    } else {
                        x += (p.cwpack >> (cls * 8u)) & 0xffu;        // byte-packed integer class weights
                    }
                }
            }
            if constexpr (TL & (TL_HIST_S | TL_HIST_G)) {
                // sink exits are the one hot bin: aggregate them with a warp vote
                const bool hot = in_sum && (knew == p.n_rows - 1);
                const unsigned hm = __ballot_sync(NRW_FULL_MASK, hot);
                if (lane == 0) sinkc += __popc(hm);
                if (in_sum && !hot) {
                    if constexpr (TL & TL_HIST_S) atomicAdd(hist_s + knew, 1u);
                    else atomicAdd(p.node_hist + knew, 1ull);
                }
            }
            if constexpr (TL & TL_CNT_S) {
                const int gq = p.per_start ? s : 0;
                int bin = -1;
                if (stepped) {
                    if (t < p.t_cap) bin = (gq * p.t_cap + t) * p.n_rows + knew; else ++novf;
                }
                if (p.cnt_match) {       // many lanes on the same (t, node) bin (short walks): one atomic per bin
                    const unsigned m = __match_any_sync(NRW_FULL_MASK, bin);
                    if (bin >= 0 && lane == __ffs(m) - 1) atomicAdd(cnt_s + bin, (uint32_t)__popc(m));
                } else if (bin >= 0) {   // lanes spread over many bins (long walks): plain shared-memory atomics
                    atomicAdd(cnt_s + bin, 1u);
                }
            }
            if constexpr (TL & TL_CNT_G) {
                const long long gq = p.per_start ? s : 0;
                long long bin = -1;
                if (stepped) {
                    if (t < p.t_cap) bin = (gq * p.t_cap + t) * (long long)p.n_rows + knew; else ++novf;
                }
                if (use_path) {
                    // Walkers of one start move almost in lock step along the start's most likely trajectory, so
                    // its (t, node) bins would take most tallies (contended atomics).  A walker still ON the path
                    // is not tallied step by step: when it leaves the path at step L (first step off it, the step
                    // after its death, or t_cap) ONE shared-memory atomic records L, and the epilogue credits the
                    // path bins 1 .. L-1 from the suffix sums of that histogram -- the same counts exactly.
                    if (stepped && onpath) {
                        const int tc = t < p.t_cap ? t : p.t_cap;
                        if (t >= p.t_cap || knew != path_s[s * p.t_cap + t]) {
                            atomicAdd(left_s + s * (p.t_cap + 1) + tc, 1u);
                            onpath = false;
                        } else {
                            bin = -1;                                  // on the path: credited in the epilogue
                        }
                    }
                    if (bin >= 0) atomicAdd(p.count_tn + bin, 1u);      // off the path: diffuse bins, plain RED
                } else if (p.ccache_bits) {
                    // Walkers of one start node move almost in lock step, so a few (t, node) bins take
                    // most of the tallies and global atomics on them serialise in L2.  Lanes of a warp on
                    // the same bin are merged by a vote; the merged count goes to a per-CTA shared-memory
                    // cache of (bin, count) slots, direct-mapped by a hash; a slot belongs to the first bin
                    // that claims it, bins that find theirs taken go to global memory directly.  The
                    // cache is added to the global counts once, in the epilogue.
                    const unsigned m = __match_any_sync(NRW_FULL_MASK, bin);
                    if (bin >= 0 && lane == __ffs(m) - 1) {
                        const uint32_t c = (uint32_t)__popc(m);
                        bool cached = false;
                        if (bin < 0xfffffffell) {
                            const uint32_t tag = (uint32_t)bin + 1u;
                            uint2* slot = ccache_s + ((tag * 2654435761u) >> (32 - p.ccache_bits));
                            uint32_t cur_tag = *reinterpret_cast<volatile uint32_t*>(&slot->x);
                            if (cur_tag == 0u && c >= 2u) {       // only bins that are hot NOW claim a slot:
                                cur_tag = atomicCAS(&slot->x, 0u, tag);   // keeps the long tail of rarely
                                if (cur_tag == 0u) cur_tag = tag;         // visited bins out of the cache
                            }
                            if (cur_tag == tag) { atomicAdd(&slot->y, c); cached = true; }
                        }
                        if (!cached) atomicAdd(p.count_tn + bin, c);
                    }
                } else if (bin >= 0) {
                    atomicAdd(p.count_tn + bin, 1u);
                }
            }

            // ---- absorption / sink exit / end of time ------------------------------------------------
            if (stepped) {
                const bool absorbed = row_n<TAB>(row) == 0u;
                if (absorbed || t >= t_last) {
                    if (!absorbed) ++nsurv;
                    if constexpr (TL & TL_CNT_G)
                        if (use_path && onpath) {                      // died on the path at step t: credited through t
                            atomicAdd(left_s + s * (p.t_cap + 1) + min(t + 1, p.t_cap), 1u);
                            onpath = false;
                        }
                    t = -1;
                    if constexpr (TL & TL_SCORE) {
                        acc += (long long)x;
                        if (p.score_sq) { const double xd = (double)x; acc2 += xd * xd; }
                    }
                }
            }
            };
        model_step(true);
        if constexpr (PAIRED) model_step(false);
    }

    // ---- epilogue: flush register tallies, merge the CTA histogram once ----------------------
    flush_score();
    if constexpr (TL & (TL_HIST_S | TL_HIST_G)) {
        if (lane == 0 && sinkc) {
            if constexpr (TL & TL_HIST_S) atomicAdd(hist_s + p.n_rows - 1, sinkc);
            else atomicAdd(p.node_hist + p.n_rows - 1, (unsigned long long)sinkc);
        }
    }
    unsigned long long v0 = nsteps;
    uint32_t v2 = nsurv, v3 = novf;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        v0 += __shfl_xor_sync(NRW_FULL_MASK, v0, o);
        v2 += __shfl_xor_sync(NRW_FULL_MASK, v2, o);
        v3 += __shfl_xor_sync(NRW_FULL_MASK, v3, o);
    }
    const unsigned long long v4 = wwalk, v1 = wwalk - v2;          // exits = walkers - survivors
    if (lane == 0 && p.counters) {
        if (v0) atomicAdd(p.counters + 0, v0);
        if (v1) atomicAdd(p.counters + 1, v1);
        if (v2) atomicAdd(p.counters + 2, (unsigned long long)v2);
        if (v3) atomicAdd(p.counters + 3, (unsigned long long)v3);
        if (v4) atomicAdd(p.counters + 4, v4);
    }
    if constexpr (TL & TL_CNT_G) {
        if (use_path) {
            // walkers on the path at step tau = those that left it later: suffix sums of the leave-time histogram
            __syncthreads();
            for (int sidx = tid; sidx < p.n_starts; sidx += WALK_THREADS) {
                const long long gq = p.per_start ? sidx : 0;
                unsigned long long on = left_s[sidx * (p.t_cap + 1) + p.t_cap];
                for (int tau = p.t_cap - 1; tau >= 1; --tau) {
                    const int node = path_s[sidx * p.t_cap + tau];
                    if (on && node >= 0)
                        atomicAdd(p.count_tn + (gq * p.t_cap + tau) * (long long)p.n_rows + node, (uint32_t)on);
                    on += left_s[sidx * (p.t_cap + 1) + tau];
                }
            }
        }
        if (p.ccache_bits) {
            __syncthreads();
            for (int i = tid; i < (1 << p.ccache_bits); i += WALK_THREADS) {
                const uint2 v = ccache_s[i];
                if (v.x && v.y) atomicAdd(p.count_tn + (v.x - 1u), v.y);
            }
        }
    }
    if constexpr (TL & (TL_HIST_S | TL_CNT_S)) {
        __syncthreads();
        if constexpr (TL & TL_HIST_S)
            for (int i = tid; i < p.n_rows; i += WALK_THREADS) {
                const uint32_t h = hist_s[i];
                if (h) atomicAdd(p.node_hist + i, (unsigned long long)h);
            }
        if constexpr (TL & TL_CNT_S)
            for (int i = tid; i < p.cnt_bins; i += WALK_THREADS) {
                const uint32_t h = cnt_s[i];
                if (h) atomicAdd(p.count_tn + i, h);
            }
    }
    walk_tail(p.tail);
}

}  // namespace nrw
