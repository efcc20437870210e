// walk_score.cuh -- K1s: the score-only fast path of the on-device walk (the fused fluence solve).
//
// Same semantics as walk_philox_kernel<TAB, TL_SCORE> (walk_philox.cuh) -- identical Philox streams,
// alias draws, tallies and counters, checked bit for bit against the numpy port -- with a hot loop
// trimmed for the resource that bounds it (issue slots and the half-rate integer ALU pipe, see
// DESIGN.md section 3):
//   * refill with ONE vote per iteration: dead lanes rank themselves with popc and take consecutive
//     ids; everything rare (range exhausted, next start node, 2^32 carry of the stream id, walkers on
//     absorbing starts) sits behind warp-uniform or [[unlikely]] branches;
//   * id ranges of GUIDED size (fetch_range, walk_philox.cuh): big ranges first, the tail of the id space in
//     phases of halving ranges, so that the warps finish together; a range that begins in another start
//     node finds it by interpolation (locate_start) instead of bisection;
//   * paired draws (common.cuh): a refill is followed by TWO steps sharing one Philox call -- lanes
//     are only refilled at the top of the pair, so every live lane is at an even step;
//   * the loop's constants (Philox round keys, threshold, row length, end-of-time bound) are read as p.x + p.zero:
//     computed values stay in registers, plain constant-bank operands were re-loaded in every iteration;
//   * STR (nrw_tables.rowinfo_str / str_thresh / str_n): on fluence matrices the alias tables have a closed form --
//     target = base + (coin <= T ? slot : same) -- so a step is ONE shared-memory look-up; with DET the row length is
//     a launch constant and the walker carries its direction (and, on factored descriptors, its cell: FACJ);
//   * one unsigned test at the head of a step pair says "the lane holds a walker" and "neither step of the pair can
//     reach the last two steps of time"; the pair near the end of time is a cold copy with the tests; the flagship
//     pair has its second step written out, so that its draw, choice and address are issued while the first step's
//     look-up is in flight; with the tables in shared memory the take is a run of selects behind one cold branch;
//   * the score is a per-lane running sum over the lane's walkers (flushed when the start node
//     changes and at every range fetch); the common step is branch-free: what arriving on a row adds to score
//     and step count sits in the top bits of its descriptor (one shift-add each), no separate path for lanes
//     whose walker ends; the step t = nt-1 (outside the solution sum, fluence_mini_app.py:428) is dealt with
//     on the cold path of walkers within two steps of the end of time;
//   * IDW: class weights equal to the class index need no weight lookup;
//   * DET: a row whose single successor is absorbing (every exit row of the fluence mesh: 45 % of
//     all walker-steps) is stepped without computing a draw -- slot 0 with threshold 2^32-1 is taken
//     whatever the draw is, so tallies and counters stay bit-identical to the always-draw port.  A
//     walker landing on such a row takes its exit step in the same iteration; walkers STARTING on
//     one never enter the step phase: an idle warp inside such a start accounts for all the ids the
//     range and the start still hold at once (one step, one tally, one exit each), lanes that pick
//     them up one by one account for them in the take.  NRW_DET=0 disables it.
// Requires nt >= 2 (the planner sends nt = 1 to the general kernel).
#pragma once
#include <type_traits>
#include "walk_philox.cuh"

namespace nrw {

// Descriptor of row k, where k may still carry the flag bits of the base it was formed from (bits 29..31 of w1:
// common.cuh): in shared memory the 32-bit address arithmetic drops them ((k << 3) wraps), elsewhere they are masked.
// The constants of fac_row (walk_philox.cuh) as VALUES the kernel computes once (walk_score_kernel adds p.zero to
// each): read as plain kernel parameters they were re-loaded from the constant bank at every step -- six loads of
// the ~30 instructions a factored descriptor costs.
struct FacK {
    uint32_t n_live, magic, shift, period, cells, exit_w0, exit_base;
    uint32_t a_dir, a_tail, a_cls;               // 32-bit shared-window addresses of the three staged tables
};
__device__ __forceinline__ uint32_t lds32(uint32_t addr) {
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr));
    return v;
}
// fac_row with those values (same arithmetic, same result)
__device__ __forceinline__ uint2 fac_row_k(const FacK& f, uint32_t k) {
    if (k >= f.n_live) return lds64(f.a_tail + ((k - f.n_live) << 3));
    const unsigned long long prod = (unsigned long long)k * f.magic;
    const uint32_t j = __funnelshift_r((uint32_t)prod, (uint32_t)(prod >> 32), f.shift);   // k / period
    const uint32_t i = k - j * f.period;
    const uint2 d = lds64(f.a_dir + (i << 3));
    const uint32_t jn = j + d.y;                                   // target cell (wraps when negative)
    const uint32_t cls = (lds32(f.a_cls + ((k >> 4) << 2)) >> ((k & 15u) * 2u)) & 3u;
    const bool leaves = jn >= f.cells;
    return make_uint2((leaves ? f.exit_w0 : d.x) + cls * 0x28000000u,
                      leaves ? f.exit_base : (jn * f.period) | 0x40000000u);
}

template <int TAB>
__device__ __forceinline__ uint2 load_row2(const WalkParams& p, uint32_t rowinfo_a, uint32_t k, const FacK& f) {
    if constexpr (TAB == TAB_SMEM) return lds64(rowinfo_a + (k << 3));
    else if constexpr (TAB == TAB_FAC) return fac_row_k(f, k & ROW_BASE_MASK);
    else return __ldg(reinterpret_cast<const uint2*>(p.rowinfo) + (k & ROW_BASE_MASK));
}

// Philox2x32-10 with the ten round keys read from the kernel parameters (constant bank operands of
// the LOP3s) instead of being re-derived from the key in uniform registers every iteration.
__device__ __forceinline__ uint2 philox2x32_10_rk(uint32_t c0, uint32_t c1, const uint32_t (&rk)[10]) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint32_t hi = __umulhi(PHILOX_M2x32, c0);
        const uint32_t lo = PHILOX_M2x32 * c0;
        c0 = hi ^ rk[r] ^ c1;
        c1 = lo;
    }
    return make_uint2(c0, c1);
}

// The same with the first round key already folded into c1 (c1k = c1 ^ rk[0]).
__device__ __forceinline__ uint2 philox2x32_10_k0(uint32_t c0, uint32_t c1k, const uint32_t (&rk)[10]) {
    {
        const uint32_t hi = __umulhi(PHILOX_M2x32, c0);
        const uint32_t lo = PHILOX_M2x32 * c0;
        c0 = hi ^ c1k;
        c1k = lo;
    }
#pragma unroll
    for (int r = 1; r < 10; ++r) {
        const uint32_t hi = __umulhi(PHILOX_M2x32, c0);
        const uint32_t lo = PHILOX_M2x32 * c0;
        c0 = hi ^ rk[r] ^ c1k;
        c1k = lo;
    }
    return make_uint2(c0, c1k);
}

// CPS = CTAs per SM: 1 -> 1024 threads (tables up to 227 KB), 2 -> two CTAs of 768 threads when the
// tables fit half an SM's shared memory (48 instead of 32 resident warps).
// HIST: also tally every in-sum step at its target node (the occupancy histogram of walk_philox.cuh,
// TL_HIST_S): shared-memory atomics for the nodes walkers spread over, a per-lane register count for
// the one node they all end on (the last row: the sink), one merge into global memory per CTA.
// CNT: instead of a score, the time-resolved counts count_tn[t, node] (the reference's tempdata,
// fluence_mini_app.py:227,245) in global memory, tallied through the likely path of every start
// (nrw_walk.count_path, see walk_philox.cuh): a walker still on its start's path is not tallied step by step --
// one shared-memory atomic records the step at which it leaves the path and the epilogue credits the path bins
// from the suffix sums of that histogram; steps off the path go to their (diffuse) bins with plain REDs.
// STR: structured draw (include/nrw.h, nrw_tables.rowinfo_str): the row descriptor carries the row's own direction
// `same` where the pattern offset was, and  target = base + (coin <= p.str_thr ? slot : same)  -- the alias
// draw's target for every (slot, coin), verified entry by entry on the host -- needs no alias look-up: one
// shared-memory load per step instead of two dependent ones, and no alias patterns to stage.
// CNTS (with CNT): the bins fit shared memory (p.sm_cnt, p.cnt_bins: small graphs over long windows -- the 3-state
// chain of simple_flux_miniapp.py:174-203 over 1 002 steps): one plain shared-memory atomic per step into the CTA's
// own histogram, merged into count_tn once; no path logic.
template <int TAB, bool IDW, bool DET, int CPS, bool HIST = false, bool CNT = false, bool STR = false, bool CNTS = false>
__global__ void __launch_bounds__(CPS == 1 ? 1024 : 768, CPS)
walk_score_kernel(const __grid_constant__ WalkParams p) {
    static_assert(tab_compact(TAB), "compact alias tables only");
    static_assert(!CNT || (!DET && !HIST), "the count walk draws every step and tallies nothing else");
    static_assert(!CNTS || CNT, "CNTS is a mode of the count walk");
    // FACJ: factored descriptors under the structured draw with deterministic exits -- the walker carries its cell
    constexpr bool FACJ = TAB == TAB_FAC && STR && DET;
    extern __shared__ __align__(16) unsigned char smem[];
    uint64_t* bar = reinterpret_cast<uint64_t*>(smem);
    const uint32_t rowinfo_a = smem_u32(smem + p.sm_rowinfo);      // 32-bit shared-window address of the descriptors
    const uint2* alias_s = reinterpret_cast<const uint2*>(smem + p.sm_alias);
    uint32_t* hist_s = reinterpret_cast<uint32_t*>(smem + p.sm_hist);
    const int tid = threadIdx.x, lane = tid & 31;
    stage_tables_issue<TAB>(p, bar, reinterpret_cast<uint2*>(smem + p.sm_rowinfo),
                            reinterpret_cast<uint2*>(smem + p.sm_alias), tid);
    if constexpr (HIST)
        for (int i = tid; i < p.n_rows; i += (int)blockDim.x) hist_s[i] = 0u;
    int32_t* path_s = reinterpret_cast<int32_t*>(smem + p.sm_path);
    uint32_t* left_s = reinterpret_cast<uint32_t*>(smem + p.sm_left);
    uint32_t* cnt_s = reinterpret_cast<uint32_t*>(smem + p.sm_cnt);
    if constexpr (CNTS) {
        for (int i = tid; i < p.cnt_bins; i += (int)blockDim.x) cnt_s[i] = 0u;
    } else if constexpr (CNT) {
        for (int i = tid; i < p.n_starts * p.t_cap; i += (int)blockDim.x) path_s[i] = __ldg(p.cnt_path + i);
        for (int i = tid; i < p.n_starts * (p.t_cap + 1); i += (int)blockDim.x) left_s[i] = 0u;
    }
    mbar_wait(bar, 0);
    __syncthreads();
    const int sink_row = p.n_rows - 1;           // HIST: the node every exit ends on is counted in registers

    const int t_last = p.t_last;                 // nt - 1: last model step (fluence_mini_app.py:236); >= 1 here
    // What the hot loop reads every iteration, as COMPUTED uniform values (p.zero is 0): a plain constant-bank operand
    // is re-loaded into a uniform register in every iteration (five loads per step pair), a computed one stays there
    const uint32_t zz = p.zero;
    uint32_t rk[10];
#pragma unroll
    for (int r = 0; r < 10; ++r) rk[r] = p.rk[r] + zz;
    const uint32_t t_fast = p.t_fast + zz, str_thr = p.str_thr + zz, str_n = p.str_n + zz;
    // CNTS: steps below t_hot are inside the count window and before the last step of time (the common step)
    const int t_hot = min(p.t_cap, p.t_last) + (int)zz;
    const uint32_t n_rows_u = (uint32_t)p.n_rows + zz;
    uint32_t cnt_a = 0;                          // CNTS: shared-window address of the lane's start's bins [t_cap][n_rows]
    uint32_t path_a = 0, left_a = 0;             // CNT: shared-window addresses of the lane's start's path / leave histogram
    uint32_t* cnt_g = nullptr;                   // CNT: the lane's start's bins in global memory
    FacK fk = {};
    if constexpr (TAB == TAB_FAC) {
        fk.n_live = (uint32_t)p.n_live + zz; fk.magic = p.fac_magic + zz; fk.shift = (uint32_t)p.fac_shift + zz;
        fk.period = (uint32_t)p.fac_period + zz; fk.cells = (uint32_t)p.fac_cells + zz;
        fk.exit_w0 = p.fac_exit_w0 + zz; fk.exit_base = p.fac_exit_base + zz;
        fk.a_dir = smem_u32(smem + p.sm_fdir) + zz; fk.a_tail = smem_u32(smem + p.sm_ftail) + zz;
        fk.a_cls = smem_u32(smem + p.sm_fcls) + zz;
    }
    const uint32_t lt_mask = lanemask_lt();

    // ---- the lane's walker --------------------------------------------------------------------------
    int t = -1;                                  // steps taken; -1 = no walker
    uint32_t c0 = 0, ra = 0, rb = 0;             // (the high counter word of the lane's walker is chi_base: a lane's walker
                                                 // always belongs to the start the lane accumulates for)
    // ---- the start the lane accumulates for (changes on the cold path only) -----------------------
    int s = -1;
    long long s_end = 0, sid_base = 0;
    uint32_t s_end_off = 0, c0_base = 0, chi_base = 0;
    int t_init = 0;
    uint32_t det_w = 0, det_n = 0;               // DET: walkers of a deterministic-exit start
    uint32_t det_hs = 0, det_hn = 0;             // HIST: their one in-sum tally lands on the sink / on det_tgt
    int det_tgt = 0;
    uint32_t nsink = 0;                          // HIST: in-sum arrivals on sink_row
    uint2 srow = make_uint2(0, 0);
    uint32_t x = 0;                              // score of the lane's walkers of start s since the last flush
                                                 // (flushed at a range fetch once it passes 2^31; planner bound: range * nt * weight < 2^31)
    uint32_t nst = 0, nsurv = 0, novf = 0;       // nst: < 2^32 between range fetches (range * nt bound)
    bool onpath = false;                         // CNT: the lane's walker has followed its start's likely path so far
    unsigned long long nsteps = 0;
    // ---- warp-uniform id range [rbase + cur, rbase + end) ------------------------------------------
    long long rbase = 0;
    uint32_t cur = 0, end = 0;
    unsigned long long wwalk = 0;
    bool exhausted = false;

    auto weight_of = [&](uint32_t cls) -> uint32_t {
        return IDW ? cls : ((p.cwpack >> (cls * 8u)) & 0xffu);
    };
    auto tally_node = [&](int k, uint32_t c) {   // HIST (callers make sure the step is inside the sum)
        if (k == sink_row) nsink += c;
        else atomicAdd(hist_s + k, c);
    };
    auto det_target = [&](uint32_t w0, uint32_t w1) -> int {     // the single successor of a detabs row
        if constexpr (STR) return (int)(row_base(w1) + row_off(w0));
        else return (int)(row_base(w1) + (alias_s[row_off(w0)].y & 0xffffu));
    };
    auto flush_score = [&]() {
        if (p.score && s >= 0 && x != 0)
            atomicAdd(reinterpret_cast<unsigned long long*>(p.score) + s, (unsigned long long)x);
        x = 0;
    };
    // after sid_base / rbase / s_end changed.  s_end_off is the first range offset the hot path may not
    // take with the lane's current bases: the end of the start node, or the offset at which the low word
    // of the stream id wraps (the take then goes through next_start, which carries into bits 32..39).
    auto rebase = [&]() {
        const long long d = s_end - rbase;
        s_end_off = d <= 0 ? 0u : (d > 0x7fffffffll ? 0x7fffffffu : (uint32_t)d);
        c0_base = (uint32_t)sid_base;
        chi_base = ((uint32_t)((unsigned long long)sid_base >> 32) << 24) ^ rk[0];     // (with the first round key: pair)
        const uint32_t wrap_off = 0u - c0_base;              // c0_base + wrap_off == 2^32 (0: no wrap ahead)
        if (wrap_off != 0u && wrap_off < s_end_off) s_end_off = wrap_off;
    };
    auto pass_wrap = [&]() {
        chi_base = ((chi_base ^ rk[0]) + (1u << 24)) ^ rk[0];
        const long long d = s_end - rbase;
        s_end_off = d > 0x7fffffffll ? 0x7fffffffu : (uint32_t)d;
    };
    // cold: the id at range offset `off` belongs to a later start than the lane's current one
    auto next_start = [&](uint32_t off) {
        const long long gnew = rbase + (long long)off;
        if (s >= 0 && gnew < s_end) {
            // same start: the low word of the stream id wraps at or before `off` (rebase() clamped
            // s_end_off there).  c0_base + off wraps by itself; carry into bits 32..39 and restore the
            // start's real end (a range is far shorter than 2^32: no second wrap before the next rebase)
            pass_wrap();
            return;
        }
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
        srow = load_row2<TAB>(p, rowinfo_a, (uint32_t)__ldg(p.start_nodes + s), fk);
        sid_base = rbase + (p.start_id_base ? (__ldg(p.start_id_base + s) - s_lo) : 0ll);
        rebase();
        if (off >= s_end_off) pass_wrap();       // the new start's ids wrapped before this lane's first one
        t_init = ((srow.x & ROW_N_MASK) == 0u) ? -1 : 0;       // started on the sink: an exit, no step
        if constexpr (CNTS) cnt_a = smem_u32(cnt_s) + (uint32_t)((p.per_start ? s : 0) * p.t_cap * p.n_rows) * 4u;
        if constexpr (CNT && !CNTS) {
            path_a = smem_u32(path_s + s * p.t_cap);
            left_a = smem_u32(left_s + s * (p.t_cap + 1));
            cnt_g = p.count_tn + (long long)(p.per_start ? s : 0) * p.t_cap * (long long)p.n_rows;
        }
        const uint32_t srow_w0 = srow.x, srow_w1 = srow.y;
        if constexpr (STR && DET) srow.x = row_off(srow_w0);    // the take hands a new walker its direction (step below)
        // ... and on factored descriptors the CELL its start row points to instead of that cell's first row (FACJ, step)
        if constexpr (FACJ) srow.y = (uint32_t)(((unsigned long long)row_base(srow_w1) * fk.magic) >> fk.shift);
        if constexpr (DET) {
            // a start whose single successor is absorbing: every walker takes that one step (t = 1)
            // and exits; no draw can change it, so the take accounts for it and the lane stays free
            const bool det = (srow_w0 & ROW_DETABS) != 0u;
            det_n = det ? 1u : 0u;
            det_w = (det && t_last >= 2) ? weight_of(row_detcls(srow_w0)) : 0u;   // t = nt-1 is outside the sum
            if (det) t_init = -1;
            if constexpr (HIST) {
                det_tgt = det ? det_target(srow_w0, srow_w1) : 0;
                const bool in_sum = det && t_last >= 2;
                det_hs = (in_sum && det_tgt == sink_row) ? 1u : 0u;
                det_hn = (in_sum && det_tgt != sink_row) ? 1u : 0u;
            }
        }
    };
    auto take = [&](uint32_t off) {
        if (off >= s_end_off) [[unlikely]] next_start(off);
        c0 = c0_base + off;
        ra = srow.x; rb = srow.y;
        t = t_init;
        if constexpr (CNT && !CNTS) onpath = true;
        if constexpr (DET) {
            x += det_w; nst += det_n;
            if constexpr (HIST) { nsink += det_hs; if (det_hn) [[unlikely]] atomicAdd(hist_s + det_tgt, 1u); }
        }
    };

    while (true) {
        // ---- refill: one vote ranks the dead lanes; they take consecutive ids of the warp's range --
        const bool dead = t < 0;
        const unsigned need = __ballot_sync(NRW_FULL_MASK, dead);
        const uint32_t off = cur + (uint32_t)__popc(need & lt_mask);
        const uint32_t ncur = cur + (uint32_t)__popc(need);
        // hot: the range holds an id for every dead lane and (DET) at least one lane is walking -- one test, one branch
        if (ncur <= end && (!DET || need != NRW_FULL_MASK)) [[likely]] {
            cur = ncur;
            if constexpr (!HIST && !CNT && TAB == TAB_SMEM) {
                // (Tables in shared memory = coarse meshes = short walks: most lanes are refilled in every iteration.
                // On the refined mesh walkers live 17 steps, most iterations refill nobody and the divergent block
                // below is skipped altogether: there the selects cost 5 % -- 4.81e11 -> 4.58e11 -- and are not used.)
                // the take as selects behind ONE divergent branch (the cold change of start): two reconvergence points
                // and a branch less than `if (dead) take(off)`, all lanes in step (C5 0.403 -> 0.4015 ms; with every
                // step drawn, where every iteration refills most lanes, 0.674 -> 0.640 ms).  Round 1 had measured the
                // opposite (+1 %) on a loop with 134 instructions; on this one the branches weigh more.
                if (dead && off >= s_end_off) [[unlikely]] next_start(off);
                c0 = dead ? c0_base + off : c0;
                ra = dead ? srow.x : ra; rb = dead ? srow.y : rb;
                t = dead ? t_init : t;
                if constexpr (DET) { x += dead ? det_w : 0u; nst += dead ? det_n : 0u; }
            } else {
                if (dead) take(off);
            }
        } else if (ncur <= end) {
            // DET: the whole warp is idle.  Inside a deterministic-exit start (nothing there ever enters the step
            // phase) it accounts for every id the range and the start still hold in one go, lane L standing for ids
            // cur + L, cur + L + 32, ...  (No `continue` after it: every lane is still dead and falls through the
            // step phase; a second loop back-edge cost 1 % -- 0.523 -> 0.517 ms on C5.)
            if (__all_sync(NRW_FULL_MASK, det_n != 0u && off < s_end_off)) {
                const uint32_t lim = min(end, s_end_off);
                const uint32_t k = (lim - off + 31u) >> 5;         // (off = cur + lane here: every lane is dead)
                x += det_w * k;
                nst += k;
                if constexpr (HIST) { nsink += det_hs * k; if (det_hn) [[unlikely]] atomicAdd(hist_s + det_tgt, k); }
                cur = lim;
            } else {
                cur = ncur;
                take(off);
            }
        } else {
            // cold (once per id range): serve what the range still holds, then fetch the next one;
            // lanes left without an id idle for this one iteration
            if (dead && off < end) take(off);
            cur = end;
            if (!exhausted) {
                nsteps += nst; nst = 0;
                if (x >= 0x80000000u) flush_score();       // keeps the 32-bit running score from wrapping
                long long nb = 0, len = 0;
                if (!fetch_range(p, lane, nb, len)) exhausted = true;
                else {
                    cur = 0;
                    end = (uint32_t)len;
                    wwalk += end;
                    sid_base += nb - rbase;
                    rbase = nb;
                    rebase();
                }
            }
            if (exhausted && !__any_sync(NRW_FULL_MASK, t >= 0)) break;
        }

        // refilled and surviving lanes step together (without this the compiler sends the two groups through
        // the step code separately: 0.52 -> 0.83 ms)
        __syncwarp();

        // ---- one model step (fluence_mini_app.py:243-245 + the fused :428-434) -----------------------
        // Cold: the walker is within two steps of the end of time (tn >= nt - 2).  Step nt - 1 is outside the
        // solution sum (:428, Q3): it is taken and counted, not scored; a walker alive after it is a survivor.
        auto near_end = [&](const uint2 row, const uint32_t knew, const int tn) {
            const bool absorbed = (row.x & ROW_N_MASK) == 0u;
            if (tn < t_last) {
                x += weight_of(row_class(row.x));
                if constexpr (HIST) tally_node((int)(knew & ROW_BASE_MASK), 1u);
            }
            ++nst;
            t = tn;
            if (absorbed) { t = -1; return; }
            if (tn >= t_last) { ++nsurv; t = -1; return; }
            if constexpr (DET) {
                // tn == nt - 2 on a row whose single successor is absorbing: its exit step is step nt - 1 --
                // a step and an exit, outside the sum
                if (row.x & ROW_DETABS) { ++nst; t = -1; }
            }
        };
        // CNT: leave event of a walker that was on its start's path until step l - 1 (l = first step off the path,
        // the step after its last one, or t_cap)
        auto leave_path = [&](int l) {
            atomicAdd(left_s + s * (p.t_cap + 1) + min(l, p.t_cap), 1u);
            onpath = false;
        };
        // NEAR: the pair may reach the last two steps of time (cold); the hot pair starts at t < p.t_fast = nt - 4, so
        // neither of its steps can, and carries no end-of-time test at all
        auto step = [&](auto near_c, uint32_t slot, uint32_t coin) -> bool {     // true: the walker goes on
            constexpr bool NEAR = decltype(near_c)::value;
            const int tn = t + 1;
            // target row; rb still carries its flag bits, which the descriptor look-up drops (load_row2)
            uint32_t knew;
            uint32_t sel = 0;
            uint2 row;
            if constexpr (FACJ) {
                // Factored descriptors, structured draw: the walker carries (direction, cell) -- rb is the CELL the row it
                // stands on points to, i.e. the cell of the row it is about to land on.  New row = (sel, rb): its index
                // needs no division (fac_row derives direction and cell from the row index: four instructions), and
                // the first row of its target cell is never needed -- only the target cell itself.  Same descriptor
                // words as fac_row_k for every live row (flags, inc, sinc); rows >= n_live (the sink) are never landed on
                // by a draw (a stay row's targets are the rows of a live cell).
                sel = (coin <= str_thr) ? slot : ra;
                const uint2 d = lds64(fk.a_dir + (sel << 3));
                knew = rb * fk.period + sel;
                const uint32_t jn = rb + d.y;                          // the new row's target cell (wraps when negative)
                const uint32_t cls = (lds32(fk.a_cls + ((knew >> 4) << 2)) >> ((knew & 15u) * 2u)) & 3u;
                const bool leaves = jn >= fk.cells;
                row.x = (leaves ? fk.exit_w0 : d.x) + cls * 0x28000000u;
                row.y = leaves ? fk.exit_base : 0x40000000u;           // flag bits (sinc, detlast); plus the exit row's base
                ra = sel; rb = jn;
            } else {
            if constexpr (STR && DET) {
                // ra is the walker's DIRECTION, not its descriptor word: the offset a walker was sent to IS the `same`
                // of the row it landed on (row = first row of the cell + direction; tables.structure_rows verifies it
                // for every row a walker can draw from), and the length is the launch constant -- nothing of w0 is
                // needed for the next draw
                sel = (coin <= str_thr) ? slot : ra;
                knew = rb + sel;
            } else if constexpr (STR) {
                knew = rb + ((coin <= str_thr) ? slot : row_off(ra));
            } else {
                const uint2 e = alias_s[row_off(ra) + slot];
                knew = rb + ((coin <= e.x) ? (e.y & 0xffffu) : (e.y >> 16));
            }
            row = load_row2<TAB>(p, rowinfo_a, knew, fk);
            ra = (STR && DET) ? sel : row.x; rb = row.y;
            }
            if constexpr (CNT) {
                const int k = (int)(knew & ROW_BASE_MASK);
                if constexpr (CNTS) {
                    if (tn < t_hot) [[likely]] {
                        // the common step: inside the window, not the last step of time -- one shared-memory atomic on
                        // the lane's start's bins, and the walk goes on unless the row is absorbing
                        red_shared_inc(cnt_a + (((uint32_t)tn * n_rows_u + (uint32_t)k) << 2));
                        ++nst;
                        const bool absorbed = (row.x & ROW_N_MASK) == 0u;
                        t = absorbed ? -1 : tn;
                        return !absorbed;
                    }
                    if (tn < p.t_cap) atomicAdd(cnt_s + ((p.per_start ? s : 0) * p.t_cap + tn) * p.n_rows + k, 1u);
                    else ++novf;
                } else if (tn < t_hot) {
                    // (NO [[likely]] on this line: with the attribute on an `else if` that continues an `if constexpr`
                    // chain, nvcc 12.9 dropped the statement that FOLLOWS the chain -- the `++nst` below -- from the
                    // instantiation in which this branch is discarded; seen in the PTX, caught by the port's step count)
                    // the common step (inside the window, before the last step of time), with the lane's start's
                    // bases kept per lane: a walker on its start's path is credited in the epilogue, one that leaves
                    // it records the step, steps off the path go to their bins
                    if (!(onpath && (uint32_t)k == lds32(path_a + ((uint32_t)tn << 2)))) {
                        if (onpath) { red_shared_inc(left_a + ((uint32_t)tn << 2)); onpath = false; }
                        atomicAdd(cnt_g + ((size_t)(uint32_t)tn * n_rows_u + (uint32_t)k), 1u);
                    }
                    ++nst;
                    const bool absorbed = (row.x & ROW_N_MASK) == 0u;
                    if (absorbed && onpath) { red_shared_inc(left_a + ((uint32_t)(tn + 1) << 2)); onpath = false; }
                    t = absorbed ? -1 : tn;
                    return !absorbed;
                } else if (tn < p.t_cap) {
                    if (!(onpath && k == path_s[s * p.t_cap + tn])) {      // on the path: credited in the epilogue
                        if (onpath) leave_path(tn);
                        const long long gq = p.per_start ? s : 0;
                        atomicAdd(p.count_tn + (gq * p.t_cap + tn) * (long long)p.n_rows + k, 1u);
                    }
                } else {
                    ++novf;
                    if (onpath) leave_path(p.t_cap);
                }
                // every step is counted as it is taken; the walk ends on an absorbing row or after step nt - 1
                ++nst;
                const bool absorbed = (row.x & ROW_N_MASK) == 0u;
                if (absorbed || tn >= t_last) {
                    if (!absorbed) ++nsurv;
                    if (onpath) leave_path(tn + 1);
                    t = -1;
                } else {
                    t = tn;
                }
                return t >= 0;
            }
            if constexpr (NEAR) {
                if (tn >= t_last - 1) { near_end(row, knew, tn); return t >= 0; }
            }
            // Hot path, branch-free: this step and -- on a row whose single successor is absorbing (DET: the
            // fluence exit rows, fluence_mini_app.py:355-357; no draw could change that step) -- the exit step it
            // implies are both inside the sum; what they add to score and step count sits in the top bits of the
            // descriptor (inc, sinc: common.cuh), one shift-add each.
            if constexpr (IDW) {
                x += DET ? row_inc(row.x) : row_class(row.x);
            } else {
                x += weight_of(row_class(row.x));
                if constexpr (DET) x += (row.x & ROW_DETABS) ? weight_of(row_detcls(row.x)) : 0u;
            }
            nst += DET ? row_sinc(row.y) : 1u;
            if constexpr (HIST) {
                tally_node((int)(knew & ROW_BASE_MASK), 1u);
                if constexpr (DET) {
                    if (row.y & ROW_DETLAST) ++nsink;                         // the exit step, onto the sink
                    else if (row.x & ROW_DETABS) [[unlikely]] tally_node(det_target(row.x, row.y), 1u);
                }
            }
            // the walk ends here when the row is absorbing or (DET) its single successor is
            const bool dead = (row.x & (DET ? ROW_TERM : ROW_N_MASK)) == (DET ? ROW_TERM : 0u);
            t = dead ? -1 : tn;
            return !dead;                        // (the branch after the first step of a pair tests the flag itself)
        };
        // one Philox call serves the two steps of a pair (common.cuh, "paired draws"): lanes are refilled
        // only here, so every live lane is at an even t and the whole warp shares the call
        // length of the row the walker draws from.  STR && DET: a launch constant (nrw_tables.str_n) -- every row
        // that is neither absorbing nor stepped without a draw has that length, so nothing is extracted from the
        // descriptor and the multiply of the second step does not wait for the first step's look-up
        auto rlen = [&](uint32_t w0) -> uint32_t {
            if constexpr (STR && DET) return str_n;
            else return row_len(w0);
        };
        auto pair = [&](auto near_c) {
            // counter = (stream id low word, id bits 32..39 << 24 | t + 1); chi_base holds the id bits XOR the first round
            // key (its low 24 bits are the key's: t + 1 < 2^24 enters by XOR as it would by OR), so the first round's
            // "^ key ^ c1" is one three-input LOP3
            const uint2 r = philox2x32_10_k0(c0, chi_base ^ (uint32_t)(t + 1), rk);
            const unsigned long long va = (unsigned long long)r.x * rlen(ra);
            if constexpr (!decltype(near_c)::value && STR && DET && IDW && TAB == TAB_SMEM && !HIST && !CNT) {
                // The flagship pair with its second step written out: draw, choice and (clamped) address of the second
                // step do not depend on the first step's look-up, so they are issued while that look-up is in flight
                // (the loop waits on fixed latencies more than on anything else: DESIGN.md section 3).  Lanes whose
                // walker ended in the first step look up row 0 and discard it.  C5 0.4005 -> 0.3935 ms.
                const bool a1 = step(near_c, (uint32_t)(va >> 32), (uint32_t)va);
                const unsigned long long vb = (unsigned long long)r.y * str_n;
                const uint32_t sel = ((uint32_t)vb <= str_thr) ? (uint32_t)(vb >> 32) : ra;
                const uint2 row = lds64(rowinfo_a + ((a1 ? rb + sel : 0u) << 3));
                if (a1) {
                    x += row_inc(row.x); nst += row_sinc(row.y);
                    ra = sel; rb = row.y;
                    t = (row.x & ROW_TERM) ? -1 : t + 1;
                }
                return;
            }
            if (step(near_c, (uint32_t)(va >> 32), (uint32_t)va)) {
                const unsigned long long vb = (unsigned long long)r.y * rlen(ra);
                step(near_c, (uint32_t)(vb >> 32), (uint32_t)vb);
            }
        };
        if constexpr (CNT) {
            if (t >= 0) pair(std::true_type{});
        } else {
            // one unsigned test: the lane holds a walker (t >= 0) AND both steps of the pair stay clear of the end of time
            if ((uint32_t)t < t_fast) pair(std::false_type{});
            else if (t >= 0) [[unlikely]] pair(std::true_type{});
        }
    }

    // ---- epilogue ------------------------------------------------------------------------------------
    flush_score();
    unsigned long long v0 = nsteps + nst;
    uint32_t v2 = nsurv;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        v0 += __shfl_xor_sync(NRW_FULL_MASK, v0, o);
        v2 += __shfl_xor_sync(NRW_FULL_MASK, v2, o);
    }
    if (lane == 0 && p.counters) {
        if (v0) atomicAdd(p.counters + 0, v0);
        if (wwalk - v2) atomicAdd(p.counters + 1, wwalk - v2);          // exits = walkers - survivors
        if (v2) atomicAdd(p.counters + 2, (unsigned long long)v2);
        if (wwalk) atomicAdd(p.counters + 4, wwalk);
    }
    if constexpr (CNT) {
        uint32_t v3 = novf;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v3 += __shfl_xor_sync(NRW_FULL_MASK, v3, o);
        if (lane == 0 && p.counters && v3) atomicAdd(p.counters + 3, (unsigned long long)v3);
        __syncthreads();
        if constexpr (CNTS) {
            for (int i = tid; i < p.cnt_bins; i += (int)blockDim.x) {
                const uint32_t h = cnt_s[i];
                if (h) atomicAdd(p.count_tn + i, h);
            }
        }
        // walkers on the path at step tau = those that left it later: suffix sums of the leave-time histogram
        for (int sidx = tid; !CNTS && sidx < p.n_starts; sidx += (int)blockDim.x) {
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
    if constexpr (HIST) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) nsink += __shfl_xor_sync(NRW_FULL_MASK, nsink, o);
        if (lane == 0 && nsink) atomicAdd(hist_s + sink_row, nsink);
        __syncthreads();
        for (int i = tid; i < p.n_rows; i += (int)blockDim.x) {
            const uint32_t h = hist_s[i];
            if (h) atomicAdd(p.node_hist + i, (unsigned long long)h);
        }
    }
    walk_tail(p.tail);
}

}  // namespace nrw
