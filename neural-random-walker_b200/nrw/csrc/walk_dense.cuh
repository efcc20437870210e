// walk_dense.cuh -- K1d: the score walk with DENSE step batches (per-warp survivor queues).
//
// Same semantics as walk_score_kernel<TAB_SMEM, IDW, DET=true> (walk_score.cuh) -- identical Philox
// streams (paired draws, common.cuh), alias draws, score tally and counters; all three kernels are
// held to the numpy port bit for bit -- but the warp no longer carries one walker per lane through a
// refill / step / step iteration.  In walk_score.cuh a pair iteration issues ~134 warp-instructions:
// ~37 of them rank the dead lanes and re-initialise them (executed with ~60 % of the lanes), the
// second step of the pair runs with the ~60 % of lanes that survived the first, and the accounting of
// the lanes that end is issued under divergence.  Here every step is issued for a FULL batch of 32
// walkers that are all at the same phase of their draw pair:
//
//   * the warp works through its id range one SEGMENT at a time (ids of one start node, not crossing a
//     2^32 boundary of the stream id), so the start row, the stream-id base and the score target are
//     warp-uniform;
//   * a FRESH batch gives lanes 0..31 the next 32 ids (no vote, no ranking), makes the Philox call of
//     the pair (t = 0) and takes the odd step on the start row;
//   * survivors are compacted (one vote + popc) into a per-warp ring in shared memory: Q1 holds walkers
//     that have taken an odd step together with word 1 of their call, Q0 walkers that have taken an
//     even step and need a new call.  An entry is (stream id low word, word 1, node, t) = 16 bytes;
//   * whenever a ring holds 32 walkers the warp pops them (one LDS.128 per lane) and steps them as a
//     dense batch: a Q1 batch takes the even step from the stored word, a Q0 batch makes the next call
//     and takes the odd step.  Rings are drained (partial batches) only at the end of a segment;
//   * rows with a single absorbing successor are stepped without a draw and starts on such rows are
//     accounted for in bulk, exactly as in walk_score.cuh (DET).
//
// Ring capacity: a fresh or Q0 batch pushes <= 32 walkers onto Q1 and runs only while Q1 holds < 32; a
// Q1 batch pushes <= 32 onto Q0 and runs only while Q0 holds < 32 (a full Q0 would have been popped
// before the batch that made Q1 full, see the schedule below) -- 64 entries per ring suffice.
// Requires TAB_SMEM tables, nt >= 2, score tally only.  Shared memory: tables + 2 KB per warp.
#pragma once
#include "walk_score.cuh"

namespace nrw {

constexpr int DQ_ENTRIES = 64;                         // per ring
constexpr int DQ_BYTES_PER_WARP = 2 * DQ_ENTRIES * 16; // Q1 then Q0

template <bool IDW, int CPS>
__global__ void __launch_bounds__(CPS == 1 ? 1024 : 768, CPS)
walk_dense_kernel(const __grid_constant__ WalkParams p) {
    extern __shared__ __align__(16) unsigned char smem[];
    uint64_t* bar = reinterpret_cast<uint64_t*>(smem);
    const uint2* rowinfo_s = reinterpret_cast<const uint2*>(smem + p.sm_rowinfo);
    const uint2* alias_s = reinterpret_cast<const uint2*>(smem + p.sm_alias);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    uint4* const q1 = reinterpret_cast<uint4*>(smem + p.sm_queue + warp * DQ_BYTES_PER_WARP);
    uint4* const q0 = q1 + DQ_ENTRIES;
    stage_tables_issue<TAB_SMEM>(p, bar, reinterpret_cast<uint2*>(smem + p.sm_rowinfo),
                                 reinterpret_cast<uint2*>(smem + p.sm_alias), tid);
    mbar_wait(bar, 0);
    __syncthreads();

    const int t_last = p.t_last;                 // nt - 1 >= 1
    const uint32_t lt_mask = lanemask_lt();

    // ---- per-lane tallies ---------------------------------------------------------------------------
    uint32_t x = 0;                              // score of the current segment's walkers stepped by this lane
    uint32_t nst = 0, nsurv = 0;                 // walker-steps (lifetimes at death) since the last range fetch
    unsigned long long nsteps = 0;
    // ---- warp-uniform state ---------------------------------------------------------------------------
    long long rbase = 0;                         // the warp's id range is [rbase + cur, rbase + end)
    uint32_t cur = 0, end = 0;
    unsigned long long wwalk = 0;
    int s = -1;                                  // start index of the current segment
    long long s_end = 0, sid_off = 0;            // its end id; stream id = global id + sid_off
    uint2 srow = make_uint2(0, 0);

    auto weight_of = [&](uint32_t cls) -> uint32_t {
        return IDW ? cls : ((p.cwpack >> (cls * 8u)) & 0xffu);
    };

    // one model step of a live walker on row (ra, rb) at step count t with draw word `word`
    // (fluence_mini_app.py:243-245 + the fused :428-434); true if the walker goes on, on row `knew`
    auto step = [&](uint32_t ra, uint32_t rb, uint32_t word, int t, int& knew) -> bool {
        const unsigned long long v = (unsigned long long)word * row_len(ra);
        const int tn = t + 1;
        const uint2 e = alias_s[row_off(ra) + (uint32_t)(v >> 32)];
        knew = (int)(rb + (((uint32_t)v <= e.x) ? (e.y & 0xffffu) : (e.y >> 16)));
        const uint2 row = rowinfo_s[knew];
        x += IDW ? (row.x >> 30) : ((p.cwpack >> ((row.x >> 27) & 0x18u)) & 0xffu);
        // the walk ends here when the row is absorbing, when its single successor is, or at the last step
        const bool terminal = (row.x & ROW_TERM) != 0u;
        if (terminal || tn >= t_last) {
            if (tn + 1 < t_last) [[likely]] {
                // absorbing row, or a row whose single successor is absorbing (fluence exit rows,
                // fluence_mini_app.py:355-357): that exit step is taken now -- no draw could change it
                if constexpr (IDW) {
                    x += row_detcls(row.x);          // 0 on rows that are not such exit rows
                    nst += (row.x >> 27) & 1u;
                } else {
                    if (row.x & ROW_DETABS) { x += weight_of(row_detcls(row.x)); ++nst; }
                }
                nst += (uint32_t)tn;
            } else {
                // within one step of the end of time: step nt-1 is outside the sum (:428, Q3)
                int tl = tn;
                bool absorbed = (row.x & ROW_N_MASK) == 0u;
                uint32_t wl = weight_of(row_class(row.x));
                if (!absorbed && tn < t_last) {      // terminal, not absorbing: the deterministic exit step
                    wl = weight_of(row_detcls(row.x));
                    x += wl;
                    tl = tn + 1;                     // == t_last
                    absorbed = true;
                }
                if (tl >= t_last) {
                    x -= wl;
                    if (!absorbed) ++nsurv;
                }
                nst += (uint32_t)tl;
            }
            return false;
        }
        return true;
    };
    // Ring heads / tails are monotone BYTE counters (16 per entry); an entry's offset is counter mod 1 KB.
    // survivors of a batch go to the back of ring q, in lane order
    auto ring_at = [](uint4* q, uint32_t byte_pos) -> uint4* {
        return reinterpret_cast<uint4*>(reinterpret_cast<unsigned char*>(q) + (byte_pos & (DQ_ENTRIES * 16 - 16)));
    };
    auto push = [&](uint4* q, uint32_t& tail, bool alive, uint32_t c0, uint32_t w1, int k, int t) {
        const unsigned b = __ballot_sync(NRW_FULL_MASK, alive);
        if (alive) *ring_at(q, tail + ((uint32_t)__popc(b & lt_mask) << 4)) = make_uint4(c0, w1, (uint32_t)k, (uint32_t)t);
        tail += (uint32_t)__popc(b) << 4;
    };
    const uint32_t lane16 = (uint32_t)lane << 4;

    while (true) {
        // ---- next id range of the warp (one atomic per range) ---------------------------------------
        if (cur >= end) {
            nsteps += nst; nst = 0;
            long long a = 0;
            if (lane == 0) a = (long long)atomicAdd(p.work_next, (unsigned long long)p.range);
            a = __shfl_sync(NRW_FULL_MASK, a, 0);
            const long long nb = p.lo + a;
            if (nb >= p.hi) break;
            rbase = nb; cur = 0;
            end = (uint32_t)min(p.range, p.hi - nb);
            wwalk += end;
        }
        // ---- next segment: ids of one start node, stream ids within one 2^32 block -------------------
        const long long g = rbase + (long long)cur;
        if (s < 0 || g >= s_end) {
            long long s_lo = s_end;
            bool found = false;
            if (s >= 0 && s + 1 < p.n_starts) {
                const long long e2 = __ldg(p.start_prefix + s + 2);
                if (g < e2) { ++s; s_end = e2; found = true; }
            }
            if (!found) {
                int lo_i = 0, hi_i = p.n_starts;
                while (hi_i - lo_i > 1) {
                    const int mid = (lo_i + hi_i) >> 1;
                    if (__ldg(p.start_prefix + mid) <= g) lo_i = mid; else hi_i = mid;
                }
                s = lo_i;
                s_lo = __ldg(p.start_prefix + s);
                s_end = __ldg(p.start_prefix + s + 1);
            }
            srow = rowinfo_s[__ldg(p.start_nodes + s)];
            sid_off = p.start_id_base ? (__ldg(p.start_id_base + s) - s_lo) : 0ll;
        }
        uint32_t seg_end = (uint32_t)min((long long)end, s_end - rbase);
        const unsigned long long sid0 = (unsigned long long)(g + sid_off);
        const uint32_t c0_first = (uint32_t)sid0;
        const uint32_t chi = (uint32_t)(sid0 >> 32) << 24;
        const uint32_t to_wrap = 0u - c0_first;              // ids before the low word wraps (0: 2^32)
        if (to_wrap != 0u && to_wrap < seg_end - cur) seg_end = cur + to_wrap;
        const uint32_t n_seg = seg_end - cur;                // >= 1

        if (row_len(srow.x) == 0u) {                         // started on an absorbing row: an exit, no step
            cur = seg_end;
            continue;
        }
        if (srow.x & ROW_DETABS) {
            // a start whose single successor is absorbing: every walker takes that one step (t = 1) and
            // exits; lane L stands for ids L, L + 32, ... of the segment
            const uint32_t k = (n_seg - (uint32_t)lane + 31u) >> 5;
            if (t_last >= 2) x += weight_of(row_detcls(srow.x)) * k;     // t = nt-1 is outside the sum
            nst += k;
        } else {
            uint32_t h1 = 0, t1 = 0, h0 = 0, t0 = 0;         // ring heads / tails in bytes (monotone; offset mod 1 KB)
            uint32_t fresh = cur;
            // The three batch kinds; FULL batches (all 32 lanes) are the steady state and carry no lane
            // predicate, partial ones only occur at the end of a segment.
            // Q1 batch: the even step of the pair, from the stored word; survivors -> Q0
            auto q1_batch = [&](auto full, uint32_t cnt) {
                constexpr bool FULL = decltype(full)::value;
                bool alive = false;
                uint32_t c0 = 0; int knew = 0, tn = 0;
                if (FULL || (uint32_t)lane < cnt) {
                    const uint4 e = *ring_at(q1, h1 + lane16);
                    const uint2 row = rowinfo_s[e.z];
                    c0 = e.x; tn = (int)e.w + 1;
                    alive = step(row.x, row.y, e.y, (int)e.w, knew);
                }
                h1 += cnt << 4;
                push(q0, t0, alive, c0, 0u, knew, tn);
            };
            // Q0 batch: the next Philox call, then the odd step; survivors -> Q1 with word 1 of the call
            auto q0_batch = [&](auto full, uint32_t cnt) {
                constexpr bool FULL = decltype(full)::value;
                bool alive = false;
                uint32_t c0 = 0, w1 = 0; int knew = 0, tn = 0;
                if (FULL || (uint32_t)lane < cnt) {
                    const uint4 e = *ring_at(q0, h0 + lane16);
                    const uint2 row = rowinfo_s[e.z];
                    c0 = e.x; tn = (int)e.w + 1;
                    const uint2 r = philox2x32_10_rk(c0, chi | (uint32_t)tn, p);
                    w1 = r.y;
                    alive = step(row.x, row.y, r.x, (int)e.w, knew);
                }
                h0 += cnt << 4;
                push(q1, t1, alive, c0, w1, knew, tn);
            };
            // fresh batch: the next <= 32 ids, first call (t = 0), odd step on the start row; survivors -> Q1
            auto fresh_batch = [&](auto full, uint32_t cnt) {
                constexpr bool FULL = decltype(full)::value;
                bool alive = false;
                const uint32_t c0 = c0_first + (fresh - cur) + (uint32_t)lane;
                uint32_t w1 = 0; int knew = 0;
                if (FULL || (uint32_t)lane < cnt) {
                    const uint2 r = philox2x32_10_rk(c0, chi | 1u, p);
                    w1 = r.y;
                    alive = step(srow.x, srow.y, r.x, 0, knew);
                }
                fresh += cnt;
                push(q1, t1, alive, c0, w1, knew, 1);
            };
            // Schedule: a full ring is popped first (Q1 before Q0), then fresh ids; partial batches drain
            // what is left.  Pushes onto Q1 (Q0 / fresh batches) happen only while Q1 holds < 32, pushes
            // onto Q0 (Q1 batches) only while Q0 holds < 32: no ring ever holds more than 63 walkers.
            while (true) {
                if (t1 - h1 >= 512u)          q1_batch(std::true_type{}, 32u);
                else if (t0 - h0 >= 512u)     q0_batch(std::true_type{}, 32u);
                else if (seg_end - fresh >= 32u) fresh_batch(std::true_type{}, 32u);
                else if (seg_end != fresh)    fresh_batch(std::false_type{}, seg_end - fresh);
                else if (t1 != h1)            q1_batch(std::false_type{}, (t1 - h1) >> 4);
                else if (t0 != h0)            q0_batch(std::false_type{}, (t0 - h0) >> 4);
                else break;                                  // segment done, rings empty
                __syncwarp();                                // ring writes visible to the lanes that pop them
            }
        }
        cur = seg_end;
        // the segment's score: one atomic per warp (the sum over the warp stays below 2^31: planner
        // bound range * nt * weight < 2^31)
        const uint32_t xs = __reduce_add_sync(NRW_FULL_MASK, x);
        x = 0;
        if (lane == 0 && xs != 0u)
            atomicAdd(reinterpret_cast<unsigned long long*>(p.score) + s, (unsigned long long)xs);
    }

    // ---- epilogue ------------------------------------------------------------------------------------
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
}

}  // namespace nrw
