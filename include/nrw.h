/* nrw.h -- C ABI of libnrw.so, the B200 (sm_100a) random-walk engine that replaces the
 * conventional-compute (`--run_mode pc`) back end of sandialabs/Neural-Random-Walker.
 *
 * The reference is pure Python and has no FFI of its own, so there is no existing binding
 * to mirror; each entry point below names the reference code it replaces
 * (paths relative to the reference root).  The Python host layer
 * (neural-random-walker_b200/nrw/engine.py) binds these with ctypes and is the only
 * caller; INTEGRATION.md shows the stub a reference maintainer would add.
 *
 * Conventions
 *   - every function returns 0 on success, a negative NRW_E* code on failure and never
 *     throws; nrw_last_error() returns a thread-local message for the last failure;
 *   - all pointers in nrw_tables / nrw_walk / nrw_replay are DEVICE pointers owned by the
 *     caller (PyTorch tensors in the Python host); the library allocates only a few
 *     hundred bytes of per-handle scratch; functions whose name ends in _host take HOST
 *     pointers and do their own staging;
 *   - `stream` is a cudaStream_t passed as void*; NULL means the legacy default stream;
 *   - nothing here synchronises unless the comment says so;
 *   - threading: an engine may be used from several host threads and several streams at once.  Every
 *     walk launch takes its own ticket slot (work hand-out counter) from a ring of NRW_TICKET_SLOTS
 *     slots that the launch itself hands back when its last block retires, so launches of one engine
 *     that overlap on different streams do not share any device state (at most NRW_TICKET_SLOTS may
 *     be in flight at once).  The *_host entries stage through per-engine buffers and serialise
 *     behind a per-engine mutex;
 *   - environment: the measurement switches (NRW_K1, NRW_DET, NRW_TAB, NRW_CPS, NRW_RANGES_PER_WARP,
 *     NRW_RANGE_TAIL, NRW_TAIL_RANGES, NRW_CNT_MATCH, NRW_CNT_CACHE, NRW_CNT_PATH, NRW_STR)
 *     are read ONCE, in nrw_create, and stored in the handle (nrw_engine_config prints them); no
 *     launch calls getenv.  None of them changes a result.
 */
#ifndef NRW_H_
#define NRW_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define NRW_ABI_VERSION 5

#if defined(__GNUC__)
#define NRW_API __attribute__((visibility("default")))
#else
#define NRW_API
#endif

#define NRW_OK            0
#define NRW_EINVAL       -1   /* bad argument / unsupported combination            */
#define NRW_ECUDA        -2   /* a CUDA runtime call failed (see nrw_last_error)    */
#define NRW_ENOSPC       -3   /* shared memory / buffer capacity exceeded           */

typedef struct nrw_engine nrw_engine;   /* opaque handle: one per (process, device) */
#define NRW_TICKET_SLOTS 1024

/* ---- table layouts -------------------------------------------------------------- */
/* alias_layout values */
#define NRW_ALIAS_NONE    0
#define NRW_ALIAS_COMPACT 1   /* rowinfo 8 B/row, alias 8 B/entry, <= 2^15 entries (shared memory) */
#define NRW_ALIAS_WIDE    2   /* rowinfo 16 B/row, alias 16 B/entry (any graph, global)  */

/* COMPACT rowinfo  (uint32 w0, w1):
 *                  w0 = off[0:15) | n[15:25) | detabs[25] | term[26] | cls[27:29) | inc[29:32)
 *                  w1 = base[0:29) | detlast[29] | sinc[30:32)
 *                  detabs = 1: the row has ONE successor and it is absorbing (an exit row of
 *                  fluence_mini_app.py:355-357); detlast = 1: that successor is row n_rows - 1;
 *                  term = (n == 0) | detabs; cls = source class of the row.  The score kernel steps detabs
 *                  rows without a draw (the draw cannot change the outcome), and inc / sinc are what
 *                  ARRIVING on the row then adds to a walker's score and step count:
 *                  inc = cls + (detabs ? class of the successor : 0), sinc = 1 + detabs.  They sit in the
 *                  top bits so that one shift-add each applies them and so that base + offset, flags
 *                  included, still addresses the right descriptor in shared memory ((k << 3) drops them).
 *         alias    (uint32 thresh, ab): target = base + ((coin <= thresh) ? ab & 0xffff : ab >> 16),
 *                  slot and coin = high and low word of w * n, w = one 32-bit Philox word (csrc/common.cuh)
 * WIDE    rowinfo  (uint32 off, n, base, cls)
 *         alias    (uint32 thresh, a, b, 0): target = base + ((r1 <= thresh) ? a : b)
 * A row with n == 0 is absorbing: a walker is tallied on arrival and never stepped
 * (the sink row of fluence_mini_app.py:241,381 and the absorbed state of
 * simple_flux_miniapp.py:285).  `off` indexes the (de-duplicated) alias pattern. */

typedef struct nrw_tables {
    int32_t  n_rows;          /* meshSize + 1 (fluence) / 3 (simple_flux)                 */
    int32_t  n_live;          /* rows < n_live are stepped (meshSize)                     */
    /* CSR with per-row sequential fp64 prefix sums == non-zeros of
     * np.cumsum(myQ, axis=1) (fluence_mini_app.py:224).  Needed by the cdf sampler and
     * by replay; may be NULL for alias-only walks.                                       */
    const int32_t* indptr;    /* [n_rows + 1]                                             */
    const int32_t* cols;      /* [nnz]                                                    */
    const double*  cum;       /* [nnz]                                                    */
    const uint8_t* node_class;/* [n_rows] source class 0..3 (cdf sampler); may be NULL    */
    int32_t  alias_layout;    /* NRW_ALIAS_*                                              */
    int32_t  n_alias;         /* entries in `alias`                                       */
    const void* rowinfo;      /* [n_rows] per alias_layout                                */
    const void* alias;        /* [n_alias] per alias_layout                               */
    /* Optional, COMPACT layout only: the row descriptors in factored form, for meshes with the
     * (direction, cell) structure of the fluence matrix (fluence_mini_app.py:331-381: live row
     * k = i + j * fac_period; nrw/tables.py::factorise_rows builds and verifies it).  When
     * `rowinfo` does not fit shared memory next to the alias patterns but these do, the walk derives
     * each descriptor from them instead of loading it from global memory:
     *   jn = j + shift[i];  inside [0, fac_cells): (dir_w0[i] | class << 30, jn * fac_period)
     *                       outside:               (fac_exit_w0 | class << 30, fac_exit_base)
     * All NULL / 0 when absent.                                                            */
    const void* fac_dir;      /* uint32 [fac_period][2]: (dir_w0, shift as int32)          */
    const void* fac_tail;     /* uint32 [fac_n_tail][2]: descriptors of rows >= n_live     */
    const void* fac_cls;      /* uint32 [(n_live + 15) / 16]: 2-bit classes of live rows   */
    int32_t  fac_period, fac_cells, fac_n_tail;
    int32_t  fac_shift;       /* k / fac_period == (k * fac_magic) >> fac_shift, k < n_live */
    uint32_t fac_magic, fac_exit_w0, fac_exit_base;
    /* Optional (NRW_SAMPLER_TREE): the probability circuit of loihirw/rw_core.py:11-156 as a device table.
     * Row r's successors (its CSR entries, padded with impossible outputs to n2 = the next power of two) are the
     * leaves of a complete binary tree; tree[tree_ptr[r] + h] is the firing threshold of the probability neuron
     * at heap index h (root 0, children 2h+1 / 2h+2): floor(128 * p_left), i.e. the 8-bit synaptic weight
     * 128*p+36 of rw_core.py:151-153 with its offset removed.  A draw walks log2(n2) levels, one 7-bit
     * uniform per level: left iff u7 < threshold.  NULL when absent.                                    */
    const uint8_t* tree;      /* [tree_ptr[n_rows]]                                        */
    const int32_t* tree_ptr;  /* [n_rows + 1]; row r owns n2(r) - 1 entries (0 for n <= 1) */
    /* Optional, COMPACT layout only: the STRUCTURED form of the same tables, for matrices whose alias tables have
     * the closed form of the fluence matrix (fluence_mini_app.py:368-371: keep the direction with probability
     * 1 - pone, else scatter uniformly): every alias entry s of a row is (str_thresh, own = s, alias = same), one
     * threshold for the whole matrix, so that  target = base + (coin <= str_thresh ? slot : same)  needs no alias
     * look-up.  rowinfo_str = rowinfo with `same` in place of the pattern offset (w0 bits [0:15));
     * nrw/tables.py::structure_rows builds it and verifies every entry, so the structured draw yields the alias
     * draw's target for every (slot, coin).  The score walk takes it when present (one shared-memory look-up per
     * step instead of two, no alias patterns staged).  str_n: the one length of every row a walker can draw from
     * when rows with a single absorbing successor are stepped without a draw -- all rows that are neither
     * absorbing nor such a row (the stay rows of the fluence matrix all hold ND-1 entries); structure_rows
     * verifies it and offers no structured form otherwise.  The walk then multiplies by a launch constant instead
     * of extracting the length from the descriptor (ABI v5; must be >= 1 when rowinfo_str is given).  fac_*_str: the factored form of rowinfo_str (period, cells,
     * shift, magic, class bits and exit base are those of fac_*).  All NULL / 0 when absent.            */
    const void* rowinfo_str;  /* [n_rows] uint32 (w0, w1)                                   */
    uint32_t str_thresh;
    uint32_t str_n;
    uint32_t fac_exit_w0_str;
    const void* fac_dir_str;  /* uint32 [fac_period][2]                                     */
    const void* fac_tail_str; /* uint32 [fac_n_tail][2]                                     */
} nrw_tables;

/* ---- Philox walk (K1) -------------------------------------------------------------- */
#define NRW_SAMPLER_ALIAS 0   /* O(1) alias draw (WIDE: slot = mulhi(r0, n), coin = r1)   */
#define NRW_SAMPLER_CDF   1   /* the reference's searchsorted(left) on the fp64 prefix
                                 sums with u = ((r0>>5)*2^26 + (r1>>6)) / 2^53            */
#define NRW_SAMPLER_TREE  2   /* Loihi-faithful: walk the row's binary tree of 8-bit Bernoulli
                                 weights (nrw_tables.tree; rows of <= 512 entries); one Philox
                                 call per step, level l uses bits [7l, 7l+7) of (r0 | r1 << 32) */

#define NRW_COUNTER_STEPS     0   /* walker-steps == uniform draws (fluence_mini_app.py:243)   */
#define NRW_COUNTER_EXITS     1   /* arrivals on absorbing rows                                */
#define NRW_COUNTER_SURVIVORS 2   /* walkers still alive after step nt-1                       */
#define NRW_COUNTER_TOVERFLOW 3   /* tallies dropped because t >= t_cap (count_tn only)        */
#define NRW_COUNTER_WALKERS   4   /* walkers started                                           */
#define NRW_N_COUNTERS        8

/* ---- the path's one collective, fused into the walk (N > 1), and the zeroing of the next tally buffer ---
 * The per-start scores and the counters of a launch live in ONE int64 buffer per rank,
 * own_buf[offset, offset + n) (nrw_walk.score and nrw_walk.counters must point into it).  When the last
 * block of the launch retires it (world > 1) publishes `epoch` to every peer, waits for theirs, sums ITS SLICE
 * (1 / world of the vector) over all ranks over NVLink peer memory, pushes the sums into every peer's
 * own_buf[out_offset, out_offset + n) and signals / awaits "delivered"; the status word
 * own_buf[out_offset + even(n)] is 0 afterwards, 1 if a peer did not arrive within 5 s (the sums are then invalid
 * and no buffer was zeroed -- test it per solve).  It also zeroes own_buf[zero_offset, zero_offset + n)
 * (zero_offset < 0: nothing).  Callers alternate two halves of own_buf by epoch parity (offset = the half of
 * this epoch, zero_offset = the other half): a half is zeroed only after the first wait of the following epoch,
 * which proves that every peer has finished reading it, so no memset is ever launched.  All offsets are even
 * and the regions hold even(n) = (n + 1) & ~1 words (an odd n carries one pad word that stays zero);
 * own_buf holds at least out_offset + even(n) + 2 words.  world == 1: no exchange; the tail only zeroes the next
 * half and the result is own_buf[offset, offset + n) itself.  peer_bufs_dev / peer_flags_dev: DEVICE arrays
 * of `world` pointers -- every rank's buffer (int64) and flag array (uint64[2 * world], zero-initialised) as
 * mapped into this process (symmetric memory / CUDA IPC; nrw/dist.py uses torch's symmetric-memory allocator);
 * `epoch` > 0 and increases by one per launch on every rank.  Replaces the all-reduce a multi-process
 * run(args) of fluence_mini_app.py:397-438 needs once per solve.                                      */
typedef struct nrw_reduce {
    int32_t  rank, world;
    const void* peer_bufs_dev;      /* NULL when world == 1                              */
    const void* peer_flags_dev;     /* NULL when world == 1                              */
    int64_t* own_buf;
    int64_t  offset, zero_offset, n;
    uint64_t epoch;
    int64_t  out_offset;            /* where the sums over ranks land in own_buf (world > 1) */
} nrw_reduce;

typedef struct nrw_walk {
    /* walkers: global walker id g in [0, start_prefix[n_starts]); walker g starts on
     * start_nodes[s] where start_prefix[s] <= g < start_prefix[s+1].  This is the
     * batched form of `initial_walkers=[[node, M]]` (fluence_mini_app.py:229-231, :401);
     * the launch walks ids [walker_lo, walker_hi) so ranks can shard by id.            */
    int32_t  n_starts;
    const int32_t* start_nodes;     /* [n_starts]                                        */
    const int64_t* start_prefix;    /* [n_starts + 1]                                    */
    int64_t  walker_lo, walker_hi;
    /* Philox stream id of walker g: start_id_base[s] + (g - start_prefix[s]), or g itself
     * when start_id_base is NULL.  Lets a rank walk a slice of EVERY start's walkers while
     * each walker keeps the stream it has in the unsharded run.  ids must be < 2^40.    */
    const int64_t* start_id_base;   /* [n_starts] or NULL                                */
    uint64_t seed;                  /* Philox key = 32-bit hash of the seed (splitmix64 finaliser folded to 32 bits, csrc/common.cuh
                                       seed_to_key): 2^32 distinct stream families; counter = (stream id, step), or
                                       (stream id, step pair) on COMPACT tables */
    int32_t  nt;                    /* neuraltime: model steps t = 1 .. nt-1 (:236)      */
    int32_t  sampler;               /* NRW_SAMPLER_*                                     */

    /* per-start score tally (fused fluence_mini_app.py:428-434): for every step
     * 1 <= t <= nt-2 that lands on a live node k, the walker's score grows by
     * llrint(time_weight[t] * class_value[cls(k)])  (time_weight NULL => 1.0, in which
     * case class_value must hold integers in [0, 255]).  score[s] += sum over the start's walkers,
     * score_sq[s] += sum of squared per-walker scores (for the 3-sigma test).           */
    int64_t* score;                 /* [n_starts] or NULL                                */
    double*  score_sq;              /* [n_starts] or NULL                                */
    double   class_value[4];
    const double* time_weight;      /* [nt] or NULL                                      */

    /* occupancy histogram over nodes, all starts together, steps 1 <= t <= nt-2
     * (per-block shared-memory histogram when it fits, merged once per block).  The per-block
     * bins are 32-bit: one launch must put fewer than 2^32 tallies on one node from one block
     * (>= 6e11 walker-steps on a single node per launch on 148 SMs); longer runs are split by
     * walker-id range (walker_lo / walker_hi) -- node_hist accumulates across launches.     */
    unsigned long long* node_hist;  /* [n_rows] or NULL                                  */

    /* time-resolved counts == tempdata (fluence_mini_app.py:227,245): group gidx of a
     * walker is its start index s if count_per_start else 0;
     * count_tn[(gidx * t_cap + t) * n_rows + k] += 1 for 1 <= t < min(nt, t_cap).
     * (t = 0 is the initial condition and is filled by nrw_expand_*).  Bins are 32-bit:
     * fewer than 2^32 walkers per group and launch.                                      */
    uint32_t* count_tn;             /* or NULL                                           */
    int32_t  t_cap;
    int32_t  count_per_start;

    unsigned long long* counters;   /* [NRW_N_COUNTERS], accumulated (caller zeroes)     */

    /* optional: the most likely trajectory of every start, count_path[s * t_cap + t] = node at step t
     * (t = 0: the start node) following each row's most probable successor.  With count_tn in global
     * memory, a walker that is still on its start's path is not tallied step by step: a per-block
     * shared-memory histogram of the steps at which walkers LEAVE the path gives the path bins exactly
     * (count on the path at t = walkers that left later), one shared-memory atomic per walker
     * instead of one contended global atomic per step.  Same counts, bit for bit.  NULL: off.     */
    const int32_t* count_path;

    const nrw_reduce* reduce;       /* HOST pointer, read at launch; NULL: no tail work  */
} nrw_walk;

/* ---- replay walk (K2): bit-exact with the reference under its own uniform stream ----- */
typedef struct nrw_replay {
    int32_t  n_starts;
    const int32_t* start_nodes;     /* [n_starts]                                        */
    const int64_t* start_counts;    /* [n_starts] M per start                            */
    int32_t  nt;
    const double* u_stream;         /* the reference's np.random.uniform() values, in draw order */
    int64_t  u_len;
    /* events: sparse tempdata, (t, node, count) sorted by (start, t, node)              */
    int32_t* ev_t; int32_t* ev_node; int64_t* ev_count;
    int64_t  ev_cap;
    int64_t* ev_offsets;            /* [n_starts + 1]                                    */
    uint32_t* scratch;              /* [3 * (n_rows + 1)] zeroed by the caller before the first call */
    /* state: [0] draws used so far, [1] next start index, [2] next t, [3] status
     * (0 done, 1 stream exhausted -> refill u_stream from position state[0] and call again,
     *  2 event capacity exceeded), [4] events written.  Caller zeroes before first call. */
    int64_t* state;                 /* [8]                                               */
    int64_t  u_base;                /* stream index of u_stream[0]                       */
} nrw_replay;

/* ---- API ------------------------------------------------------------------------------ */
NRW_API int  nrw_abi_version(void);
NRW_API const char* nrw_last_error(void);

/* reads the NRW_* measurement switches from the environment (once) */
NRW_API int  nrw_create(int device, nrw_engine** out);
/* the switches as resolved at nrw_create, "name=value ..." (for logs and bench lines) */
NRW_API int  nrw_engine_config(nrw_engine* e, char* buf, int buf_len);
NRW_API int  nrw_destroy(nrw_engine* e);
/* number of SMs, max opt-in shared memory per block, of the engine's device */
NRW_API int  nrw_device_info(nrw_engine* e, int* n_sms, int* max_smem_bytes);

/* K4: per-row sequential fp64 prefix sums of CSR values
 * (replaces np.cumsum(myQ, axis=1), fluence_mini_app.py:224).                             */
NRW_API int  nrw_build_cum(nrw_engine* e, int32_t n_rows, const int32_t* indptr,
                   const double* vals, double* cum, void* stream);

/* K1: replaces the loops of run_pc, fluence_mini_app.py:236-245 and
 * simple_flux_miniapp.py:189-201, for all start nodes of run() at once (:397-407).        */
NRW_API int  nrw_walk_philox(nrw_engine* e, const nrw_tables* t, const nrw_walk* w, void* stream);
/* which kernel variant nrw_walk_philox would launch: writes a short name and the launch
 * geometry (for bench / profiles); no launch.                                             */
NRW_API int  nrw_walk_philox_plan(nrw_engine* e, const nrw_tables* t, const nrw_walk* w,
                          char* name, int name_len, int* grid, int* block, int* smem_bytes);

/* K2: same loops, consuming the reference's own uniform stream in the reference's order
 * (t, node ascending, slot) -- fluence_mini_app.py:236-243, Q5/Q6 of SURVEY.md.           */
NRW_API int  nrw_walk_replay(nrw_engine* e, const nrw_tables* t, const nrw_replay* r, void* stream);

/* K3: fluence_mini_app.py:422-438 in the reference's (t, k) order, one start per thread:
 * out[s] = (sum over events of (((count*tw[t])*v)*src[k])*dt) / M[s], 1 <= t <= nt-2, k < n_live. */
NRW_API int  nrw_assemble_events(nrw_engine* e, int32_t n_starts, const int64_t* ev_offsets,
                         const int32_t* ev_t, const int32_t* ev_node, const int64_t* ev_count,
                         const double* time_weight, const double* src, double v, double dt,
                         const int64_t* start_counts, int32_t nt, int32_t n_live,
                         double* out, void* stream);

/* K5: sparse events / dense uint32 counts -> the dense float64 [n_rows, nt] array that
 * run_pc returns (fluence_mini_app.py:227,247).  `out` is fully written.                  */
NRW_API int  nrw_expand_events(nrw_engine* e, int64_t n_events, const int32_t* ev_t,
                       const int32_t* ev_node, const int64_t* ev_count,
                       int32_t n_rows, int32_t nt, double* out, void* stream);
NRW_API int  nrw_expand_counts(nrw_engine* e, const uint32_t* count_tn, int32_t t_cap, int32_t n_rows,
                       int32_t nt, int32_t start_node, int64_t start_count,
                       double* out, void* stream);

/* End-to-end convenience with HOST buffers (what `bench.py` times as e2e): copies the
 * compact alias tables and the start list to the device, runs K1 in score mode, copies
 * score / counters back and synchronises `stream`.  `reduce` NULL: single GPU, the engine's own
 * tally buffer.  With `reduce` (N > 1; n must be n_starts + NRW_N_COUNTERS): the walk tallies into
 * reduce->own_buf + offset, the sums over ranks come back from reduce->own_buf + out_offset, and the call
 * fails with NRW_ECUDA if a peer timed out.  start_id_base_host may be NULL.  rowinfo_str_host / str_thresh / str_n:
 * the structured form of the same tables (nrw_tables.rowinfo_str, .str_thresh, .str_n; nrw/tables.py::structure_rows) or NULL / 0 / 0;
 * when given and the planner takes the structured walk, ONLY it is copied (n_rows * 8 bytes instead of
 * descriptors + alias patterns) -- same results.                                              */
NRW_API int  nrw_fluence_scores_host(nrw_engine* e,
                             int32_t n_rows, int32_t n_live, int32_t n_alias,
                             const void* rowinfo_host, const void* alias_host,
                             const void* rowinfo_str_host, uint32_t str_thresh, uint32_t str_n,
                             int32_t n_starts, const int32_t* start_nodes_host,
                             const int64_t* start_prefix_host, const int64_t* start_id_base_host,
                             int64_t walker_lo, int64_t walker_hi, uint64_t seed, int32_t nt,
                             const double class_value[4],
                             int64_t* score_host, unsigned long long* counters_host,
                             const nrw_reduce* reduce, void* stream);

/* nrw_fluence_scores_host WITHOUT any allocation, for callers that cannot call cudaMalloc (code under CUDA
 * graph capture, tight serving loops, harnesses that forbid it).  Same arguments and results as
 * nrw_fluence_scores_host with reduce == NULL (one GPU), bit for bit; but every device-side buffer the call needs
 * -- the staged copies of the tables and of the start list, the tally (scores + counters) and the launch's ticket
 * slot -- lives in ONE file-scope `static __device__` array of NRW_STATIC_SCRATCH_BYTES (1 MiB) inside libnrw.so
 * instead of cudaMalloc'ed buffers that grow on demand.  Every GPU that loads the library holds that array until
 * the process exits.  The call (and nothing it calls) uses cudaMalloc, cudaFree or any allocating routine; the
 * array's device address is resolved once per device with cudaGetSymbolAddress on the first call for that device.
 * Limit: 256 (ticket slot) + the 256-byte-rounded sizes of (n_starts + 8) * 8 [tally], n_rows * 8 [row
 * descriptors], n_alias * 8 [alias entries], n_starts * 4, (n_starts + 1) * 8 and (with start_id_base_host)
 * n_starts * 8 bytes must fit NRW_STATIC_SCRATCH_BYTES -- the 10 001-row C5 mesh with all 10 000 starts needs
 * 0.42 MiB; larger problems are refused with NRW_ENOSPC and a message that names both numbers.
 * ALL calls of a process on one device share the array, which keeps the previous call's contents: the call
 * clears what must start at zero (tally, ticket slot) itself, and two calls that overlap on the GPU (two streams,
 * two threads) would race on it -- the library serialises callers behind a mutex and returns only after `stream`
 * has been synchronised, so calls through this entry never overlap; do not use the array in any other way.
 * No counterpart in the reference. */
#define NRW_STATIC_SCRATCH_BYTES (1 << 20)
NRW_API int  nrw_fluence_scores_host_static(nrw_engine* e,
                             int32_t n_rows, int32_t n_live, int32_t n_alias,
                             const void* rowinfo_host, const void* alias_host,
                             const void* rowinfo_str_host, uint32_t str_thresh, uint32_t str_n,
                             int32_t n_starts, const int32_t* start_nodes_host,
                             const int64_t* start_prefix_host, const int64_t* start_id_base_host,
                             int64_t walker_lo, int64_t walker_hi, uint64_t seed, int32_t nt,
                             const double class_value[4],
                             int64_t* score_host, unsigned long long* counters_host,
                             void* stream);

/* The path's one collective over peer memory (no NCCL) as a stand-alone launch: afterwards
 * peer_bufs[q][out_offset + i] = sum over ranks r of peer_bufs[r][offset + i] on every rank q (i < n).  Same
 * protocol, buffer layout and alternation rule as nrw_reduce above (the walk runs this exchange in its own tail;
 * this entry serves tallies no single walk launch produced).  *err_flag (device int) is set to 1 if a peer does
 * not arrive within 5 s (%globaltimer); the output region is then incomplete.  Replaces the all-reduce a
 * multi-process run(args) of fluence_mini_app.py:397-438 needs once per solve. */
NRW_API int  nrw_p2p_allreduce_i64(nrw_engine* e, const void* peer_bufs_dev, const void* peer_flags_dev,
                                   int64_t offset, int64_t zero_offset, int64_t out_offset, int64_t n,
                                   int32_t rank, int32_t world, uint64_t epoch, int32_t* err_flag, void* stream);

/* Roofline denominators measured on the engine's device at the walk's own launch geometry (one
 * 1024-thread CTA per SM): kind 0 = random atomicAdd on a per-CTA shared-memory histogram of n bins,
 * 1 = dependent random 8-byte shared-memory loads from n entries, 2 = dependent random 8-byte loads
 * from an n-entry global table (device pointer `table`, L2-resident for n*8 B << 126 MB), 3 = random
 * 64-bit atomicAdd on a global histogram `hist` of n bins.  Every thread performs `iters`
 * operations; time the launch with events on `stream`.  `scratch` = 8 device bytes.               */
NRW_API int  nrw_microbench(nrw_engine* e, int32_t kind, int32_t n, int32_t iters, const void* table,
                    unsigned long long* hist, unsigned long long* scratch, void* stream);

/* The walk's work distribution, exposed for tests (no device needed): warps draw TICKETS from one counter
 * and ticket t of phase j is the id range [start[j] + (t - ticket[j]) * range_of_phase[j], ... ) clipped at
 * start[j+1].  Phase 0 hands out ranges of `range` ids; the last `tail_ranges` such ranges per warp of the
 * n ids go out in phases of halving range length down to `range_tail`.  The phases tile [0, n) exactly
 * once.  Arrays: start, ticket hold NRW_MAX_PHASES + 1 entries, range_of_phase NRW_MAX_PHASES.
 * Replaces the reference's serial `for i in range(meshSize)` (fluence_mini_app.py:397) as the unit of
 * work hand-out.                                                                                    */
#define NRW_MAX_PHASES 8
NRW_API int  nrw_guided_phases(int64_t n, int64_t range, int64_t warps, int64_t range_tail,
                               int64_t tail_ranges, int64_t* start, int64_t* ticket,
                               int64_t* range_of_phase, int32_t* n_phase);

/* Host only (no device needed): the text of a result file, np.savetxt(path, a, delimiter=',') of
 * fluence_mini_app.py:446 / simple_flux_miniapp.py:358 -- '%.18e' per value, ',' between the values of a row, a
 * newline after each row; the same bytes.  a: [nrows][ncols] row-major; out needs nrows * ncols * 28 bytes; *len
 * receives the length written.  The rows are converted by up to n_threads host threads (a correctly rounded 19-digit
 * conversion costs ~0.3 us per value: on the 10 000-cell mesh the result file was 45 % of a warm run(args)).       */
NRW_API int  nrw_format_csv_e18(const double* a, int64_t nrows, int64_t ncols, char* out, int64_t cap,
                                int64_t* len, int32_t n_threads);

#ifdef __cplusplus
}
#endif
#endif /* NRW_H_ */
