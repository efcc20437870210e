// nrw_capi.cu -- the C ABI of libnrw.so (declared in include/nrw.h).  Plain pointers and sizes;
// no torch types; errors are returned, never thrown.
#include "../../../include/nrw.h"

#include <clocale>
#include <cstdarg>
#include <cstdio>
#include <algorithm>
#include <atomic>
#include <cstring>
#include <mutex>
#include <thread>
#include <unordered_map>
#include <vector>

#include "common.cuh"
#include "microbench.cuh"
#include "p2p_reduce.cuh"
#include "post.cuh"
#include "walk_philox.cuh"
#include "walk_replay.cuh"
#include "walk_score.cuh"

#include <cstdlib>

using namespace nrw;

static thread_local char g_err[512] = "";

static int fail(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}
#define CUDA_TRY(expr)                                                                        \
    do {                                                                                      \
        cudaError_t _e = (expr);                                                              \
        if (_e != cudaSuccess)                                                                \
            return fail(NRW_ECUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e),    \
                        __FILE__, __LINE__);                                                  \
    } while (0)

// The measurement switches, read from the environment ONCE in nrw_create.
struct nrw_config {
    int k1_general;        // NRW_K1=general: score walks take the general kernel
    int det;               // NRW_DET=0: draw on rows with a single absorbing successor too
    int tab;               // NRW_TAB: 0 auto, 1 hyb (descriptors L2-resident), 2 fac (factored descriptors)
    int cps;               // NRW_CPS=2: two CTAs of 768 threads per SM when the tables leave room
    long long ranges_per_warp, range_tail, tail_ranges;
    int cnt_match;         // NRW_CNT_MATCH: -1 auto, 0 plain, 1 match_any aggregation
    int cnt_cache;         // NRW_CNT_CACHE=0: no CTA-level cache of hot bins
    int cnt_path;          // NRW_CNT_PATH=0: ignore nrw_walk.count_path
    int str;               // NRW_STR=0: ignore nrw_tables.rowinfo_str (alias look-ups in the score walk)
    int block;             // NRW_BLOCK: threads per CTA of the trimmed walk (1024; fewer = an occupancy experiment)
};

struct nrw_engine {
    int device;
    int n_sms;
    int max_smem;
    nrw_config cfg;
    // ring of ticket slots {work_next, done}: one per launch, handed back by the launch's last CTA (walk_tail.cuh)
    unsigned long long* slots;
    std::atomic<unsigned long long> next_slot;
    std::mutex mu_host;                                       // the *_host staging buffers (calls serialise)
    // grow-only staging buffers of nrw_fluence_scores_host
    void* d_rowinfo; size_t cap_rowinfo;
    void* d_alias; size_t cap_alias;
    void* d_starts; size_t cap_starts;
    void* d_prefix; size_t cap_prefix;
    void* d_idbase; size_t cap_idbase;
    void* d_tally; size_t cap_tally;                          // scores, then NRW_N_COUNTERS counters
    long long* h_tail;                                        // pinned: counters (+ exchange status) of a host-buffer solve
};

static int env_int(const char* name, int dflt) {
    const char* v = getenv(name);
    return v && *v ? atoi(v) : dflt;
}
static nrw_config config_from_env() {
    nrw_config c;
    const char* k1 = getenv("NRW_K1");
    c.k1_general = k1 && strcmp(k1, "general") == 0;
    c.det = env_int("NRW_DET", 1) != 0;
    const char* tab = getenv("NRW_TAB");
    c.tab = tab && strcmp(tab, "hyb") == 0 ? 1 : (tab && strcmp(tab, "fac") == 0 ? 2 : 0);
    c.cps = env_int("NRW_CPS", 1) == 2 ? 2 : 1;
    c.ranges_per_warp = env_int("NRW_RANGES_PER_WARP", 6);
    if (c.ranges_per_warp < 1) c.ranges_per_warp = 6;
    c.range_tail = env_int("NRW_RANGE_TAIL", 1024);
    c.tail_ranges = env_int("NRW_TAIL_RANGES", 2);
    c.cnt_match = getenv("NRW_CNT_MATCH") ? (env_int("NRW_CNT_MATCH", 1) != 0) : -1;
    c.cnt_cache = env_int("NRW_CNT_CACHE", 1) != 0;
    c.cnt_path = env_int("NRW_CNT_PATH", 1) != 0;
    c.str = env_int("NRW_STR", 1) != 0;
    c.block = env_int("NRW_BLOCK", WALK_THREADS);
    if (c.block < 32 || c.block > WALK_THREADS || c.block % 32) c.block = WALK_THREADS;
    return c;
}

extern "C" int nrw_abi_version(void) { return NRW_ABI_VERSION; }
extern "C" const char* nrw_last_error(void) { return g_err; }

extern "C" int nrw_create(int device, nrw_engine** out) {
    if (!out) return fail(NRW_EINVAL, "nrw_create: out is NULL");
    *out = nullptr;
    int n = 0;
    CUDA_TRY(cudaGetDeviceCount(&n));
    if (device < 0 || device >= n) return fail(NRW_EINVAL, "nrw_create: device %d of %d", device, n);
    CUDA_TRY(cudaSetDevice(device));
    cudaDeviceProp prop;
    CUDA_TRY(cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10)
        return fail(NRW_EINVAL, "nrw_create: device %d is sm_%d%d; libnrw is built for sm_100a only",
                    device, prop.major, prop.minor);
    nrw_engine* e = new nrw_engine();
    e->device = device;
    e->n_sms = prop.multiProcessorCount;
    e->max_smem = (int)prop.sharedMemPerBlockOptin;
    e->cfg = config_from_env();
    e->slots = nullptr;
    e->next_slot = 0;
    e->d_rowinfo = e->d_alias = e->d_starts = e->d_prefix = e->d_idbase = e->d_tally = nullptr;
    e->cap_rowinfo = e->cap_alias = e->cap_starts = e->cap_prefix = e->cap_idbase = e->cap_tally = 0;
    e->h_tail = nullptr;
    const size_t slot_bytes = (size_t)NRW_TICKET_SLOTS * TICKET_WORDS * sizeof(unsigned long long);
    cudaError_t err = cudaMalloc(&e->slots, slot_bytes);
    if (err == cudaSuccess) err = cudaMemset(e->slots, 0, slot_bytes);
    if (err == cudaSuccess) err = cudaHostAlloc((void**)&e->h_tail, (NRW_N_COUNTERS + 2) * sizeof(long long), cudaHostAllocDefault);
    if (err != cudaSuccess) {
        nrw_destroy(e);                                   // frees whatever was allocated
        return fail(NRW_ECUDA, "nrw_create: cudaMalloc failed: %s", cudaGetErrorString(err));
    }
    *out = e;
    return NRW_OK;
}

extern "C" int nrw_destroy(nrw_engine* e) {
    if (!e) return NRW_OK;
    cudaSetDevice(e->device);
    cudaFree(e->slots);
    if (e->h_tail) cudaFreeHost(e->h_tail);
    cudaFree(e->d_rowinfo); cudaFree(e->d_alias); cudaFree(e->d_starts);
    cudaFree(e->d_prefix); cudaFree(e->d_idbase); cudaFree(e->d_tally);
    delete e;
    return NRW_OK;
}

extern "C" int nrw_engine_config(nrw_engine* e, char* buf, int buf_len) {
    if (!e || !buf || buf_len <= 0) return fail(NRW_EINVAL, "nrw_engine_config: bad argument");
    const nrw_config& c = e->cfg;
    snprintf(buf, (size_t)buf_len,
             "k1=%s det=%d tab=%s cps=%d ranges_per_warp=%lld range_tail=%lld tail_ranges=%lld cnt_match=%d "
             "cnt_cache=%d cnt_path=%d str=%d block=%d", c.k1_general ? "general" : "auto", c.det,
             c.tab == 1 ? "hyb" : (c.tab == 2 ? "fac" : "auto"), c.cps, c.ranges_per_warp, c.range_tail,
             c.tail_ranges, c.cnt_match, c.cnt_cache, c.cnt_path, c.str, c.block);
    return NRW_OK;
}

extern "C" int nrw_device_info(nrw_engine* e, int* n_sms, int* max_smem_bytes) {
    if (!e) return fail(NRW_EINVAL, "nrw_device_info: engine is NULL");
    if (n_sms) *n_sms = e->n_sms;
    if (max_smem_bytes) *max_smem_bytes = e->max_smem;
    return NRW_OK;
}

// ---- K4 ---------------------------------------------------------------------------------------
extern "C" int nrw_build_cum(nrw_engine* e, int32_t n_rows, const int32_t* indptr,
                             const double* vals, double* cum, void* stream) {
    if (!e || !indptr || !vals || !cum || n_rows < 0) return fail(NRW_EINVAL, "nrw_build_cum: bad argument");
    if (n_rows == 0) return NRW_OK;
    CUDA_TRY(cudaSetDevice(e->device));
    build_cum_kernel<<<(n_rows + 255) / 256, 256, 0, (cudaStream_t)stream>>>(n_rows, indptr, vals, cum);
    CUDA_TRY(cudaGetLastError());
    return NRW_OK;
}

// ---- K3 ---------------------------------------------------------------------------------------
extern "C" int nrw_assemble_events(nrw_engine* e, int32_t n_starts, const int64_t* ev_offsets,
                                   const int32_t* ev_t, const int32_t* ev_node,
                                   const int64_t* ev_count, const double* time_weight,
                                   const double* src, double v, double dt,
                                   const int64_t* start_counts, int32_t nt, int32_t n_live,
                                   double* out, void* stream) {
    if (!e || !ev_offsets || !time_weight || !src || !start_counts || !out || n_starts < 0)
        return fail(NRW_EINVAL, "nrw_assemble_events: bad argument");
    if (n_starts == 0) return NRW_OK;
    CUDA_TRY(cudaSetDevice(e->device));
    assemble_events_kernel<<<(n_starts + 127) / 128, 128, 0, (cudaStream_t)stream>>>(
        n_starts, (const long long*)ev_offsets, ev_t, ev_node, (const long long*)ev_count,
        time_weight, src, v, dt, (const long long*)start_counts, nt, n_live, out);
    CUDA_TRY(cudaGetLastError());
    return NRW_OK;
}

// ---- K5 ---------------------------------------------------------------------------------------
extern "C" int nrw_expand_events(nrw_engine* e, int64_t n_events, const int32_t* ev_t,
                                 const int32_t* ev_node, const int64_t* ev_count, int32_t n_rows,
                                 int32_t nt, double* out, void* stream) {
    if (!e || !out || n_rows <= 0 || nt <= 0 || n_events < 0)
        return fail(NRW_EINVAL, "nrw_expand_events: bad argument");
    CUDA_TRY(cudaSetDevice(e->device));
    CUDA_TRY(cudaMemsetAsync(out, 0, (size_t)n_rows * nt * sizeof(double), (cudaStream_t)stream));
    if (n_events > 0) {
        const long long blocks = (n_events + 255) / 256;
        expand_events_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(
            n_events, ev_t, ev_node, (const long long*)ev_count, n_rows, nt, out);
        CUDA_TRY(cudaGetLastError());
    }
    return NRW_OK;
}

extern "C" int nrw_expand_counts(nrw_engine* e, const uint32_t* count_tn, int32_t t_cap,
                                 int32_t n_rows, int32_t nt, int32_t start_node,
                                 int64_t start_count, double* out, void* stream) {
    if (!e || !count_tn || !out || n_rows <= 0 || nt <= 0 || t_cap <= 0)
        return fail(NRW_EINVAL, "nrw_expand_counts: bad argument");
    CUDA_TRY(cudaSetDevice(e->device));
    dim3 grid((nt + 31) / 32, (n_rows + 31) / 32), block(32, 8);
    if (grid.y > 65535) return fail(NRW_EINVAL, "nrw_expand_counts: n_rows too large for a dense tempdata");
    expand_counts_kernel<<<grid, block, 0, (cudaStream_t)stream>>>(count_tn, t_cap, n_rows, nt,
                                                                   start_node, start_count, out);
    CUDA_TRY(cudaGetLastError());
    return NRW_OK;
}

// ---- K1: plan + launch ------------------------------------------------------------------------
struct WalkPlan {
    int tab, tl, grid, smem, block;
    WalkParams p;
    const void* fn;
    const char* name;
};

#define NRW_WALK_VARIANTS(X)                                                                   \
    X(TAB_SMEM, 1) X(TAB_SMEM, 3) X(TAB_SMEM, 4) X(TAB_SMEM, 5) X(TAB_SMEM, 7) X(TAB_SMEM, 8)  \
    X(TAB_SMEM, 9) X(TAB_SMEM, 11) X(TAB_SMEM, 16) X(TAB_SMEM, 32)                             \
    X(TAB_HYB, 1) X(TAB_HYB, 3) X(TAB_HYB, 4) X(TAB_HYB, 5) X(TAB_HYB, 8) X(TAB_HYB, 9)        \
    X(TAB_HYB, 32)                                                                             \
    X(TAB_FAC, 1) X(TAB_FAC, 3) X(TAB_FAC, 4) X(TAB_FAC, 5) X(TAB_FAC, 8) X(TAB_FAC, 9)        \
    X(TAB_FAC, 32)                                                                             \
    X(TAB_GLB, 1) X(TAB_GLB, 3) X(TAB_GLB, 8) X(TAB_GLB, 9) X(TAB_GLB, 16) X(TAB_GLB, 32)      \
    X(TAB_CDF, 1) X(TAB_CDF, 3) X(TAB_CDF, 8) X(TAB_CDF, 9) X(TAB_CDF, 16) X(TAB_CDF, 32)     \
    X(TAB_TREE, 1) X(TAB_TREE, 3) X(TAB_TREE, 8) X(TAB_TREE, 9) X(TAB_TREE, 16) X(TAB_TREE, 32)

static int align16(int x) { return (x + 15) & ~15; }
// table sizes in bytes, clamped so that tables far beyond any shared memory cannot wrap an int
static int table_bytes(long long n, int elem) {
    const long long b = (n * elem + 15) & ~15ll;
    return b > (1ll << 29) ? (1 << 29) : (int)b;
}

// Guided range sizes: phase 0 = ranges of `range` ids; the last `tail_ranges` big ranges per warp of the id
// space go out in phases of ranges half as long each (two per warp and phase) down to `range_tail` ids (a
// multiple of 32, >= 32).  Phase j covers ids [start[j], start[j+1]) with tickets ticket[j] .. ticket[j+1]-1;
// the phases tile [0, n) exactly once.  Returns the number of phases (<= NRW_MAX_PHASES).
static int guided_phases(long long n, long long range, long long warps, long long range_tail,
                         long long tail_ranges, long long* start, long long* ticket, long long* rng) {
    long long rt = (range_tail / 32) * 32;
    if (rt < 32) rt = 32;
    if (rt > range) rt = range;
    long long tail_ids = tail_ranges <= 0 ? 0 : tail_ranges * range * warps;
    if (tail_ids > n || tail_ids < 0) tail_ids = n;
    int np = 0;
    long long at = 0, tk = 0, r = range;
    while (true) {
        const bool last = np == NRW_MAX_PHASES - 1 || r <= rt || tail_ids == 0;
        const long long stop = last ? n : n - tail_ids;         // this phase covers [at, stop)
        if (stop > at || last) {
            start[np] = at; ticket[np] = tk; rng[np] = r;
            tk += (stop - at + r - 1) / r;
            at = stop;
            ++np;
        }
        if (last) break;
        r = std::max(rt, ((r / 2) / 32) * 32);
        tail_ids /= 2;
    }
    start[np] = n; ticket[np] = tk;
    return np;
}

extern "C" int nrw_guided_phases(int64_t n, int64_t range, int64_t warps, int64_t range_tail, int64_t tail_ranges,
                                 int64_t* start, int64_t* ticket, int64_t* range_of_phase, int32_t* n_phase) {
    if (n < 0 || range < 1 || warps < 1 || !start || !ticket || !range_of_phase || !n_phase)
        return fail(NRW_EINVAL, "nrw_guided_phases: bad argument");
    long long st[NRW_MAX_PHASES + 1], tk[NRW_MAX_PHASES + 1], rg[NRW_MAX_PHASES];
    const int np = guided_phases(n, range, warps, range_tail, tail_ranges, st, tk, rg);
    for (int i = 0; i <= np; ++i) { start[i] = st[i]; ticket[i] = tk[i]; }
    for (int i = 0; i < np; ++i) range_of_phase[i] = rg[i];
    *n_phase = np;
    return NRW_OK;
}

static int make_plan(nrw_engine* e, const nrw_tables* t, const nrw_walk* w, WalkPlan* plan, bool allow_str = true) {
    if (!e || !t || !w) return fail(NRW_EINVAL, "nrw_walk_philox: NULL argument");
    if (t->n_rows <= 0 || t->n_live < 0 || t->n_live > t->n_rows)
        return fail(NRW_EINVAL, "nrw_walk_philox: n_rows=%d n_live=%d", t->n_rows, t->n_live);
    if (w->n_starts <= 0 || !w->start_nodes || !w->start_prefix)
        return fail(NRW_EINVAL, "nrw_walk_philox: empty start list");
    if (w->walker_lo < 0 || w->walker_hi < w->walker_lo || w->walker_hi >= (1ll << 40))
        return fail(NRW_EINVAL, "nrw_walk_philox: walker id range [%lld, %lld) (ids must be < 2^40)",
                    (long long)w->walker_lo, (long long)w->walker_hi);
    if (w->nt < 1 || w->nt >= (1 << 24))
        return fail(NRW_EINVAL, "nrw_walk_philox: nt=%d outside [1, 2^24)", w->nt);
    if (!w->score && !w->node_hist && !w->count_tn)
        return fail(NRW_EINVAL, "nrw_walk_philox: no tally requested");
    if (w->time_weight && !w->score)
        return fail(NRW_EINVAL, "nrw_walk_philox: time_weight without score");

    WalkParams& p = plan->p;
    memset(&p, 0, sizeof(p));
    p.n_rows = t->n_rows; p.n_live = t->n_live; p.n_alias = t->n_alias;
    p.indptr = t->indptr; p.cols = t->cols; p.cum = t->cum; p.node_class = t->node_class;
    p.rowinfo = t->rowinfo; p.alias = t->alias;
    p.tree = t->tree; p.tree_ptr = t->tree_ptr;
    p.n_starts = w->n_starts; p.start_nodes = w->start_nodes;
    p.start_prefix = (const long long*)w->start_prefix;
    p.start_id_base = (const long long*)w->start_id_base;
    p.lo = w->walker_lo; p.hi = w->walker_hi;
    p.key = seed_to_key(w->seed); p.nt = w->nt;
    for (int r = 0; r < 10; ++r) p.rk[r] = p.key + (uint32_t)r * PHILOX_W32;
    p.t_last = w->nt - 1;
    p.t_fast = w->nt > 4 ? (uint32_t)(w->nt - 4) : 0u;
    p.score = (long long*)w->score; p.score_sq = w->score_sq; p.tw = w->time_weight;
    p.cwpack = 0;
    for (int i = 0; i < 4; ++i) {
        p.cv[i] = w->class_value[i];
        if (!w->time_weight && w->score) {
            const long long q = (long long)w->class_value[i];
            if ((double)q != w->class_value[i] || q < 0 || q > 255)
                return fail(NRW_EINVAL, "nrw_walk_philox: class_value[%d]=%g must be an integer in "
                            "[0, 255] when time_weight is NULL", i, w->class_value[i]);
            p.cwpack |= (uint32_t)q << (8 * i);
        }
    }
    p.node_hist = w->node_hist; p.count_tn = w->count_tn;
    p.t_cap = w->t_cap; p.per_start = w->count_per_start;
    // warp-aggregated (match_any) or plain shared-memory atomics for the time-resolved counts:
    // aggregation pays when walks are short and lanes share bins; NRW_CNT_MATCH=0/1 overrides
    p.cnt_match = e->cfg.cnt_match >= 0 ? e->cfg.cnt_match : (w->t_cap <= 256);
    p.counters = w->counters; p.work_next = nullptr;          // the ticket slot is assigned at launch
    p.tail.zero_off = -1; p.tail.world = 1;
    p.tail.timeout_ns = 5000000000ull;                        // 5 s without a peer: give up (status word)
    if (const nrw_reduce* r = w->reduce) {
        if (r->world < 1 || r->world > 64 || r->rank < 0 || r->rank >= r->world || r->n <= 0 || r->offset < 0 ||
            !r->own_buf || r->epoch == 0)
            return fail(NRW_EINVAL, "nrw_walk_philox: bad reduce descriptor");
        if (r->world > 1 && (!r->peer_bufs_dev || !r->peer_flags_dev || r->out_offset < 0))
            return fail(NRW_EINVAL, "nrw_walk_philox: reduce over %d ranks needs peer buffers, flags and out_offset", r->world);
        if (((r->offset | r->out_offset | (r->zero_offset < 0 ? 0 : r->zero_offset)) & 1) != 0)
            return fail(NRW_EINVAL, "nrw_walk_philox: reduce offsets must be even (16-byte aligned halves)");
        const long long* lo = (const long long*)r->own_buf + r->offset;
        const long long* hi = lo + r->n;
        if (w->score && ((const long long*)w->score < lo || (const long long*)w->score + w->n_starts > hi))
            return fail(NRW_EINVAL, "nrw_walk_philox: score is not inside the reduce buffer");
        if (w->counters && ((const long long*)w->counters < lo || (const long long*)w->counters + NRW_N_COUNTERS > hi))
            return fail(NRW_EINVAL, "nrw_walk_philox: counters are not inside the reduce buffer");
        p.tail.world = r->world; p.tail.rank = r->rank;
        p.tail.bufs = (long long* const*)r->peer_bufs_dev;
        p.tail.flags = (unsigned long long* const*)r->peer_flags_dev;
        p.tail.own = (long long*)r->own_buf;
        p.tail.off = r->offset; p.tail.zero_off = r->zero_offset; p.tail.out_off = r->out_offset; p.tail.n = r->n;
        p.tail.epoch = r->epoch;
        p.tail.status = r->world > 1 ? (long long*)r->own_buf + r->out_offset + ((r->n + 1) & ~1ll) : nullptr;
    }
    if (w->count_tn && w->t_cap < 1) return fail(NRW_EINVAL, "nrw_walk_philox: t_cap=%d", w->t_cap);
    if (w->count_tn && w->walker_hi - w->walker_lo >= (1ll << 32))
        return fail(NRW_EINVAL, "nrw_walk_philox: count_tn bins are 32-bit; walk fewer than 2^32 ids per launch");

    // Structured draw (nrw_tables.rowinfo_str): taken by the trimmed score walk when the tables carry it.  Decided
    // here because it changes what is staged (no alias patterns); if the launch then turns out not to qualify for
    // the trimmed kernel (histogram too large for shared memory ...) the plan is redone without it.
    bool k1s_horizon_ok;
    {   // the trimmed kernel keeps 32-bit running sums per lane: walks so long that even a 32-id range could wrap
        // them stay with the general kernel
        long long mw = 3;
        if (p.cwpack != 0x03020100u) for (int i = 0; i < 4; ++i) mw = std::max(mw, (long long)((p.cwpack >> (8 * i)) & 0xffu));
        k1s_horizon_ok = 32ll * (long long)w->nt * mw < (1ll << 31);
    }
    const bool str = allow_str && e->cfg.str && w->sampler == NRW_SAMPLER_ALIAS && t->alias_layout == NRW_ALIAS_COMPACT &&
                     t->rowinfo_str && t->str_n >= 1u && (w->score || w->node_hist) && !w->count_tn && !w->score_sq && !w->time_weight &&
                     w->nt >= 2 && !e->cfg.k1_general && e->cfg.cps != 2 && k1s_horizon_ok;

    // table location
    int smem = 16, tab;
    if (w->sampler == NRW_SAMPLER_CDF) {
        if (!t->indptr || !t->cols || !t->cum)
            return fail(NRW_EINVAL, "nrw_walk_philox: cdf sampler needs indptr/cols/cum");
        tab = TAB_CDF;
    } else if (w->sampler == NRW_SAMPLER_TREE) {
        if (!t->indptr || !t->cols || !t->tree || !t->tree_ptr)
            return fail(NRW_EINVAL, "nrw_walk_philox: tree sampler needs indptr/cols/tree/tree_ptr");
        tab = TAB_TREE;
    } else if (w->sampler == NRW_SAMPLER_ALIAS) {
        if (!t->rowinfo || !t->alias || t->n_alias <= 0)
            return fail(NRW_EINVAL, "nrw_walk_philox: alias sampler needs rowinfo/alias tables");
        if (t->alias_layout == NRW_ALIAS_WIDE) {
            tab = TAB_GLB;
        } else if (t->alias_layout == NRW_ALIAS_COMPACT) {
            if (t->n_alias > (1 << 15))
                return fail(NRW_EINVAL, "nrw_walk_philox: compact layout holds <= 2^15 alias entries");
            const int ab = str ? 0 : table_bytes(t->n_alias, 8), rb = table_bytes(t->n_rows, 8);
            // factored row descriptors (fluence-structured meshes): taken when the descriptor table
            // does not fit shared memory but the three small tables do; NRW_TAB=hyb / fac overrides
            const bool have_fac = (str ? (t->fac_dir_str && t->fac_tail_str) : (t->fac_dir && t->fac_tail)) && t->fac_cls && t->fac_period > 1 &&
                                  t->fac_cells > 0 && t->fac_n_tail == t->n_rows - t->n_live &&
                                  (long long)t->fac_period * t->fac_cells == t->n_live;
            const long long fbl = have_fac ? (long long)table_bytes(t->fac_period, 8) + table_bytes(t->fac_n_tail, 8) +
                                             table_bytes((t->n_live + 15) / 16, 4) : 0;
            const int fb = fbl > (1ll << 29) ? (1 << 29) : (int)fbl;
            const bool force_fac = e->cfg.tab == 2, no_fac = e->cfg.tab == 1;
            if (force_fac && !(have_fac && 16 + ab + fb <= e->max_smem))
                return fail(NRW_EINVAL, "nrw_walk_philox: NRW_TAB=fac but the tables carry no factored rows that fit");
            if (force_fac) tab = TAB_FAC;
            else if (16 + ab + rb <= e->max_smem && !no_fac) tab = TAB_SMEM;
            else if (have_fac && !no_fac && 16 + ab + fb <= e->max_smem) tab = TAB_FAC;
            else if (16 + ab <= e->max_smem) tab = TAB_HYB;
            else return fail(NRW_ENOSPC, "nrw_walk_philox: %d compact alias entries exceed shared "
                             "memory (%d B); use the wide layout", t->n_alias, e->max_smem);
            p.sm_alias = smem; smem += ab;
            if (tab == TAB_SMEM) { p.sm_rowinfo = smem; smem += rb; }
            if (tab == TAB_FAC) {
                p.fac_dir = t->fac_dir; p.fac_tail = t->fac_tail; p.fac_cls = t->fac_cls;
                p.fac_period = t->fac_period; p.fac_cells = t->fac_cells; p.fac_n_tail = t->fac_n_tail;
                p.fac_shift = t->fac_shift; p.fac_magic = t->fac_magic;
                p.fac_exit_w0 = t->fac_exit_w0; p.fac_exit_base = t->fac_exit_base;
                if (t->fac_shift < 0 || t->fac_shift > 31) return fail(NRW_EINVAL, "nrw_walk_philox: fac_shift=%d", t->fac_shift);
                p.sm_fdir = smem; smem += align16(t->fac_period * 8);
                p.sm_ftail = smem; smem += align16(t->fac_n_tail * 8);
                p.sm_fcls = smem; smem += align16(((t->n_live + 15) / 16) * 4);
            }
            if (str) {           // the structured descriptors in place of the alias ones; no alias patterns are staged
                p.rowinfo = t->rowinfo_str; p.n_alias = 0; p.str_thr = t->str_thresh; p.str_n = t->str_n;
                if (tab == TAB_FAC) { p.fac_dir = t->fac_dir_str; p.fac_tail = t->fac_tail_str; p.fac_exit_w0 = t->fac_exit_w0_str; }
            }
        } else {
            return fail(NRW_EINVAL, "nrw_walk_philox: alias_layout=%d", t->alias_layout);
        }
    } else {
        return fail(NRW_EINVAL, "nrw_walk_philox: sampler=%d", w->sampler);
    }

    // tallies
    int tl = 0;
    if (w->score) tl |= TL_SCORE | (w->time_weight ? TL_TW : 0);
    if (w->node_hist) {
        const int hb = table_bytes(t->n_rows, 4);
        if (tab_compact(tab) && smem + hb <= e->max_smem) {
            tl |= TL_HIST_S; p.sm_hist = smem; smem += hb;
        } else {
            tl |= TL_HIST_G;
        }
    }
    if (w->count_tn) {
        const long long groups = w->count_per_start ? w->n_starts : 1;
        const long long bins = groups * (long long)w->t_cap * t->n_rows;
        if (bins <= 0 || bins >= (1ll << 40)) return fail(NRW_EINVAL, "nrw_walk_philox: %lld count bins", bins);
        if (tab != TAB_HYB && tab != TAB_FAC && bins * 4 + smem <= e->max_smem && bins <= 32768) {
            tl |= TL_CNT_S; p.sm_cnt = smem; smem += align16((int)bins * 4); p.cnt_bins = (int)bins;
        } else {
            tl |= TL_CNT_G;
            // the likely path of every start, when the caller supplies it and its two small tables fit shared memory
            // (NRW_CNT_PATH=0 ignores it): walkers on the path are tallied through a leave-time histogram
            const long long n_walkers = w->walker_hi - w->walker_lo;
            const long long path_cells = (long long)w->n_starts * w->t_cap;
            const long long path_bytes = (path_cells * 4 + 15) / 16 * 16, left_bytes = ((path_cells + w->n_starts) * 4 + 15) / 16 * 16;
            if (w->count_path && e->cfg.cnt_path && n_walkers < (1ll << 32) && smem + path_bytes + left_bytes <= e->max_smem) {
                p.cnt_path = w->count_path;
                p.sm_path = smem; smem += (int)path_bytes;
                p.sm_left = smem; smem += (int)left_bytes;
            } else if (e->cfg.cnt_cache && n_walkers < (1ll << 32)) {
                // spare shared memory becomes a cache of hot bins (walk_philox.cuh); a slot's 32-bit count
                // cannot wrap while the launch has fewer than 2^32 walker-steps per bin: bound by walkers (NRW_CNT_CACHE=0 turns it off)
                int bits = 14;                                        // 16384 slots x 8 B = 128 KB at most
                while (bits >= 10 && smem + (8 << bits) > e->max_smem) --bits;
                if (bits >= 10) { p.sm_ccache = smem; p.ccache_bits = bits; smem += 8 << bits; }
            }
        }
    }
    plan->tab = tab; plan->tl = tl; plan->smem = smem;

    const long long n = w->walker_hi - w->walker_lo;
    long long grid = (n + WALK_THREADS - 1) / WALK_THREADS;
    if (grid < 1) grid = 1;
    if (grid > e->n_sms) grid = e->n_sms;
    plan->grid = (int)grid;
    // big ranges per warp: 6 (measured on C5 / C1 at 1.8e7 .. 1.35e8 walkers with the guided tail below:
    // 4..8 within 1 % on the large launches, 6 best on the mid-size one; NRW_RANGES_PER_WARP overrides)
    const long long rpw = e->cfg.ranges_per_warp;
    long long range = n / (grid * (WALK_THREADS / 32) * rpw);
    range = (range / 32) * 32;
    if (range < 32) range = 32;
    if (range > 4096) range = 4096;
    p.range = range;

    plan->fn = nullptr; plan->name = ""; plan->block = WALK_THREADS;
#define X(TAB, TL)                                                           \
    if (tab == TAB && tl == TL) {                                            \
        plan->fn = (const void*)walk_philox_kernel<TAB, TL>;                 \
        plan->name = "walk_philox<" #TAB "," #TL ">";                        \
    }
    NRW_WALK_VARIANTS(X)
#undef X
    // score-only walks on compact tables take the trimmed kernel (walk_score.cuh); the environment
    // variable NRW_K1=general forces the general kernel (A/B measurements, parity tests of both)
    if ((tl == TL_SCORE || tl == TL_HIST_S || tl == (TL_SCORE | TL_HIST_S)) && !w->score_sq && w->nt >= 2 &&
        tab_compact(tab)) {
        const bool hist = (tl & TL_HIST_S) != 0;
        const bool general = e->cfg.k1_general != 0 || !k1s_horizon_ok;
        const bool idw = p.cwpack == 0x03020100u;
        const bool det = e->cfg.det != 0;                    // NRW_DET=0: draw on deterministic rows too
        if (!general) {
            // a lane's 32-bit step and score sums are flushed at every range fetch: keep
            // range * nt * (largest per-step weight) < 2^31
            long long maxw = 3;                                  // IDW: the weight is the class index
            if (!idw)
                for (int i = 0; i < 4; ++i) {
                    const long long wi = (long long)((p.cwpack >> (8 * i)) & 0xffu);
                    if (wi > maxw) maxw = wi;
                }
            while (p.range > 32 && p.range * (long long)w->nt * maxw >= (1ll << 31)) p.range = ((p.range / 2) / 32) * 32;
            // one CTA of 1024 threads per SM.  NRW_CPS=2: two CTAs of 768 threads when the tables leave room
            // (48 instead of 32 warps per SM) -- +1.5 % before the range sizes were guided, 4-40 % slower
            // since (C1 mesh, 9e5 .. 1.35e8 walkers): more warps draw more tickets for no gain in issue rate
            const int cps = (plan->smem <= 112 * 1024 && e->cfg.cps == 2) ? 2 : 1;
#define Y(TAB, IDW, DET, CPS)                                                       \
    if (tab == TAB && idw == IDW && det == DET && cps == CPS && str && CPS == 1) {  \
        if (hist) {                                                                 \
            plan->fn = (const void*)walk_score_kernel<TAB, IDW, DET, 1, true, false, true>;  \
            plan->name = "walk_score_hist_str<" #TAB "," #IDW "," #DET ",cps=1>";   \
        } else {                                                                    \
            plan->fn = (const void*)walk_score_kernel<TAB, IDW, DET, 1, false, false, true>; \
            plan->name = "walk_score_str<" #TAB "," #IDW "," #DET ",cps=1>";        \
        }                                                                           \
    }                                                                               \
    if (tab == TAB && idw == IDW && det == DET && cps == CPS && !str) {             \
        if (hist) {                                                                 \
            plan->fn = (const void*)walk_score_kernel<TAB, IDW, DET, CPS, true>;    \
            plan->name = "walk_score_hist<" #TAB "," #IDW "," #DET ",cps=" #CPS ">"; \
        } else {                                                                    \
            plan->fn = (const void*)walk_score_kernel<TAB, IDW, DET, CPS, false>;   \
            plan->name = "walk_score<" #TAB "," #IDW "," #DET ",cps=" #CPS ">";     \
        }                                                                           \
    }
#define Y2(TAB, IDW, DET) Y(TAB, IDW, DET, 1) Y(TAB, IDW, DET, 2)
            Y2(TAB_SMEM, true, true) Y2(TAB_SMEM, false, true) Y2(TAB_HYB, true, true) Y2(TAB_HYB, false, true)
            Y2(TAB_SMEM, true, false) Y2(TAB_SMEM, false, false) Y2(TAB_HYB, true, false) Y2(TAB_HYB, false, false)
            Y2(TAB_FAC, true, true) Y2(TAB_FAC, false, true) Y2(TAB_FAC, true, false) Y2(TAB_FAC, false, false)
#undef Y2
            if (cps == 1 && e->cfg.block != WALK_THREADS) {        // occupancy experiment: fewer warps per SM
                plan->block = e->cfg.block;
                long long r1 = n / ((long long)plan->grid * (plan->block / 32) * rpw);
                r1 = (r1 / 32) * 32;
                p.range = r1 < 32 ? 32 : (r1 > 4096 ? 4096 : r1);
                while (p.range > 32 && p.range * (long long)w->nt * maxw >= (1ll << 31)) p.range = ((p.range / 2) / 32) * 32;
            }
            if (cps == 2) {
                plan->block = 768;
                long long g2 = (n + 767) / 768;
                if (g2 > 2ll * e->n_sms) g2 = 2ll * e->n_sms;
                plan->grid = (int)(g2 < 1 ? 1 : g2);
                long long r2 = n / ((long long)plan->grid * 24 * rpw);
                r2 = (r2 / 32) * 32;
                p.range = r2 < 32 ? 32 : (r2 > 4096 ? 4096 : r2);
                while (p.range > 32 && p.range * (long long)w->nt * maxw >= (1ll << 31)) p.range = ((p.range / 2) / 32) * 32;
            }
#undef Y
        }
    }
    // the structured descriptors were staged for the trimmed walk; any other kernel needs the alias tables
    if (str && strncmp(plan->name, "walk_score", 10) != 0) return make_plan(e, t, w, plan, false);
    // time-resolved counts through the likely paths, nothing else tallied, compact tables: the trimmed walk too
    if (tl == TL_CNT_G && p.cnt_path && tab_compact(tab) && w->nt >= 2 && !e->cfg.k1_general &&
        32ll * (long long)w->nt < (1ll << 31)) {
        while (p.range > 32 && p.range * (long long)w->nt >= (1ll << 31)) p.range = ((p.range / 2) / 32) * 32;
#define Z(TAB)                                                                              \
    if (tab == TAB) {                                                                       \
        plan->fn = (const void*)walk_score_kernel<TAB, true, false, 1, false, true>;        \
        plan->name = "walk_count<" #TAB ">";                                                \
    }
        Z(TAB_SMEM) Z(TAB_HYB) Z(TAB_FAC)
#undef Z
    }
    // time-resolved counts in the CTA's own shared-memory histogram with plain atomics (small graphs over long windows:
    // the 3-state chain of simple_flux over 1 002 steps), nothing else tallied, compact tables: the trimmed walk too
    if (tl == TL_CNT_S && !p.cnt_match && tab == TAB_SMEM && w->nt >= 2 && !e->cfg.k1_general &&
        32ll * (long long)w->nt < (1ll << 31)) {
        while (p.range > 32 && p.range * (long long)w->nt >= (1ll << 31)) p.range = ((p.range / 2) / 32) * 32;
        plan->fn = (const void*)walk_score_kernel<TAB_SMEM, true, false, 1, false, true, false, true>;
        plan->name = "walk_count_smem<TAB_SMEM>";
    }
    {   // guided range sizes (fetch_range, walk_philox.cuh).  NRW_RANGE_TAIL / NRW_TAIL_RANGES override
        // (0 ranges = one size)
        const long long warps = (long long)plan->grid * (plan->block / 32);
        p.n_phase = guided_phases(n, p.range, warps, e->cfg.range_tail, e->cfg.tail_ranges,
                                  p.ph_start, p.ph_ticket, p.ph_range);
    }
    if (!plan->fn)
        return fail(NRW_EINVAL, "nrw_walk_philox: tally combination %d is not built for table mode %d "
                    "(score=1 time_weight=2 hist_smem=4 hist_global=8 count_smem=16 count_global=32)",
                    tl, tab);
    return NRW_OK;
}

extern "C" int nrw_walk_philox_plan(nrw_engine* e, const nrw_tables* t, const nrw_walk* w, char* name,
                                    int name_len, int* grid, int* block, int* smem_bytes) {
    WalkPlan plan;
    const int rc = make_plan(e, t, w, &plan);
    if (rc) return rc;
    if (name && name_len > 0) { strncpy(name, plan.name, name_len - 1); name[name_len - 1] = 0; }
    if (grid) *grid = plan.grid;
    if (block) *block = plan.block;
    if (smem_bytes) *smem_bytes = plan.smem;
    return NRW_OK;
}

// Opt a kernel in to `smem` bytes of dynamic shared memory once (the attribute call costs microseconds per launch).
// The attribute belongs to the function on a device, not to an engine: the cache is per (device, kernel), shared by
// every engine of the process, and only ever raises the limit.
static int ensure_smem(nrw_engine* e, const void* fn, int smem) {
    static std::mutex mu;
    static std::unordered_map<unsigned long long, int> opted;          // (device, kernel) -> bytes opted in
    const unsigned long long key = (unsigned long long)(uintptr_t)fn * 64ull + (unsigned long long)e->device;
    std::lock_guard<std::mutex> lock(mu);
    auto it = opted.find(key);
    if (it != opted.end() && it->second >= smem) return NRW_OK;
    CUDA_TRY(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    opted[key] = smem;
    return NRW_OK;
}

extern "C" int nrw_walk_philox(nrw_engine* e, const nrw_tables* t, const nrw_walk* w, void* stream) {
    WalkPlan plan;
    int rc = make_plan(e, t, w, &plan);
    if (rc) return rc;
    if (w->walker_hi == w->walker_lo && !w->reduce) return NRW_OK;
    CUDA_TRY(cudaSetDevice(e->device));
    if ((rc = ensure_smem(e, plan.fn, plan.smem))) return rc;
    // this launch's ticket slot (walk_tail.cuh): all zero when free; the launch's last tail CTA zeroes it again
    unsigned long long* slot = e->slots + TICKET_WORDS * (e->next_slot.fetch_add(1) % NRW_TICKET_SLOTS);
    plan.p.work_next = slot;
    plan.p.tail.slot = slot;
    void* args[] = {&plan.p};
    CUDA_TRY(cudaLaunchKernel(plan.fn, dim3(plan.grid), dim3(plan.block), args, plan.smem,
                              (cudaStream_t)stream));
    return NRW_OK;
}

// ---- K2 ---------------------------------------------------------------------------------------
extern "C" int nrw_walk_replay(nrw_engine* e, const nrw_tables* t, const nrw_replay* r, void* stream) {
    if (!e || !t || !r) return fail(NRW_EINVAL, "nrw_walk_replay: NULL argument");
    if (!t->indptr || !t->cols || !t->cum) return fail(NRW_EINVAL, "nrw_walk_replay: needs indptr/cols/cum");
    if (r->n_starts <= 0 || !r->start_nodes || !r->start_counts || r->nt < 1)
        return fail(NRW_EINVAL, "nrw_walk_replay: bad start list / nt");
    if (!r->u_stream || r->u_len < 0 || !r->ev_t || !r->ev_node || !r->ev_count || !r->ev_offsets ||
        !r->scratch || !r->state)
        return fail(NRW_EINVAL, "nrw_walk_replay: NULL buffer");
    ReplayParams p;
    p.n_rows = t->n_rows; p.n_live = t->n_live;
    p.indptr = t->indptr; p.cols = t->cols; p.cum = t->cum;
    p.n_starts = r->n_starts; p.start_nodes = r->start_nodes;
    p.start_counts = (const long long*)r->start_counts; p.nt = r->nt;
    p.u = r->u_stream; p.u_len = r->u_len; p.u_base = r->u_base;
    p.ev_t = r->ev_t; p.ev_node = r->ev_node; p.ev_count = (long long*)r->ev_count;
    p.ev_cap = r->ev_cap; p.ev_offsets = (long long*)r->ev_offsets;
    p.scratch = r->scratch; p.state = (long long*)r->state;
    CUDA_TRY(cudaSetDevice(e->device));
    walk_replay_kernel<<<1, REPLAY_THREADS, 0, (cudaStream_t)stream>>>(p);
    CUDA_TRY(cudaGetLastError());
    return NRW_OK;
}

// ---- end-to-end with host buffers -------------------------------------------------------------
static int grow(void** buf, size_t* cap, size_t need) {
    if (need <= *cap) return NRW_OK;
    if (*buf) cudaFree(*buf);
    *buf = nullptr; *cap = 0;
    size_t want = need + need / 4 + 256;
    cudaError_t err = cudaMalloc(buf, want);
    if (err != cudaSuccess) return fail(NRW_ECUDA, "cudaMalloc(%zu) failed: %s", want, cudaGetErrorString(err));
    *cap = want;
    return NRW_OK;
}

// H2D of the transition tables of a host-buffer solve: a dry plan says which form the walk will read -- the
// structured descriptors alone (t->rowinfo_str, staged where the descriptors go) or descriptors + alias patterns --
// and only that form is copied; t->rowinfo_str is cleared when the alias form is taken.
static int stage_tables_host(nrw_engine* e, nrw_tables* t, const nrw_walk* w, const void* rowinfo_host,
                             const void* alias_host, const void* rowinfo_str_host, cudaStream_t st) {
    bool use_str = false;
    if (rowinfo_str_host) {
        WalkPlan probe;
        const int rc = make_plan(e, t, w, &probe);
        if (rc) return rc;
        use_str = strncmp(probe.name, "walk_score_str", 14) == 0 || strncmp(probe.name, "walk_score_hist_str", 19) == 0;
    }
    if (use_str) {
        CUDA_TRY(cudaMemcpyAsync((void*)t->rowinfo_str, rowinfo_str_host, (size_t)t->n_rows * 8, cudaMemcpyHostToDevice, st));
    } else {
        t->rowinfo_str = nullptr;
        CUDA_TRY(cudaMemcpyAsync((void*)t->rowinfo, rowinfo_host, (size_t)t->n_rows * 8, cudaMemcpyHostToDevice, st));
        CUDA_TRY(cudaMemcpyAsync((void*)t->alias, alias_host, (size_t)t->n_alias * 8, cudaMemcpyHostToDevice, st));
    }
    return NRW_OK;
}

extern "C" int nrw_fluence_scores_host(nrw_engine* e, int32_t n_rows, int32_t n_live, int32_t n_alias,
                                       const void* rowinfo_host, const void* alias_host,
                                       const void* rowinfo_str_host, uint32_t str_thresh, uint32_t str_n,
                                       int32_t n_starts, const int32_t* start_nodes_host,
                                       const int64_t* start_prefix_host, const int64_t* start_id_base_host,
                                       int64_t walker_lo, int64_t walker_hi, uint64_t seed, int32_t nt,
                                       const double class_value[4], int64_t* score_host,
                                       unsigned long long* counters_host, const nrw_reduce* reduce, void* stream) {
    if (!e || !rowinfo_host || !alias_host || !start_nodes_host || !start_prefix_host || !score_host ||
        !class_value || n_rows <= 0 || n_alias <= 0 || n_starts <= 0)
        return fail(NRW_EINVAL, "nrw_fluence_scores_host: bad argument");
    const long long n_tally = (long long)n_starts + NRW_N_COUNTERS;
    if (reduce && (reduce->n != n_tally || !reduce->own_buf))
        return fail(NRW_EINVAL, "nrw_fluence_scores_host: reduce->n must be n_starts + %d", NRW_N_COUNTERS);
    cudaStream_t st = (cudaStream_t)stream;
    std::lock_guard<std::mutex> lock(e->mu_host);              // the staging buffers below are per engine
    CUDA_TRY(cudaSetDevice(e->device));
    int rc;
    if ((rc = grow(&e->d_rowinfo, &e->cap_rowinfo, (size_t)n_rows * 8 + 16))) return rc;
    if ((rc = grow(&e->d_alias, &e->cap_alias, (size_t)n_alias * 8 + 16))) return rc;
    if ((rc = grow(&e->d_starts, &e->cap_starts, (size_t)n_starts * 4))) return rc;
    if ((rc = grow(&e->d_prefix, &e->cap_prefix, (size_t)(n_starts + 1) * 8))) return rc;
    if (start_id_base_host && (rc = grow(&e->d_idbase, &e->cap_idbase, (size_t)n_starts * 8))) return rc;
    long long* tally;                                           // scores, then the counters: one buffer
    if (reduce) {
        tally = (long long*)reduce->own_buf + reduce->offset;  // zeroed by the tail of the previous launch
    } else {
        if ((rc = grow(&e->d_tally, &e->cap_tally, (size_t)n_tally * 8))) return rc;
        tally = (long long*)e->d_tally;
        CUDA_TRY(cudaMemsetAsync(tally, 0, (size_t)n_tally * 8, st));
    }
    nrw_tables t;
    memset(&t, 0, sizeof(t));
    t.n_rows = n_rows; t.n_live = n_live; t.alias_layout = NRW_ALIAS_COMPACT; t.n_alias = n_alias;
    t.rowinfo = e->d_rowinfo; t.alias = e->d_alias;
    // the structured form of the tables, when the caller has it (nrw_tables.rowinfo_str): it is staged in the
    // descriptor buffer INSTEAD of descriptors + alias patterns if the planner takes the structured walk
    if (rowinfo_str_host) { t.rowinfo_str = e->d_rowinfo; t.str_thresh = str_thresh; t.str_n = str_n; }
    nrw_walk w;
    memset(&w, 0, sizeof(w));
    w.n_starts = n_starts; w.start_nodes = (const int32_t*)e->d_starts;
    w.start_prefix = (const int64_t*)e->d_prefix;
    w.start_id_base = start_id_base_host ? (const int64_t*)e->d_idbase : nullptr;
    w.walker_lo = walker_lo; w.walker_hi = walker_hi; w.seed = seed; w.nt = nt;
    w.sampler = NRW_SAMPLER_ALIAS; w.score = (int64_t*)tally;
    // the start list goes first: these copies run while the host looks at the classes and plans the launch
    CUDA_TRY(cudaMemcpyAsync(e->d_starts, start_nodes_host, (size_t)n_starts * 4, cudaMemcpyHostToDevice, st));
    CUDA_TRY(cudaMemcpyAsync(e->d_prefix, start_prefix_host, (size_t)(n_starts + 1) * 8, cudaMemcpyHostToDevice, st));
    if (start_id_base_host)
        CUDA_TRY(cudaMemcpyAsync(e->d_idbase, start_id_base_host, (size_t)n_starts * 8, cudaMemcpyHostToDevice, st));
    {   // weights of classes no row carries are free: if the used ones equal the class index, pass the
        // identity vector so that the walk takes the IDW kernel (walk_score.cuh)
        uint32_t present = 0;
        const uint32_t* ri = (const uint32_t*)rowinfo_host;
        for (int32_t r = 0; r < n_rows; ++r) present |= 1u << row_class(ri[2 * r]);
        bool ident = true;
        for (uint32_t c = 0; c < 4; ++c) if ((present >> c) & 1u) ident = ident && class_value[c] == (double)c;
        for (int i = 0; i < 4; ++i) w.class_value[i] = ident ? (double)i : class_value[i];
    }
    w.counters = (unsigned long long*)(tally + n_starts);
    w.reduce = reduce;
    if ((rc = stage_tables_host(e, &t, &w, rowinfo_host, alias_host, rowinfo_str_host, st))) return rc;
    if ((rc = nrw_walk_philox(e, &t, &w, stream))) return rc;
    const long long* result = (reduce && reduce->world > 1) ? (const long long*)reduce->own_buf + reduce->out_offset : tally;
    // counters (+ the exchange's status word: it follows the sums, after the pad word of an odd n -- n_tally is
    // n_starts + 8, so the status sits at n_tally or n_tally + 1) into the engine's PINNED tail buffer (a copy to a
    // pageable stack array goes through the driver's staging path), then the scores
    long long* tail_host = e->h_tail;
    const bool want_status = reduce && reduce->world > 1;
    if (counters_host || want_status)
        CUDA_TRY(cudaMemcpyAsync(tail_host, result + n_starts, (size_t)(NRW_N_COUNTERS + (want_status ? 2 : 0)) * 8,
                                 cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaMemcpyAsync(score_host, result, (size_t)n_starts * 8, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaStreamSynchronize(st));
    if (counters_host) memcpy(counters_host, tail_host, NRW_N_COUNTERS * sizeof(unsigned long long));
    if (want_status && tail_host[NRW_N_COUNTERS + (n_tally & 1)] != 0)
        return fail(NRW_ECUDA, "nrw_fluence_scores_host: a peer did not reach the reduction within 5 s");
    return NRW_OK;
}

// ---- the host-buffer entry without any allocation (no counterpart in the reference; operators' request) -------
// ALL device-side buffers of the call live in this file-scope array: every GPU that loads the library holds its
// 1 MiB until the process exits.  In HOST code the name `g_static_scratch` is a host-side stand-in, not a device
// address: the address a kernel or a copy may use comes from cudaGetSymbolAddress, resolved once per device with
// that device current (static_scratch_base).  Calls share the array and it keeps the last call's contents.
static __device__ __align__(256) unsigned char g_static_scratch[NRW_STATIC_SCRATCH_BYTES];

static std::mutex g_static_mu;                               // one caller at a time: the array is shared
static unsigned char* g_static_base[64] = {nullptr};         // device address of g_static_scratch per device

// Device address of the static scratch on `device` (which the caller has made current); resolved once per device.
static int static_scratch_base(int device, unsigned char** base) {
    if (device < 0 || device >= 64) return fail(NRW_EINVAL, "nrw_fluence_scores_host_static: device %d", device);
    if (!g_static_base[device]) {
        void* ptr = nullptr;
        CUDA_TRY(cudaGetSymbolAddress(&ptr, g_static_scratch));
        if (!ptr) return fail(NRW_ECUDA, "nrw_fluence_scores_host_static: cudaGetSymbolAddress returned NULL");
        g_static_base[device] = (unsigned char*)ptr;
    }
    *base = g_static_base[device];
    return NRW_OK;
}

extern "C" int nrw_fluence_scores_host_static(nrw_engine* e, int32_t n_rows, int32_t n_live, int32_t n_alias,
                                              const void* rowinfo_host, const void* alias_host,
                                              const void* rowinfo_str_host, uint32_t str_thresh, uint32_t str_n,
                                              int32_t n_starts, const int32_t* start_nodes_host,
                                              const int64_t* start_prefix_host, const int64_t* start_id_base_host,
                                              int64_t walker_lo, int64_t walker_hi, uint64_t seed, int32_t nt,
                                              const double class_value[4], int64_t* score_host,
                                              unsigned long long* counters_host, void* stream) {
    if (!e || !rowinfo_host || !alias_host || !start_nodes_host || !start_prefix_host || !score_host ||
        !class_value || n_rows <= 0 || n_alias <= 0 || n_starts <= 0)
        return fail(NRW_EINVAL, "nrw_fluence_scores_host_static: bad argument");
    // carve-up of the static array (256-byte aligned pieces); sizes beyond it are refused before anything is touched
    auto up256 = [](size_t b) { return (b + 255) & ~(size_t)255; };
    const size_t n_tally = (size_t)n_starts + NRW_N_COUNTERS;
    const size_t o_slot = 0;
    const size_t o_tally = o_slot + 256;
    const size_t o_row = o_tally + up256(n_tally * 8);
    const size_t o_alias = o_row + up256((size_t)n_rows * 8);
    const size_t o_starts = o_alias + up256((size_t)n_alias * 8);
    const size_t o_prefix = o_starts + up256((size_t)n_starts * 4);
    const size_t o_idbase = o_prefix + up256(((size_t)n_starts + 1) * 8);
    const size_t need = o_idbase + (start_id_base_host ? up256((size_t)n_starts * 8) : 0);
    if (need > (size_t)NRW_STATIC_SCRATCH_BYTES)
        return fail(NRW_ENOSPC, "nrw_fluence_scores_host_static: n_rows=%d n_alias=%d n_starts=%d need %zu bytes of scratch, "
                    "the static array holds %d (use nrw_fluence_scores_host)", n_rows, n_alias, n_starts, need,
                    (int)NRW_STATIC_SCRATCH_BYTES);
    cudaStream_t st = (cudaStream_t)stream;
    std::lock_guard<std::mutex> lock(g_static_mu);
    CUDA_TRY(cudaSetDevice(e->device));
    unsigned char* base = nullptr;
    int rc;
    if ((rc = static_scratch_base(e->device, &base))) return rc;
    unsigned long long* slot = (unsigned long long*)(base + o_slot);
    long long* tally = (long long*)(base + o_tally);
    // the array keeps the previous call's contents: ticket slot and tally must start at zero
    CUDA_TRY(cudaMemsetAsync(base + o_slot, 0, o_row, st));
    CUDA_TRY(cudaMemcpyAsync(base + o_starts, start_nodes_host, (size_t)n_starts * 4, cudaMemcpyHostToDevice, st));
    CUDA_TRY(cudaMemcpyAsync(base + o_prefix, start_prefix_host, ((size_t)n_starts + 1) * 8, cudaMemcpyHostToDevice, st));
    if (start_id_base_host)
        CUDA_TRY(cudaMemcpyAsync(base + o_idbase, start_id_base_host, (size_t)n_starts * 8, cudaMemcpyHostToDevice, st));
    nrw_tables t;
    memset(&t, 0, sizeof(t));
    t.n_rows = n_rows; t.n_live = n_live; t.alias_layout = NRW_ALIAS_COMPACT; t.n_alias = n_alias;
    t.rowinfo = base + o_row; t.alias = base + o_alias;
    if (rowinfo_str_host) { t.rowinfo_str = base + o_row; t.str_thresh = str_thresh; t.str_n = str_n; }
    nrw_walk w;
    memset(&w, 0, sizeof(w));
    w.n_starts = n_starts; w.start_nodes = (const int32_t*)(base + o_starts);
    w.start_prefix = (const int64_t*)(base + o_prefix);
    w.start_id_base = start_id_base_host ? (const int64_t*)(base + o_idbase) : nullptr;
    w.walker_lo = walker_lo; w.walker_hi = walker_hi; w.seed = seed; w.nt = nt;
    w.sampler = NRW_SAMPLER_ALIAS; w.score = (int64_t*)tally;
    {   // as in nrw_fluence_scores_host: identity weights on the classes in use select the IDW kernel
        uint32_t present = 0;
        const uint32_t* ri = (const uint32_t*)rowinfo_host;
        for (int32_t r = 0; r < n_rows; ++r) present |= 1u << row_class(ri[2 * r]);
        bool ident = true;
        for (uint32_t c = 0; c < 4; ++c) if ((present >> c) & 1u) ident = ident && class_value[c] == (double)c;
        for (int i = 0; i < 4; ++i) w.class_value[i] = ident ? (double)i : class_value[i];
    }
    w.counters = (unsigned long long*)(tally + n_starts);
    if ((rc = stage_tables_host(e, &t, &w, rowinfo_host, alias_host, rowinfo_str_host, st))) return rc;
    // plan + launch as nrw_walk_philox does, with the ticket slot in the static array instead of the engine's ring
    WalkPlan plan;
    if ((rc = make_plan(e, &t, &w, &plan))) return rc;
    if (w.walker_hi > w.walker_lo) {
        if ((rc = ensure_smem(e, plan.fn, plan.smem))) return rc;
        plan.p.work_next = slot;
        plan.p.tail.slot = slot;
        void* args[] = {&plan.p};
        CUDA_TRY(cudaLaunchKernel(plan.fn, dim3(plan.grid), dim3(plan.block), args, plan.smem, st));
    }
    CUDA_TRY(cudaMemcpyAsync(score_host, tally, (size_t)n_starts * 8, cudaMemcpyDeviceToHost, st));
    if (counters_host)
        CUDA_TRY(cudaMemcpyAsync(counters_host, tally + n_starts, NRW_N_COUNTERS * sizeof(unsigned long long),
                                 cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaStreamSynchronize(st));
    return NRW_OK;
}

// ---- the path's collective over peer memory ------------------------------------------------------
extern "C" int nrw_p2p_allreduce_i64(nrw_engine* e, const void* peer_bufs_dev, const void* peer_flags_dev,
                                     int64_t offset, int64_t zero_offset, int64_t out_offset, int64_t n,
                                     int32_t rank, int32_t world, uint64_t epoch, int32_t* err_flag, void* stream) {
    if (!e || !peer_bufs_dev || !peer_flags_dev || !err_flag || n <= 0 || offset < 0 || out_offset < 0 || world < 1 ||
        world > 64 || rank < 0 || rank >= world || epoch == 0 ||
        ((offset | out_offset | (zero_offset < 0 ? 0 : zero_offset)) & 1) != 0)
        return fail(NRW_EINVAL, "nrw_p2p_allreduce_i64: bad argument");
    CUDA_TRY(cudaSetDevice(e->device));
    TailParams r;
    memset(&r, 0, sizeof(r));
    r.world = world; r.rank = rank;
    r.bufs = (long long* const*)peer_bufs_dev;
    r.flags = (unsigned long long* const*)peer_flags_dev;
    r.own = nullptr;                                            // resolved on the device: bufs[rank]
    r.off = offset; r.zero_off = zero_offset; r.out_off = out_offset; r.n = n; r.epoch = epoch;
    r.status = nullptr;
    r.timeout_ns = 5000000000ull;                               // 5 s without a peer: give up
    // 80 KB per peer: latency-, not bandwidth-bound -- one 16-byte pair per thread, every load in flight at once
    long long grid = ((n + 1) / 2 + P2P_THREADS - 1) / P2P_THREADS;
    if (grid > P2P_CTAS) grid = P2P_CTAS;
    p2p_allreduce_i64_kernel<<<(int)grid, P2P_THREADS, 0, (cudaStream_t)stream>>>(r, err_flag);
    CUDA_TRY(cudaGetLastError());
    return NRW_OK;
}

// ---- roofline denominators ----------------------------------------------------------------------
extern "C" int nrw_microbench(nrw_engine* e, int32_t kind, int32_t n, int32_t iters, const void* table,
                              unsigned long long* hist, unsigned long long* scratch, void* stream) {
    if (!e || n <= 0 || n > (1 << 26) || iters <= 0 || !scratch) return fail(NRW_EINVAL, "nrw_microbench: bad argument");
    cudaStream_t st = (cudaStream_t)stream;
    CUDA_TRY(cudaSetDevice(e->device));
    const int grid = e->n_sms;
    int smem = 0;
    const void* fn = nullptr;
    switch (kind) {
        case MB_SMEM_ATOMIC: fn = (const void*)microbench_kernel<MB_SMEM_ATOMIC>; smem = align16(n * 4); break;
        case MB_SMEM_LOAD:   fn = (const void*)microbench_kernel<MB_SMEM_LOAD>;   smem = align16(n * 8); break;
        case MB_L2_LOAD:     fn = (const void*)microbench_kernel<MB_L2_LOAD>;
            if (!table) return fail(NRW_EINVAL, "nrw_microbench: L2 load needs a table"); break;
        case MB_GLOBAL_ATOMIC: fn = (const void*)microbench_kernel<MB_GLOBAL_ATOMIC>;
            if (!hist) return fail(NRW_EINVAL, "nrw_microbench: global atomics need a histogram"); break;
        default: return fail(NRW_EINVAL, "nrw_microbench: kind=%d", kind);
    }
    if (smem > e->max_smem) return fail(NRW_ENOSPC, "nrw_microbench: %d B of shared memory", smem);
    if (int rc = ensure_smem(e, fn, smem > 0 ? smem : 16)) return rc;
    const uint2* tab = (const uint2*)table;
    void* args[] = {&n, &iters, &tab, &hist, &scratch};
    CUDA_TRY(cudaLaunchKernel(fn, dim3(grid), dim3(1024), args, smem, st));
    return NRW_OK;
}

// ---- result-file text (host only): np.savetxt(path, a, delimiter=',') of fluence_mini_app.py:446 -------------
// '%.18e' per value, ',' between the values of a row, '\n' after each row -- the bytes np.savetxt writes.  Correctly
// rounded 19-digit conversion costs ~0.3 us per value in any libc (and in CPython): on the 10 000-cell mesh that was
// 2.6 ms of a 5.9 ms run(args), so the rows are converted by a few threads, each into its own slice of `out` (a value
// takes at most 27 bytes with its delimiter; 28 are reserved), and compacted in place afterwards.
static size_t format_rows_e18(const double* a, int64_t r0, int64_t r1, int64_t ncols, char* dst) {
    char* p = dst;
    for (int64_t r = r0; r < r1; ++r) {
        for (int64_t c = 0; c < ncols; ++c) {
            const double x = a[r * ncols + c];
            if (x != x) { memcpy(p, "nan", 3); p += 3; }                 // numpy / CPython print no sign on a nan
            else p += snprintf(p, 32, "%.18e", x);
            *p++ = (c + 1 < ncols) ? ',' : '\n';
        }
    }
    return (size_t)(p - dst);
}

extern "C" int nrw_format_csv_e18(const double* a, int64_t nrows, int64_t ncols, char* out, int64_t cap,
                                  int64_t* len, int32_t n_threads) {
    if (!a || !out || !len || nrows < 0 || ncols <= 0) return fail(NRW_EINVAL, "nrw_format_csv_e18: bad argument");
    {   // np.savetxt's text does not depend on the locale; snprintf's decimal point does
        const struct lconv* lc = localeconv();
        if (lc && lc->decimal_point && strcmp(lc->decimal_point, ".") != 0)
            return fail(NRW_EINVAL, "nrw_format_csv_e18: LC_NUMERIC has decimal point '%s'; format in the caller", lc->decimal_point);
    }
    const int64_t per_value = 28;                                      // -d.{18}e+XXX (26) + delimiter, rounded up
    if (cap < nrows * ncols * per_value) return fail(NRW_ENOSPC, "nrw_format_csv_e18: out needs %lld bytes, has %lld",
                                                     (long long)(nrows * ncols * per_value), (long long)cap);
    int nt = n_threads < 1 ? 1 : (n_threads > 64 ? 64 : n_threads);
    if ((int64_t)nt > nrows) nt = (int)(nrows < 1 ? 1 : nrows);
    if (nrows * ncols < 2048) nt = 1;
    std::vector<size_t> lens((size_t)nt, 0);
    std::vector<int64_t> first((size_t)nt + 1, 0);
    for (int t = 0; t <= nt; ++t) first[(size_t)t] = nrows * t / nt;
    auto work = [&](int t) {
        lens[(size_t)t] = format_rows_e18(a, first[(size_t)t], first[(size_t)t + 1], ncols,
                                          out + first[(size_t)t] * ncols * per_value);
    };
    std::vector<std::thread> th;
    for (int t = 1; t < nt; ++t) th.emplace_back(work, t);
    work(0);
    for (auto& x : th) x.join();
    size_t total = lens[0];
    for (int t = 1; t < nt; ++t) {                                     // compact the slices (they only move left)
        memmove(out + total, out + first[(size_t)t] * ncols * per_value, lens[(size_t)t]);
        total += lens[(size_t)t];
    }
    *len = (int64_t)total;
    return NRW_OK;
}
