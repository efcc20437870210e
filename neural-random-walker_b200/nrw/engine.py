"""ctypes binding of libnrw.so (include/nrw.h) + the PyTorch-side buffer management.

PyTorch owns every device buffer (tables, start lists, tallies, uniform streams); the C library
gets raw pointers and a stream.  There is NO CPU fallback: importing this module is harmless
without a GPU (the CLI can still parse arguments and build matrices), but constructing an
``Engine`` raises unless libnrw.so is built and a CUDA device is present.
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass

import numpy as np

from . import tables as _tables
from .mesh import TransitionCSR

_HERE = os.path.dirname(os.path.abspath(__file__))
# NRW_LIB: another build of the same library (A/B measurements of kernel changes); normally the in-tree one
LIB_PATH = os.environ.get("NRW_LIB") or os.path.join(_HERE, "csrc", "libnrw.so")

NRW_N_COUNTERS = 8
COUNTER_NAMES = ("steps", "exits", "survivors", "t_overflow", "walkers")
SAMPLER = {"alias": 0, "cdf": 1, "tree": 2}
ALIAS_LAYOUT = {"none": 0, "compact": 1, "wide": 2}


class NrwError(RuntimeError):
    pass


class _Tables(C.Structure):
    _fields_ = [("n_rows", C.c_int32), ("n_live", C.c_int32),
                ("indptr", C.c_void_p), ("cols", C.c_void_p), ("cum", C.c_void_p),
                ("node_class", C.c_void_p),
                ("alias_layout", C.c_int32), ("n_alias", C.c_int32),
                ("rowinfo", C.c_void_p), ("alias", C.c_void_p),
                ("fac_dir", C.c_void_p), ("fac_tail", C.c_void_p), ("fac_cls", C.c_void_p),
                ("fac_period", C.c_int32), ("fac_cells", C.c_int32), ("fac_n_tail", C.c_int32),
                ("fac_shift", C.c_int32), ("fac_magic", C.c_uint32), ("fac_exit_w0", C.c_uint32),
                ("fac_exit_base", C.c_uint32),
                ("tree", C.c_void_p), ("tree_ptr", C.c_void_p),
                ("rowinfo_str", C.c_void_p), ("str_thresh", C.c_uint32), ("str_n", C.c_uint32),
                ("fac_exit_w0_str", C.c_uint32),
                ("fac_dir_str", C.c_void_p), ("fac_tail_str", C.c_void_p)]


class _Reduce(C.Structure):
    """nrw_reduce (include/nrw.h): the fused all-reduce / next-buffer zeroing a walk runs in its tail."""
    _fields_ = [("rank", C.c_int32), ("world", C.c_int32),
                ("peer_bufs_dev", C.c_void_p), ("peer_flags_dev", C.c_void_p),
                ("own_buf", C.c_void_p),
                ("offset", C.c_int64), ("zero_offset", C.c_int64), ("n", C.c_int64),
                ("epoch", C.c_uint64), ("out_offset", C.c_int64)]


class _Walk(C.Structure):
    _fields_ = [("n_starts", C.c_int32), ("start_nodes", C.c_void_p), ("start_prefix", C.c_void_p),
                ("walker_lo", C.c_int64), ("walker_hi", C.c_int64), ("start_id_base", C.c_void_p),
                ("seed", C.c_uint64),
                ("nt", C.c_int32), ("sampler", C.c_int32),
                ("score", C.c_void_p), ("score_sq", C.c_void_p),
                ("class_value", C.c_double * 4), ("time_weight", C.c_void_p),
                ("node_hist", C.c_void_p),
                ("count_tn", C.c_void_p), ("t_cap", C.c_int32), ("count_per_start", C.c_int32),
                ("counters", C.c_void_p),
                ("count_path", C.c_void_p),
                ("reduce", C.POINTER(_Reduce))]


class _Replay(C.Structure):
    _fields_ = [("n_starts", C.c_int32), ("start_nodes", C.c_void_p), ("start_counts", C.c_void_p),
                ("nt", C.c_int32), ("u_stream", C.c_void_p), ("u_len", C.c_int64),
                ("ev_t", C.c_void_p), ("ev_node", C.c_void_p), ("ev_count", C.c_void_p),
                ("ev_cap", C.c_int64), ("ev_offsets", C.c_void_p), ("scratch", C.c_void_p),
                ("state", C.c_void_p), ("u_base", C.c_int64)]


_lib = None

# every symbol include/nrw.h declares; tests/test_capi.py checks the library exports them all
EXPORTS = ("nrw_abi_version", "nrw_last_error", "nrw_create", "nrw_destroy", "nrw_device_info", "nrw_engine_config",
           "nrw_build_cum", "nrw_walk_philox", "nrw_walk_philox_plan", "nrw_walk_replay",
           "nrw_assemble_events", "nrw_expand_events", "nrw_expand_counts",
           "nrw_fluence_scores_host", "nrw_fluence_scores_host_static", "nrw_microbench", "nrw_p2p_allreduce_i64", "nrw_guided_phases",
           "nrw_format_csv_e18")


def load_library():
    """dlopen libnrw.so and declare prototypes.  Raises NrwError when it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.isfile(LIB_PATH):
        raise NrwError("libnrw.so not found at %s -- build it with "
                       "neural-random-walker_b200/nrw/csrc/build.sh (or __graft_entry__.build()); "
                       "there is no CPU fallback" % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    vp, i32, i64, u64, dbl = C.c_void_p, C.c_int32, C.c_int64, C.c_uint64, C.c_double
    lib.nrw_abi_version.restype = C.c_int
    lib.nrw_last_error.restype = C.c_char_p
    lib.nrw_create.argtypes = [C.c_int, C.POINTER(vp)]
    lib.nrw_destroy.argtypes = [vp]
    lib.nrw_device_info.argtypes = [vp, C.POINTER(C.c_int), C.POINTER(C.c_int)]
    lib.nrw_engine_config.argtypes = [vp, C.c_char_p, C.c_int]
    lib.nrw_build_cum.argtypes = [vp, i32, vp, vp, vp, vp]
    lib.nrw_walk_philox.argtypes = [vp, C.POINTER(_Tables), C.POINTER(_Walk), vp]
    lib.nrw_walk_philox_plan.argtypes = [vp, C.POINTER(_Tables), C.POINTER(_Walk), C.c_char_p,
                                         C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int),
                                         C.POINTER(C.c_int)]
    lib.nrw_walk_replay.argtypes = [vp, C.POINTER(_Tables), C.POINTER(_Replay), vp]
    lib.nrw_assemble_events.argtypes = [vp, i32, vp, vp, vp, vp, vp, vp, dbl, dbl, vp, i32, i32, vp, vp]
    lib.nrw_expand_events.argtypes = [vp, i64, vp, vp, vp, i32, i32, vp, vp]
    lib.nrw_expand_counts.argtypes = [vp, vp, i32, i32, i32, i32, i64, vp, vp]
    lib.nrw_fluence_scores_host.argtypes = [vp, i32, i32, i32, vp, vp, vp, C.c_uint32, C.c_uint32, i32, vp, vp, vp, i64, i64, u64,
                                            i32, C.POINTER(dbl * 4), vp, vp, C.POINTER(_Reduce), vp]
    lib.nrw_fluence_scores_host_static.argtypes = [vp, i32, i32, i32, vp, vp, vp, C.c_uint32, C.c_uint32, i32, vp, vp, vp, i64, i64,
                                                   u64, i32, C.POINTER(dbl * 4), vp, vp, vp]
    lib.nrw_microbench.argtypes = [vp, i32, i32, i32, vp, vp, vp, vp]
    lib.nrw_p2p_allreduce_i64.argtypes = [vp, vp, vp, i64, i64, i64, i64, i32, i32, u64, vp, vp]
    lib.nrw_guided_phases.argtypes = [i64, i64, i64, i64, i64, vp, vp, vp, C.POINTER(i32)]
    lib.nrw_format_csv_e18.argtypes = [vp, i64, i64, vp, i64, C.POINTER(i64), i32]
    for name in EXPORTS:
        fn = getattr(lib, name)
        if name != "nrw_last_error":
            fn.restype = C.c_int
    _lib = lib
    return lib


def identity_weights_if_equivalent(cv, max_class):
    """Weights of classes no row carries cannot change any tally; when the weights of the classes
    in use equal the class index, hand the kernel the full identity vector [0, 1, 2, 3] -- it then
    reads the weight straight out of the row descriptor (the IDW variants of walk_score.cuh)."""
    cv = np.asarray(cv, dtype=np.float64).copy()
    k = int(max_class) + 1
    if np.array_equal(cv[:k], np.arange(k, dtype=np.float64)):
        cv[:] = np.arange(4, dtype=np.float64)
    return cv


def _ptr(t):
    return None if t is None else C.c_void_p(t.data_ptr())


@dataclass
class DeviceTables:
    """Device-resident transition tables (torch tensors keep the memory alive)."""
    csr: TransitionCSR
    indptr: object
    cols: object
    vals: object
    cum: object
    node_class: object
    alias_host: _tables.AliasTables | None
    rowinfo: object
    alias: object
    struct: _Tables
    max_class: int = 0          # largest source class any row carries (weights of unused classes are free)
    factored: tuple = ()        # device tensors of the factored row descriptors (tables.FactoredRows), if any
    tree: tuple = ()            # (thresholds uint8, tree_ptr int32) device tensors of the tree sampler, built on demand
    tree_host: tuple = ()

    @property
    def n_rows(self):
        return self.csr.n_rows

    @property
    def n_live(self):
        return self.csr.n_live


@dataclass
class WalkResult:
    """Device tensors written by one K1 launch (nothing here synchronises)."""
    counters: object
    score: object = None
    score_sq: object = None
    node_hist: object = None
    count_tn: object = None
    t_cap: int = 0
    kernel: str = ""

    def counters_dict(self) -> dict:
        c = self.counters.cpu().numpy().astype(np.uint64)
        return {k: int(c[i]) for i, k in enumerate(COUNTER_NAMES)}


@dataclass
class ReplayResult:
    ev_t: object
    ev_node: object
    ev_count: object
    ev_offsets: object          # int64 [n_starts + 1] (device)
    n_events: int
    draws_used: int


class PreparedWalk:
    """A K1 launch descriptor built once: ``pw()`` re-zeroes the tallies and launches again,
    ``pw.launch_kernel()`` launches without touching the tallies (what bench.py times)."""

    def __init__(self, engine, tables, wstruct, res):
        self.engine, self.tables, self.w, self.res = engine, tables, wstruct, res
        self.tallies = [t for t in (res.counters, res.score, res.score_sq, res.node_hist, res.count_tn)
                        if t is not None]

    def zero(self):
        for t in self.tallies:
            t.zero_()

    def set_seed(self, seed):
        """Same launch, other Philox key (a fresh, independent set of trajectories)."""
        self.w.seed = int(seed) & 0xFFFFFFFFFFFFFFFF

    def launch_kernel(self):
        e = self.engine
        e._check(e.lib.nrw_walk_philox(e.handle, C.byref(self.tables.struct), C.byref(self.w), e.stream))

    def __call__(self):
        self.zero()
        self.launch_kernel()
        return self.res


class Engine:
    """One engine per (process, device).  All launches go to torch's current stream.

    ``switches`` (measurement / test only): NRW_* environment switches this engine is
    created with -- the library reads them once, in nrw_create, so an A/B tool or a parity test that wants
    another code path creates another engine: ``Engine(0, switches={"NRW_DET": "0"})``."""

    def __init__(self, device=None, switches=None):
        import torch
        self.torch = torch
        self.lib = load_library()
        if not torch.cuda.is_available():
            raise NrwError("no CUDA device: the pc path of this build runs on a B200 only "
                           "(there is deliberately no CPU fallback)")
        if device is None:
            device = torch.cuda.current_device()
        self.device = torch.device("cuda", device if isinstance(device, int) else device.index)
        h = C.c_void_p()
        saved = {k: os.environ.get(k) for k in (switches or {})}
        try:
            for k, v in (switches or {}).items():
                if v is None:
                    os.environ.pop(k, None)
                else:
                    os.environ[k] = str(v)
            self._check(self.lib.nrw_create(self.device.index, C.byref(h)))
        finally:
            for k, v in saved.items():
                if v is None:
                    os.environ.pop(k, None)
                else:
                    os.environ[k] = v
        self.handle = h
        a, b = C.c_int(), C.c_int()
        self._check(self.lib.nrw_device_info(self.handle, C.byref(a), C.byref(b)))
        self.n_sms, self.max_smem = a.value, b.value
        buf = C.create_string_buffer(256)
        self._check(self.lib.nrw_engine_config(self.handle, buf, 256))
        self.config = buf.value.decode()

    def close(self):
        if getattr(self, "handle", None):
            self.lib.nrw_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ------------------------------------------------------------------ helpers
    def _check(self, rc):
        if rc != 0:
            raise NrwError("libnrw error %d: %s" % (rc, self.lib.nrw_last_error().decode()))

    @property
    def stream(self):
        return C.c_void_p(self.torch.cuda.current_stream(self.device).cuda_stream)

    def _dev(self, arr, dtype):
        a = np.ascontiguousarray(arr, dtype=dtype)
        if not a.flags.writeable:
            a = a.copy()
        return self.torch.from_numpy(a).to(self.device)

    # ------------------------------------------------------------------ tables
    def upload(self, csr: TransitionCSR, node_class=None, alias=True, alias_layout="auto") -> DeviceTables:
        """CSR -> device: indptr/cols/vals, cum by kernel K4, alias tables (host-built, de-duplicated)."""
        torch = self.torch
        if csr.nnz >= 2 ** 31:
            raise NrwError("nnz %d does not fit int32 indptr" % csr.nnz)
        indptr = self._dev(csr.indptr, np.int32)
        cols = self._dev(csr.cols if csr.nnz else np.zeros(1, np.int32), np.int32)
        vals = self._dev(csr.vals if csr.nnz else np.zeros(1), np.float64)
        cum = torch.empty_like(vals)
        self._check(self.lib.nrw_build_cum(self.handle, csr.n_rows, _ptr(indptr), _ptr(vals),
                                           _ptr(cum), self.stream))
        nc = None
        if node_class is not None:
            nc = self._dev(node_class, np.uint8)
        st = _Tables()
        st.n_rows, st.n_live = csr.n_rows, csr.n_live
        st.indptr, st.cols, st.cum = _ptr(indptr), _ptr(cols), _ptr(cum)
        st.node_class = _ptr(nc)
        at = rowinfo = alias_t = None
        keep_fac = ()
        st.alias_layout, st.n_alias = ALIAS_LAYOUT["none"], 0
        if alias:
            at = _tables.build_alias(csr, node_class=node_class, layout=alias_layout)
            if at.layout == "compact":
                # the compact layout is only useful when the patterns fit one SM's shared memory
                if 16 + ((at.n_alias * 8 + 15) & ~15) > self.max_smem:
                    at = _tables.build_alias(csr, node_class=node_class, layout="wide")
            rowinfo = self._dev(at.rowinfo, np.uint32)
            alias_t = self._dev(at.alias, np.uint32)
            st.alias_layout, st.n_alias = ALIAS_LAYOUT[at.layout], at.n_alias
            st.rowinfo, st.alias = _ptr(rowinfo), _ptr(alias_t)
            # meshes with the (direction, cell) structure of the fluence matrix: factored row descriptors,
            # used by the walk when the descriptor table itself does not fit shared memory
            fr = _tables.factorise_rows(at, csr.n_live) if at.layout == "compact" else None
            if fr is not None:
                fac = (self._dev(fr.dirtab, np.uint32), self._dev(fr.tail, np.uint32), self._dev(fr.cls_bits, np.uint32))
                st.fac_dir, st.fac_tail, st.fac_cls = (_ptr(x) for x in fac)
                st.fac_period, st.fac_cells, st.fac_n_tail = fr.period, fr.cells, fr.tail.shape[0]
                st.fac_shift, st.fac_magic = fr.shift, fr.magic
                st.fac_exit_w0, st.fac_exit_base = fr.exit_w0, fr.exit_base
                keep_fac = fac
            # structured form of the same tables (fluence matrices): the score walk then draws without alias look-ups
            sr = at.structured() if at.layout == "compact" else None
            if sr is not None:
                ri_str = self._dev(sr[0], np.uint32)
                st.rowinfo_str, st.str_thresh, st.str_n = _ptr(ri_str), int(sr[1]), int(sr[2])
                keep_fac = keep_fac + (ri_str,)
                fs = _tables.factorise_rows(at, csr.n_live, rowinfo=sr[0]) if fr is not None else None
                if fs is not None and (fs.period, fs.cells, fs.shift, fs.magic, fs.exit_base) == (
                        fr.period, fr.cells, fr.shift, fr.magic, fr.exit_base) and np.array_equal(fs.cls_bits, fr.cls_bits):
                    fac_s = (self._dev(fs.dirtab, np.uint32), self._dev(fs.tail, np.uint32))
                    st.fac_dir_str, st.fac_tail_str, st.fac_exit_w0_str = _ptr(fac_s[0]), _ptr(fac_s[1]), fs.exit_w0
                    keep_fac = keep_fac + fac_s
        max_class = 0 if node_class is None else int(np.asarray(node_class).max(initial=0))
        return DeviceTables(csr, indptr, cols, vals, cum, nc, at, rowinfo, alias_t, st, max_class, keep_fac)

    # ------------------------------------------------------------------ K1
    def _walk_struct(self, tables, start_nodes, start_counts, nt, seed, walker_range, sampler,
                     score, score_sq, class_value, time_weight, node_hist, count_tn, t_cap,
                     count_per_start, start_id_base=None, counters=None, count_path=False):
        torch = self.torch
        start_nodes = np.asarray(start_nodes, dtype=np.int64).reshape(-1)
        n_starts = len(start_nodes)
        if n_starts == 0:
            raise NrwError("empty start list")
        if start_nodes.min() < 0 or start_nodes.max() >= tables.n_rows:
            raise IndexError("start node out of range")
        counts = np.broadcast_to(np.asarray(start_counts, dtype=np.int64), (n_starts,))
        if counts.min() < 0:
            raise ValueError("negative walker count")
        prefix = np.zeros(n_starts + 1, dtype=np.int64)
        prefix[1:] = np.cumsum(counts)
        lo, hi = (0, int(prefix[-1])) if walker_range is None else (int(walker_range[0]), int(walker_range[1]))
        if not (0 <= lo <= hi <= int(prefix[-1])):
            raise ValueError("walker_range outside [0, total walkers]")
        keep = dict(start_nodes=self._dev(start_nodes, np.int32), prefix=self._dev(prefix, np.int64))
        w = _Walk()
        w.n_starts = n_starts
        w.start_nodes, w.start_prefix = _ptr(keep["start_nodes"]), _ptr(keep["prefix"])
        w.walker_lo, w.walker_hi = lo, hi
        if start_id_base is not None:
            idb = np.asarray(start_id_base, dtype=np.int64).reshape(-1)
            if idb.shape != (n_starts,) or (idb < 0).any() or (idb + counts).max() >= (1 << 40):
                raise ValueError("start_id_base must hold one id in [0, 2^40) per start")
            keep["id_base"] = self._dev(idb, np.int64)
            w.start_id_base = _ptr(keep["id_base"])
        w.seed = int(seed) & 0xFFFFFFFFFFFFFFFF
        w.nt, w.sampler = int(nt), SAMPLER[sampler]
        if counters is None:
            counters = torch.zeros(NRW_N_COUNTERS, dtype=torch.int64, device=self.device)
        elif (counters.dtype != torch.int64 or counters.numel() != NRW_N_COUNTERS or not counters.is_contiguous()
                or counters.device != self.device):
            raise ValueError("counters buffer must be a contiguous int64[%d] tensor on the engine's device"
                             % NRW_N_COUNTERS)
        res = WalkResult(counters=counters)
        w.counters = _ptr(res.counters)
        cv = np.zeros(4) if class_value is None else np.asarray(class_value, dtype=np.float64).copy()
        cv = identity_weights_if_equivalent(cv, tables.max_class)
        w.class_value = (C.c_double * 4)(*cv.tolist())
        if score is not None and score is not False:
            if score is True:
                res.score = torch.zeros(n_starts, dtype=torch.int64, device=self.device)
            else:                                   # caller's buffer (e.g. symmetric memory, nrw/dist.py)
                if (score.dtype != torch.int64 or score.numel() != n_starts or not score.is_contiguous()
                        or score.device != self.device):
                    raise ValueError("score buffer must be a contiguous int64[n_starts] tensor on the engine's device")
                res.score = score
            w.score = _ptr(res.score)
            if score_sq:
                res.score_sq = torch.zeros(n_starts, dtype=torch.float64, device=self.device)
                w.score_sq = _ptr(res.score_sq)
            if time_weight is not None:
                tw = np.asarray(time_weight, dtype=np.float64)
                if tw.shape != (int(nt),):
                    raise ValueError("time_weight must have nt entries")
                keep["tw"] = self._dev(tw, np.float64)
                w.time_weight = _ptr(keep["tw"])
        if node_hist:
            res.node_hist = torch.zeros(tables.n_rows, dtype=torch.int64, device=self.device)
            w.node_hist = _ptr(res.node_hist)
        if count_tn:
            t_cap = int(nt if t_cap is None else min(t_cap, nt))
            groups = n_starts if count_per_start else 1
            res.count_tn = torch.zeros((groups, t_cap, tables.n_rows), dtype=torch.int32, device=self.device)
            res.t_cap = t_cap
            w.count_tn, w.t_cap, w.count_per_start = _ptr(res.count_tn), t_cap, int(bool(count_per_start))
            if count_path and n_starts * t_cap <= (1 << 13):
                # most likely trajectory of every start (include/nrw.h, nrw_walk.count_path): walkers still on
                # it are tallied through a leave-time histogram instead of one contended atomic per step
                keep["path"] = self._dev(_tables.likely_paths(tables.csr, start_nodes, t_cap), np.int32)
                w.count_path = _ptr(keep["path"])
        return w, res, keep

    def _ensure_tree(self, tables: DeviceTables):
        """Tree-sampler tables (loihirw.rw_core probability circuit, 8-bit weights) -- built and uploaded on first use."""
        if not tables.tree:
            thr, ptr = _tables.build_tree_tables(tables.csr)
            tables.tree_host = (thr, ptr)
            tables.tree = (self._dev(thr if len(thr) else np.zeros(1, np.uint8), np.uint8), self._dev(ptr, np.int32))
            tables.struct.tree, tables.struct.tree_ptr = _ptr(tables.tree[0]), _ptr(tables.tree[1])

    def walk(self, tables: DeviceTables, start_nodes, start_counts, nt, seed, *, walker_range=None,
             sampler="alias", score=False, score_sq=False, class_value=None, time_weight=None,
             node_hist=False, count_tn=False, t_cap=None, count_per_start=False,
             start_id_base=None, count_path=False) -> WalkResult:
        """One K1 launch on the current stream.  Walker g (global id over the concatenated start
        list) starts on its start node; ``walker_range`` restricts the launch to ids [lo, hi)."""
        if sampler == "tree":
            self._ensure_tree(tables)
        w, res, keep = self._walk_struct(tables, start_nodes, start_counts, nt, seed, walker_range,
                                         sampler, score, score_sq, class_value, time_weight,
                                         node_hist, count_tn, t_cap, count_per_start, start_id_base,
                                         count_path=count_path)
        name = C.create_string_buffer(96)
        self._check(self.lib.nrw_walk_philox_plan(self.handle, C.byref(tables.struct), C.byref(w), name,
                                                  96, None, None, None))
        res.kernel = name.value.decode()
        self._check(self.lib.nrw_walk_philox(self.handle, C.byref(tables.struct), C.byref(w), self.stream))
        res._keep = keep            # start list / weights stay alive until the result is dropped
        return res

    def prepare_walk(self, tables, start_nodes, start_counts, nt, seed, **kw):
        """Pre-build the launch descriptor once; returns ``launch()`` that re-zeroes the tallies and
        launches K1 again (what bench.py times: no host-side table or descriptor work per step)."""
        args = dict(walker_range=None, sampler="alias", score=False, score_sq=False, class_value=None,
                    time_weight=None, node_hist=False, count_tn=False, t_cap=None, count_per_start=False,
                    start_id_base=None, counters=None, count_path=False)
        args.update(kw)
        w, res, keep = self._walk_struct(tables, start_nodes, start_counts, nt, seed, **args)
        name = C.create_string_buffer(96)
        grid, block, smem = C.c_int(), C.c_int(), C.c_int()
        self._check(self.lib.nrw_walk_philox_plan(self.handle, C.byref(tables.struct), C.byref(w), name,
                                                  96, C.byref(grid), C.byref(block), C.byref(smem)))
        res.kernel = name.value.decode()
        res._keep = keep
        res.launch_geometry = (grid.value, block.value, smem.value)
        return PreparedWalk(self, tables, w, res), res

    # ------------------------------------------------------------------ K2
    def replay(self, tables: DeviceTables, start_nodes, start_counts, nt, *, draw=None,
               rewind=None, chunk=None) -> ReplayResult:
        """Bit-exact walk on the reference's uniform stream.

        ``draw(n)`` returns the next n uniforms of the stream; the default draws from the global
        legacy ``np.random`` state exactly as the reference does (fluence_mini_app.py:243) and, when
        the walk is over, leaves that state advanced by exactly the number of uniforms the reference
        would have consumed -- so chained calls and later user code see the reference's stream
        position.  Start nodes are replayed in order on one stream."""
        torch = self.torch
        start_nodes = np.asarray(start_nodes, dtype=np.int64).reshape(-1)
        n_starts = len(start_nodes)
        counts = np.ascontiguousarray(np.broadcast_to(np.asarray(start_counts, dtype=np.int64), (n_starts,)))
        if n_starts == 0:
            raise NrwError("empty start list")
        if start_nodes.min() < 0 or start_nodes.max() >= tables.n_rows:
            raise IndexError("start node out of range")
        if counts.max(initial=0) >= 2 ** 32:
            raise ValueError("replay supports < 2^32 walkers per start")
        use_global = draw is None
        if use_global:
            saved = np.random.get_state()
            draw = lambda n: np.random.uniform(size=n)

        def advance(k):
            np.random.set_state(saved)
            while k > 0:
                m = min(k, 1 << 22)
                np.random.uniform(size=m)
                k -= m

        total = int(counts.sum())
        chunk = int(chunk) if chunk else int(min(max(1 << 16, 8 * total), 1 << 25))
        ev_cap = int(max(1 << 16, 16 * total + 2 * n_starts))
        d_nodes, d_counts = self._dev(start_nodes, np.int32), self._dev(counts, np.int64)
        while True:                                   # restarted only on event-capacity overflow
            ev_t = torch.empty(ev_cap, dtype=torch.int32, device=self.device)
            ev_node = torch.empty(ev_cap, dtype=torch.int32, device=self.device)
            ev_count = torch.empty(ev_cap, dtype=torch.int64, device=self.device)
            ev_off = torch.zeros(n_starts + 1, dtype=torch.int64, device=self.device)
            scratch = torch.zeros(3 * (tables.n_rows + 1), dtype=torch.int32, device=self.device)
            state = torch.zeros(8, dtype=torch.int64, device=self.device)
            r = _Replay()
            r.n_starts, r.start_nodes, r.start_counts, r.nt = n_starts, _ptr(d_nodes), _ptr(d_counts), int(nt)
            r.ev_t, r.ev_node, r.ev_count, r.ev_cap = _ptr(ev_t), _ptr(ev_node), _ptr(ev_count), ev_cap
            r.ev_offsets, r.scratch, r.state = _ptr(ev_off), _ptr(scratch), _ptr(state)
            base = 0                                  # stream index of host_buf[0]
            host_buf = np.zeros(0)
            generated = 0
            this_chunk = chunk
            while True:
                fresh = draw(this_chunk)
                generated += this_chunk
                host_buf = np.concatenate([host_buf, fresh]) if len(host_buf) else fresh
                d_u = self._dev(host_buf, np.float64)
                r.u_stream, r.u_len, r.u_base = _ptr(d_u), len(host_buf), base
                self._check(self.lib.nrw_walk_replay(self.handle, C.byref(tables.struct), C.byref(r), self.stream))
                st = state.cpu().numpy()
                status, pos = int(st[3]), int(st[0])
                if status != 1:
                    break
                host_buf = host_buf[pos - base:]      # keep the unused tail, extend the stream
                base = pos
                this_chunk = min(this_chunk * 2, 1 << 26)
            if status == 0:
                break
            ev_cap *= 4
            if use_global:
                np.random.set_state(saved)
            elif rewind is not None:
                rewind()
            else:
                raise NrwError("event capacity exceeded and the custom draw() cannot be rewound")
        if use_global:
            advance(pos)
        return ReplayResult(ev_t, ev_node, ev_count, ev_off, int(st[4]), pos)

    # ------------------------------------------------------------------ K3 / K5
    def assemble_events(self, rep: ReplayResult, time_weight, src, v, dt, start_counts, nt, n_live):
        """K3: rwSol value per start node from replay events, in the reference's (t, k) order."""
        torch = self.torch
        n_starts = rep.ev_offsets.numel() - 1
        tw = self._dev(time_weight, np.float64)
        sv = self._dev(src, np.float64)
        cnt = self._dev(np.broadcast_to(np.asarray(start_counts, dtype=np.int64), (n_starts,)), np.int64)
        out = torch.empty(n_starts, dtype=torch.float64, device=self.device)
        self._check(self.lib.nrw_assemble_events(
            self.handle, n_starts, _ptr(rep.ev_offsets), _ptr(rep.ev_t), _ptr(rep.ev_node),
            _ptr(rep.ev_count), _ptr(tw), _ptr(sv), float(v), float(dt), _ptr(cnt), int(nt),
            int(n_live), _ptr(out), self.stream))
        return out

    def expand_events(self, rep: ReplayResult, start_index, n_rows, nt):
        """K5a: dense float64 tempdata[n_rows, nt] of one start from the event list."""
        torch = self.torch
        off = rep.ev_offsets.cpu().numpy()
        lo, hi = int(off[start_index]), int(off[start_index + 1])
        out = torch.empty((n_rows, nt), dtype=torch.float64, device=self.device)
        self._check(self.lib.nrw_expand_events(
            self.handle, hi - lo, _ptr(rep.ev_t[lo:]) if hi > lo else None,
            _ptr(rep.ev_node[lo:]) if hi > lo else None, _ptr(rep.ev_count[lo:]) if hi > lo else None,
            int(n_rows), int(nt), _ptr(out), self.stream))
        return out

    def expand_counts(self, count_tn, t_cap, n_rows, nt, start_node, start_count):
        """K5b: dense uint32 [t_cap, n_rows] counts -> float64 tempdata[n_rows, nt]."""
        out = self.torch.empty((n_rows, nt), dtype=self.torch.float64, device=self.device)
        self._check(self.lib.nrw_expand_counts(self.handle, _ptr(count_tn), int(t_cap), int(n_rows),
                                               int(nt), int(start_node), int(start_count), _ptr(out),
                                               self.stream))
        return out

    # ------------------------------------------------------------------ roofline denominators
    MICROBENCH_KINDS = {"smem_atomic": 0, "smem_load": 1, "l2_load": 2, "global_atomic": 3}

    def microbench(self, kind: str, n: int, iters: int = 2000, reps: int = 5) -> float:
        """Operations per second of one resource at the walk's launch geometry (see include/nrw.h,
        nrw_microbench): best of ``reps`` event-timed launches."""
        torch = self.torch
        k = self.MICROBENCH_KINDS[kind]
        table = hist = None
        if kind == "l2_load":
            g = torch.Generator(device="cpu").manual_seed(1)
            table = torch.randint(0, 2 ** 31 - 1, (int(n), 2), dtype=torch.int32, generator=g).to(self.device)
        if kind == "global_atomic":
            hist = torch.zeros(int(n), dtype=torch.int64, device=self.device)
        scratch = torch.zeros(1, dtype=torch.int64, device=self.device)
        best = float("inf")
        for _ in range(reps + 1):                              # first launch = warm-up
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            self._check(self.lib.nrw_microbench(self.handle, k, int(n), int(iters), _ptr(table), _ptr(hist),
                                                _ptr(scratch), self.stream))
            b.record()
            b.synchronize()
            best = min(best, a.elapsed_time(b))
        return self.n_sms * 1024 * float(iters) / (best * 1e-3)

    # ------------------------------------------------------------------ host-buffer end-to-end
    def fluence_scores_host(self, alias_tables, n_rows, n_live, start_nodes, start_prefix, walker_lo,
                            walker_hi, seed, nt, class_value, score_out, counters_out):
        """nrw_fluence_scores_host: numpy (ideally pinned) buffers in, numpy buffers out, synchronous."""
        self.prepare_fluence_scores_host(alias_tables, n_rows, n_live, start_nodes, start_prefix, walker_lo,
                                         walker_hi, nt, class_value, score_out, counters_out)(seed)

    def prepare_fluence_scores_host(self, alias_tables, n_rows, n_live, start_nodes, start_prefix, walker_lo,
                                    walker_hi, nt, class_value, score_out, counters_out, start_id_base=None,
                                    exchange=None):
        """The same call with its arguments converted once: returns ``solve(seed)`` for a loop of solves
        over the same buffers (a CUDA-graph replay of the copy / walk / copy-back sequence was measured
        and gave nothing: the per-call cost beyond the kernel, ~50 us, is DMA and synchronisation latency).
        ``exchange`` (nrw.dist.TallyExchange, N > 1): the walk tallies into the exchange's buffer and sums it
        over ranks in its own tail; the reduced scores come back to ``score_out``."""
        if alias_tables.layout != "compact":
            raise NrwError("the host entry takes the compact alias layout")
        if exchange is not None and not exchange.fused:
            raise NrwError("the host entry needs the fused exchange (peer memory on every rank)")
        cv = (C.c_double * 4)(*[float(x) for x in class_value])
        keep = (alias_tables.rowinfo, alias_tables.alias, start_nodes, start_prefix, score_out, counters_out, cv,
                start_id_base, exchange)
        fn, check, handle, stream = self.lib.nrw_fluence_scores_host, self._check, self.handle, self.stream
        # the structured form of the tables, when they have one (`rowinfo_str` / `str_thresh` / `str_n` attributes of
        # the view, or AliasTables.structured()): the walk then needs half the table bytes on the device
        ri_str, thr = getattr(alias_tables, "rowinfo_str", None), int(getattr(alias_tables, "str_thresh", 0))
        n_uni = int(getattr(alias_tables, "str_n", 0))
        if ri_str is None and hasattr(alias_tables, "structured") and alias_tables.structured() is not None:
            ri_str, thr, n_uni = alias_tables.structured()
        keep = keep + (ri_str,)
        head = (handle, int(n_rows), int(n_live), int(alias_tables.n_alias),
                C.c_void_p(alias_tables.rowinfo.ctypes.data), C.c_void_p(alias_tables.alias.ctypes.data),
                C.c_void_p(ri_str.ctypes.data) if ri_str is not None else None, thr, n_uni,
                len(start_nodes), C.c_void_p(start_nodes.ctypes.data), C.c_void_p(start_prefix.ctypes.data),
                C.c_void_p(start_id_base.ctypes.data) if start_id_base is not None else None,
                int(walker_lo), int(walker_hi))
        mid = (int(nt), C.byref(cv), C.c_void_p(score_out.ctypes.data),
               C.c_void_p(counters_out.ctypes.data) if counters_out is not None else None)

        def solve(seed):
            red = exchange.arm()[1] if exchange is not None else None
            check(fn(*head, int(seed) & 0xFFFFFFFFFFFFFFFF, *mid, red, stream))
        solve.keepalive = keep
        return solve


_engines: dict = {}


def get_engine(device=None) -> Engine:
    """Process-wide engine per device (the handle the drop-in entry points share)."""
    import torch
    if not torch.cuda.is_available():
        raise NrwError("no CUDA device: run_mode 'pc' of this build needs a B200 "
                       "(there is deliberately no CPU fallback)")
    idx = torch.cuda.current_device() if device is None else int(device)
    if idx not in _engines:
        _engines[idx] = Engine(idx)
    return _engines[idx]
