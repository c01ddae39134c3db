"""ctypes binding of the C-ABI in include/sietill_b200.h (libsietill_b200.so).

This is plumbing for tests and bench.py; the product is the shared library.  There is no CPU fallback:
if the library is missing, or no B200 is present, calls fail loudly.
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libsietill_b200.so")

T_NAMES = ["h2d", "score", "viterbi", "rescore", "sort", "reduce", "allreduce", "d2h", "score_filter", "score_refine", "mstep"]
TDP_SOURCE, TDP_DESTINATION = 0, 1

ALI_DTYPE = np.dtype([("count", np.uint16), ("state", np.uint16), ("weight", np.float32)])  # Types.hpp:29-38

# every symbol include/sietill_b200.h declares (checked by tests/test_capi_symbols.py)
SYMBOLS = [
    "sgmm_create", "sgmm_destroy", "sgmm_last_error", "sgmm_version", "sgmm_set_model", "sgmm_set_features",
    "sgmm_set_corpus", "sgmm_set_tdp", "sgmm_set_beam", "sgmm_set_lexicon", "sgmm_recognize", "sgmm_score_all", "sgmm_align", "sgmm_set_alignment", "sgmm_accumulate",
    "sgmm_accumulate_soft", "sgmm_process_features", "sgmm_align_accumulate",
    "sgmm_am_score_sum", "sgmm_last_frame_results", "sgmm_comm_unique_id", "sgmm_comm_init", "sgmm_stage_ms",
    "sgmm_launch_count", "sgmm_timer_start", "sgmm_timer_stop", "sgmm_host_alloc", "sgmm_host_free",
    "sgmm_set_option", "sgmm_get_counter", "sgmm_mstep", "sgmm_get_model", "sgmm_get_statistics", "sgmm_debug_table", "sgmm_score_sum_static",
]

_lib = None


def load():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(make -C speechrecognition_b200/csrc).  There is no CPU fallback.")
        L = C.CDLL(LIB_PATH)
        L.sgmm_last_error.restype = C.c_char_p
        L.sgmm_version.restype = C.c_char_p
        L.sgmm_launch_count.restype = C.c_uint64
        L.sgmm_last_error.argtypes = [C.c_void_p]
        L.sgmm_destroy.argtypes = [C.c_void_p]
        L.sgmm_launch_count.argtypes = [C.c_void_p]
        L.sgmm_host_alloc.restype = C.c_void_p
        L.sgmm_host_alloc.argtypes = [C.c_size_t]
        L.sgmm_host_free.argtypes = [C.c_void_p]
        _lib = L
    return _lib


def _ptr(a, ctype):
    return None if a is None else a.ctypes.data_as(C.POINTER(ctype))


class SgmmError(RuntimeError):
    pass


class WordLists:
    """The recognised word sequences of a batch: ``hyp[u]`` is utterance u's word-id array (a view into one buffer)."""

    def __init__(self, words, off):
        self.words, self.off = words, off

    def __len__(self):
        return len(self.off) - 1

    def __getitem__(self, u):
        if isinstance(u, slice):
            return [self[i] for i in range(*u.indices(len(self)))]
        if u < 0:
            u += len(self)
        if not 0 <= u < len(self):
            raise IndexError(u)
        return self.words[int(self.off[u]):int(self.off[u + 1])]

    def __iter__(self):
        return (self[u] for u in range(len(self)))


class Context:
    """One sgmm_ctx (one GPU).  All arrays are numpy host arrays, as the C-ABI takes host buffers."""

    def __init__(self, device=0):
        self.L = load()
        h = C.c_void_p()
        if self.L.sgmm_create(int(device), C.byref(h)):
            raise SgmmError(self.L.sgmm_last_error(None).decode())
        self.h = h
        self.dim = self.n_states = self.D = 0
        self.n_frames = self.n_utts = 0

    def close(self):
        if getattr(self, "h", None):
            self.L.sgmm_destroy(self.h)
            self.h = None

    __del__ = close

    def _ck(self, rc):
        if rc:
            raise SgmmError(self.L.sgmm_last_error(self.h).decode())

    # ---- setup
    def set_model(self, dim, state_off, mean, inv_var, norm, log_weight):
        state_off = np.ascontiguousarray(state_off, np.uint32)
        mean = np.ascontiguousarray(mean, np.float64).reshape(-1)
        inv_var = np.ascontiguousarray(inv_var, np.float64).reshape(-1)
        norm = np.ascontiguousarray(norm, np.float64)
        log_weight = np.ascontiguousarray(log_weight, np.float64)
        n = len(state_off) - 1
        D = int(state_off[-1])
        assert mean.size == D * dim and inv_var.size == D * dim and norm.size == D and log_weight.size == D
        self._ck(self.L.sgmm_set_model(self.h, C.c_uint32(dim), C.c_uint32(n), _ptr(state_off, C.c_uint32),
                                       _ptr(mean, C.c_double), _ptr(inv_var, C.c_double), _ptr(norm, C.c_double),
                                       _ptr(log_weight, C.c_double)))
        self.dim, self.n_states, self.D = dim, n, D

    def set_model_tables(self, dim, t):
        """From a reference-layout table dict (see oracle.binding / host MixtureModel): resolves
        mixtures_[s][k] -> (mean_idx, var_idx) into the state-major packed tables of sgmm_set_model."""
        mi, vi = t["mean_idx"].astype(np.int64), t["var_idx"].astype(np.int64)
        self.set_model(dim, t["mix_offsets"], t["means"].reshape(-1, dim)[mi], t["vars_inv"].reshape(-1, dim)[vi],
                       t["norm"][vi], t["mean_wlog"][mi])

    POOLING = {"none": 0, "mixture": 1, "global": 2}

    def mstep(self, pooling, var_seed=1e-4):
        """MixtureModel::finalize on the device from the resident statistics; the new model replaces the current one."""
        self._ck(self.L.sgmm_mstep(self.h, int(self.POOLING.get(pooling, pooling)), C.c_double(var_seed)))

    def get_model(self, want_var=False):
        D, d = self.D, self.dim
        mean, inv = np.empty((D, d), np.float64), np.empty((D, d), np.float64)
        norm, logw = np.empty(D, np.float64), np.empty(D, np.float64)
        var = np.empty((D, d), np.float64) if want_var else None
        w = np.empty(D, np.float64) if want_var else None
        self._ck(self.L.sgmm_get_model(self.h, _ptr(mean, C.c_double), _ptr(inv, C.c_double), _ptr(norm, C.c_double),
                                       _ptr(logw, C.c_double), _ptr(var, C.c_double), _ptr(w, C.c_double)))
        return {"mean": mean, "inv_var": inv, "norm": norm, "log_weight": logw, "var": var, "weight": w}

    def get_statistics(self):
        D, d = self.D, self.dim
        cnt, sx, sxx = np.empty(D, np.float64), np.empty((D, d), np.float64), np.empty((D, d), np.float64)
        ssum = np.zeros(1, np.float64)
        self._ck(self.L.sgmm_get_statistics(self.h, _ptr(cnt, C.c_double), _ptr(sx, C.c_double), _ptr(sxx, C.c_double),
                                            _ptr(ssum, C.c_double)))
        return cnt, sx, sxx, float(ssum[0])

    def debug_table(self, name):
        """Raw bytes of one derived model table (tests: device builder vs host builder)."""
        n = C.c_size_t(0)
        self._ck(self.L.sgmm_debug_table(self.h, name.encode(), None, C.c_size_t(0), C.byref(n)))
        out = np.empty(n.value, np.uint8)
        if n.value:
            self._ck(self.L.sgmm_debug_table(self.h, name.encode(), out.ctypes.data_as(C.c_void_p), C.c_size_t(n.value), C.byref(n)))
        return out

    def set_features(self, x):
        x = np.ascontiguousarray(x, np.float32)
        assert x.ndim == 2
        self._ck(self.L.sgmm_set_features(self.h, _ptr(x, C.c_float), C.c_uint64(x.shape[0]), C.c_uint32(x.shape[1])))
        self.n_frames = x.shape[0]

    def process_features(self, raw, frame_off, n_first=12, n_second=1, deriv_step=3, mean=None, stddev=None,
                         energy_max_norm=True, want=True):
        """SignalAnalysis::process_features on the device; the result becomes the resident feature matrix."""
        raw = np.ascontiguousarray(raw, np.float32)
        assert raw.ndim == 2
        fo = np.ascontiguousarray(frame_off, np.uint64)
        nt = raw.shape[1] + n_first + n_second
        mean = None if mean is None else np.ascontiguousarray(mean, np.float64)
        stddev = None if stddev is None else np.ascontiguousarray(stddev, np.float64)
        out = np.empty((raw.shape[0], nt), np.float32) if want else None
        self._ck(self.L.sgmm_process_features(self.h, _ptr(raw, C.c_float), C.c_uint64(raw.shape[0]),
                                              C.c_uint32(raw.shape[1]), C.c_uint32(n_first), C.c_uint32(n_second),
                                              C.c_uint32(deriv_step), C.c_uint32(len(fo) - 1), _ptr(fo, C.c_uint64),
                                              _ptr(mean, C.c_double), _ptr(stddev, C.c_double), int(energy_max_norm),
                                              _ptr(out, C.c_float)))
        self.n_frames = raw.shape[0]
        return out

    def set_corpus(self, frame_off, automata, aut_off):
        frame_off = np.ascontiguousarray(frame_off, np.uint64)
        automata = np.ascontiguousarray(automata, np.uint32)
        aut_off = np.ascontiguousarray(aut_off, np.uint32)
        self._ck(self.L.sgmm_set_corpus(self.h, C.c_uint32(len(frame_off) - 1), _ptr(frame_off, C.c_uint64),
                                        _ptr(automata, C.c_uint32), _ptr(aut_off, C.c_uint32)))
        self.n_utts = len(frame_off) - 1

    def set_tdp(self, loop, forward, skip, silence_state=0):
        self._ck(self.L.sgmm_set_tdp(self.h, C.c_double(loop), C.c_double(forward), C.c_double(skip),
                                     C.c_uint32(silence_state)))

    def set_beam(self, pruning_threshold=float("inf")):
        self._ck(self.L.sgmm_set_beam(self.h, C.c_double(pruning_threshold)))

    # ---- compute
    def score_all(self, f0=0, f1=None, want_argmin=True):
        f1 = self.n_frames if f1 is None else f1
        scores = np.empty((f1 - f0, self.n_states), np.float64)
        argmin = np.empty((f1 - f0, self.n_states), np.uint32) if want_argmin else None
        self._ck(self.L.sgmm_score_all(self.h, C.c_uint64(f0), C.c_uint64(f1), _ptr(scores, C.c_double),
                                       _ptr(argmin, C.c_uint32)))
        return scores, argmin

    def align(self, convention=TDP_SOURCE, want_alignment=True, want_costs=True, out=None, out_costs=None):
        ali = out if out is not None else (np.empty(self.n_frames, ALI_DTYPE) if want_alignment else None)
        costs = out_costs if out_costs is not None else (np.empty(self.n_utts, np.float64) if want_costs else None)
        self._ck(self.L.sgmm_align(self.h, int(convention), _ptr(ali, C.c_uint8) if ali is not None else None,
                                   _ptr(costs, C.c_double)))
        return ali, costs

    def set_lexicon(self, automata, silence_word=0):
        """automata: one array of state ids per word (positions of its linear HMM), e.g. synth.Lexicon().automata."""
        word_off = np.concatenate([[0], np.cumsum([len(a) for a in automata])]).astype(np.uint32)
        word_states = np.ascontiguousarray(np.concatenate(automata), np.uint32)
        self._ck(self.L.sgmm_set_lexicon(self.h, C.c_uint32(len(automata)), _ptr(word_off, C.c_uint32),
                                         _ptr(word_states, C.c_uint32), C.c_uint32(silence_word)))

    def recognize(self, frame_off, word_penalty=10.0, am_threshold=20.0):
        """Word-loop recognition of every utterance of the resident features -> (list of word-id arrays, scores)."""
        frame_off = np.ascontiguousarray(frame_off, np.uint64)
        U = len(frame_off) - 1
        words = np.empty(int(frame_off[-1]), np.uint32)      # upper bound (one word end per frame); only the words are written
        off = np.zeros(U + 1, np.uint32)
        scores = np.zeros(U, np.float64)
        self._ck(self.L.sgmm_recognize(self.h, C.c_uint32(U), _ptr(frame_off, C.c_uint64), C.c_double(word_penalty),
                                       C.c_double(am_threshold), _ptr(words, C.c_uint32), C.c_uint64(len(words)),
                                       _ptr(off, C.c_uint32), _ptr(scores, C.c_double)))
        return WordLists(words[:int(off[U])], off), scores

    def set_alignment(self, states):
        if states.dtype != ALI_DTYPE:
            a = np.zeros(len(states), ALI_DTYPE)
            a["state"] = states
            a["count"] = 1
            a["weight"] = 1.0
            states = a
        states = np.ascontiguousarray(states)
        self._ck(self.L.sgmm_set_alignment(self.h, _ptr(states, C.c_uint8), C.c_uint64(len(states))))

    def accumulate(self, first_pass=False, want=True, out=None):
        """out = (count[D], sum_x[D, d], sum_xx[D, d]) f64 arrays to fill (e.g. pinned_empty buffers)."""
        D, d = self.D, self.dim
        if out is not None:
            cnt, sx, sxx = out
            assert cnt.shape == (D,) and sx.shape == (D, d) and sxx.shape == (D, d)
            assert all(a.dtype == np.float64 and a.flags.c_contiguous for a in out)
        else:
            cnt = np.empty(D, np.float64) if want else None
            sx = np.empty((D, d), np.float64) if want else None
            sxx = np.empty((D, d), np.float64) if want else None
        ssum = np.zeros(1, np.float64)
        self._ck(self.L.sgmm_accumulate(self.h, int(first_pass), _ptr(cnt, C.c_double), _ptr(sx, C.c_double),
                                        _ptr(sxx, C.c_double), _ptr(ssum, C.c_double)))
        return cnt, sx, sxx, float(ssum[0])

    def align_accumulate(self, out_ali, out_costs, out_stats, convention=TDP_SOURCE):
        """sgmm_align + sgmm_accumulate(first_pass=False) in one call; all outputs into caller buffers."""
        cnt, sx, sxx = out_stats
        ssum = np.zeros(1, np.float64)
        self._ck(self.L.sgmm_align_accumulate(self.h, int(convention), out_ali.ctypes.data_as(C.c_void_p),
                                              _ptr(out_costs, C.c_double), _ptr(cnt, C.c_double), _ptr(sx, C.c_double),
                                              _ptr(sxx, C.c_double), _ptr(ssum, C.c_double)))
        return cnt, sx, sxx, float(ssum[0])

    def accumulate_soft(self):
        """Posterior-weighted E-step (Mixtures.cpp:300-357): (weight[D], sum_x[D, d], sum_xx[D, d], sum of sum scores)."""
        D, d = self.D, self.dim
        cnt = np.empty(D, np.float64)
        sx = np.empty((D, d), np.float64)
        sxx = np.empty((D, d), np.float64)
        ssum = np.zeros(1, np.float64)
        self._ck(self.L.sgmm_accumulate_soft(self.h, _ptr(cnt, C.c_double), _ptr(sx, C.c_double),
                                             _ptr(sxx, C.c_double), _ptr(ssum, C.c_double)))
        return cnt, sx, sxx, float(ssum[0])

    def score_sum_static(self, d_values_ptr, n, d_sum_ptr, stream_ptr=0):
        """sgmm_score_sum_static: DEVICE pointers (ints), enqueue only — the caller synchronises `stream`."""
        self._ck(self.L.sgmm_score_sum_static(self.h, C.c_void_p(d_values_ptr), C.c_uint64(n), C.c_void_p(d_sum_ptr),
                                              C.c_void_p(stream_ptr)))

    def am_score_sum(self):
        ssum = np.zeros(1, np.float64)
        self._ck(self.L.sgmm_am_score_sum(self.h, _ptr(ssum, C.c_double)))
        return float(ssum[0])

    def last_frame_results(self):
        dens = np.empty(self.n_frames, np.uint32)
        score = np.empty(self.n_frames, np.float64)
        self._ck(self.L.sgmm_last_frame_results(self.h, _ptr(dens, C.c_uint32), _ptr(score, C.c_double)))
        return dens, score

    # ---- multi-GPU
    @staticmethod
    def comm_unique_id():
        buf = np.zeros(128, np.uint8)
        if load().sgmm_comm_unique_id(_ptr(buf, C.c_uint8), C.c_size_t(128)):
            raise SgmmError(load().sgmm_last_error(None).decode())
        return buf

    def comm_init(self, rank, world, uid):
        uid = np.ascontiguousarray(uid, np.uint8)
        self._ck(self.L.sgmm_comm_init(self.h, int(rank), int(world), _ptr(uid, C.c_uint8), C.c_size_t(128)))

    # ---- instrumentation
    def stage_ms(self):
        ms = np.zeros(len(T_NAMES), np.float32)
        self._ck(self.L.sgmm_stage_ms(self.h, _ptr(ms, C.c_float)))
        return dict(zip(T_NAMES, (float(v) for v in ms)))

    def timer_start(self):
        self._ck(self.L.sgmm_timer_start(self.h))

    def timer_stop(self):
        ms = C.c_float(0)
        self._ck(self.L.sgmm_timer_stop(self.h, C.byref(ms)))
        return float(ms.value)

    def set_option(self, key, value):
        self._ck(self.L.sgmm_set_option(self.h, key.encode(), C.c_int64(int(value))))

    def get_counter(self, key):
        v = C.c_int64(0)
        self._ck(self.L.sgmm_get_counter(self.h, key.encode(), C.byref(v)))
        return int(v.value)

    def launch_count(self):
        return int(self.L.sgmm_launch_count(self.h))


_pinned_keepalive = []


def pinned_empty(shape, dtype):
    """numpy array over cudaHostAlloc'ed (page-locked) memory; lives until the process exits."""
    dtype = np.dtype(dtype)
    n = max(int(np.prod(shape)) * dtype.itemsize, 1)
    p = load().sgmm_host_alloc(C.c_size_t(n))
    if not p:
        raise SgmmError("cudaHostAlloc failed")
    buf = (C.c_uint8 * n).from_address(p)
    _pinned_keepalive.append(buf)
    return np.frombuffer(buf, dtype=np.uint8, count=int(np.prod(shape)) * dtype.itemsize).view(dtype).reshape(shape)
