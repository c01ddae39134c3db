"""TEST INFRASTRUCTURE — ctypes bindings for the two checkers.

* ``Oracle``    -> oracle/libsietill_oracle.so  (plain-C restatement, oracle/sietill_oracle.c)
* ``Reference`` -> oracle/_ref/libsietill_ref.so (the UNMODIFIED reference sources + oracle/ref_harness.cpp)

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this.
Both classes expose the same methods and the same model-table dictionary so tests can diff them.
"""
import ctypes as C
import json
import os
import subprocess
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ORACLE_SO = os.path.join(HERE, "libsietill_oracle.so")
REF_SO = os.path.join(HERE, "_ref", "libsietill_ref.so")
REF_BIN = os.path.join(HERE, "_ref", "sietill_ref")

POOLING = {"none": 0, "mixture": 1, "global": 2}
TABLES = ["means", "mean_acc", "mean_w", "mean_wlog", "mean_wacc", "mean_refs", "vars", "vars_inv",
          "var_acc", "var_wacc", "var_refs", "norm", "mix_offsets", "mean_idx", "var_idx"]


def build_port():
    subprocess.check_call(["make", "-s", "-C", HERE, "port"])


def build_ref():
    subprocess.check_call(["make", "-s", "-C", HERE, "ref"])


def have_ref():
    return os.path.exists(REF_SO)


def _p(a, t):
    return a.ctypes.data_as(C.POINTER(t)) if a is not None else None


def _f32(x):
    return np.ascontiguousarray(x, dtype=np.float32)


def _alloc_tables(sizes):
    d, n, M, V, D = [int(v) for v in sizes]
    f8, u8, u4 = np.float64, np.uint64, np.uint32
    return {
        "means": np.zeros(M * d, f8), "mean_acc": np.zeros(M * d, f8), "mean_w": np.zeros(M, f8),
        "mean_wlog": np.zeros(M, f8), "mean_wacc": np.zeros(M, f8), "mean_refs": np.zeros(M, u8),
        "vars": np.zeros(V * d, f8), "vars_inv": np.zeros(V * d, f8), "var_acc": np.zeros(V * d, f8),
        "var_wacc": np.zeros(V, f8), "var_refs": np.zeros(V, u8), "norm": np.zeros(V, f8),
        "mix_offsets": np.zeros(n + 1, u4), "mean_idx": np.zeros(D, u4), "var_idx": np.zeros(D, u4),
    }


def _table_args(t):
    ct = {np.dtype(np.float64): C.c_double, np.dtype(np.uint64): C.c_uint64, np.dtype(np.uint32): C.c_uint32}
    return [_p(t[k], ct[t[k].dtype]) for k in TABLES]


class Oracle:
    """The C restatement.  States / automata / argmins are uint32 arrays."""

    def __init__(self, dim, n_mix, pooling="none", max_approx=True, tdp=(0.0, 0.0, 0.0), silence=0):
        if not os.path.exists(ORACLE_SO):
            build_port()
        L = self.L = C.CDLL(ORACLE_SO)
        L.orc_model_create.restype = C.c_void_p
        for fn in ("orc_density_score", "orc_min_score", "orc_sum_score", "orc_score", "orc_am_score",
                   "orc_align_full", "orc_align_pruned", "orc_align_batch", "orc_tdp_score"):
            getattr(L, fn).restype = C.c_double
        L.orc_num_densities.restype = C.c_uint64
        self.dim, self.n_mix, self.pooling = dim, n_mix, POOLING[pooling]
        self.max_approx, self.tdp, self.silence = int(max_approx), tuple(float(v) for v in tdp), silence
        self.h = C.c_void_p(L.orc_model_create(dim, n_mix, self.pooling))

    def __del__(self):
        if getattr(self, "h", None):
            self.L.orc_model_destroy(self.h)
            self.h = None

    def sizes(self):
        out = np.zeros(5, np.uint64)
        self.L.orc_model_sizes(self.h, _p(out, C.c_uint64))
        return out

    def tables(self):
        t = _alloc_tables(self.sizes())
        self.L.orc_model_export(self.h, *_table_args(t))
        return t

    def load_tables(self, t):
        t = {k: np.ascontiguousarray(v) for k, v in t.items()}
        self.L.orc_model_import(self.h, C.c_uint32(len(t["mean_w"])), C.c_uint32(len(t["var_wacc"])),
                                *_table_args(t))

    def accumulate(self, states, x, first_pass=False, max_approx=None):
        states = np.ascontiguousarray(states, np.uint32)
        x = _f32(x)
        argmin = np.zeros(len(states), np.uint32)
        ma = self.max_approx if max_approx is None else int(max_approx)
        self.L.orc_accumulate(self.h, _p(states, C.c_uint32), _p(x, C.c_float), C.c_uint64(len(states)),
                              int(first_pass), ma, _p(argmin, C.c_uint32))
        return argmin

    def finalize(self):
        self.L.orc_finalize(self.h)

    def split(self, min_obs):
        self.L.orc_split(self.h, C.c_uint64(int(min_obs)))

    def eliminate(self, min_obs):
        self.L.orc_eliminate(self.h, C.c_double(float(min_obs)))

    def num_densities(self):
        return int(self.L.orc_num_densities(self.h))

    def min_score(self, x, states=None):
        x = _f32(x).reshape(-1, self.dim)
        n = self.n_mix if states is None else len(states)
        st = None if states is None else np.ascontiguousarray(states, np.uint32)
        scores = np.zeros((len(x), n), np.float64)
        argmin = np.zeros((len(x), n), np.uint32)
        self.L.orc_min_score_batch(self.h, _p(x, C.c_float), C.c_uint64(len(x)), _p(st, C.c_uint32),
                                   C.c_uint64(n), _p(scores, C.c_double), _p(argmin, C.c_uint32))
        return scores, argmin

    def score(self, x):
        """FeatureScorer::score of every (frame, state): min or sum score per self.max_approx."""
        x = _f32(x).reshape(-1, self.dim)
        scores = np.zeros((len(x), self.n_mix), np.float64)
        self.L.orc_score_batch(self.h, _p(x, C.c_float), C.c_uint64(len(x)), self.max_approx, _p(scores, C.c_double))
        return scores

    def am_score(self, states, x):
        states = np.ascontiguousarray(states, np.uint32)
        x = _f32(x)
        return float(self.L.orc_am_score(self.h, _p(states, C.c_uint32), _p(x, C.c_float),
                                         C.c_uint64(len(states)), self.max_approx))

    def _tdp_args(self):
        return [C.c_double(self.tdp[0]), C.c_double(self.tdp[1]), C.c_double(self.tdp[2]),
                C.c_uint32(self.silence), self.max_approx]

    def align_full(self, x, automaton, want_margin=False):
        x = _f32(x).reshape(-1, self.dim)
        aut = np.ascontiguousarray(automaton, np.uint32)
        T = len(x)
        states, pos = np.zeros(T, np.uint32), np.zeros(T, np.uint32)
        margin = np.zeros(T, np.float64) if want_margin else None
        cost = self.L.orc_align_full(self.h, *self._tdp_args(), _p(x, C.c_float), C.c_uint64(T),
                                     _p(aut, C.c_uint32), C.c_uint64(len(aut)), _p(states, C.c_uint32),
                                     _p(pos, C.c_uint32), _p(margin, C.c_double))
        return (float(cost), states, pos, margin) if want_margin else (float(cost), states, pos)

    def align_pruned(self, x, automaton, threshold):
        x = _f32(x).reshape(-1, self.dim)
        aut = np.ascontiguousarray(automaton, np.uint32)
        T = len(x)
        states, pos = np.zeros(T, np.uint32), np.zeros(T, np.uint32)
        cost = self.L.orc_align_pruned(self.h, *self._tdp_args(), _p(x, C.c_float), C.c_uint64(T),
                                       _p(aut, C.c_uint32), C.c_uint64(len(aut)), C.c_double(threshold),
                                       _p(states, C.c_uint32), _p(pos, C.c_uint32))
        return float(cost), states, pos

    def align_batch(self, x, frame_off, automata, aut_off, threads=1):
        x = _f32(x)
        frame_off = np.ascontiguousarray(frame_off, np.uint64)
        automata = np.ascontiguousarray(automata, np.uint32)
        aut_off = np.ascontiguousarray(aut_off, np.uint32)
        U = len(frame_off) - 1
        states = np.zeros(int(frame_off[-1]), np.uint32)
        costs = np.zeros(U, np.float64)
        self.L.orc_align_batch(self.h, *self._tdp_args(), _p(x, C.c_float), C.c_uint64(U),
                               _p(frame_off, C.c_uint64), _p(automata, C.c_uint32), _p(aut_off, C.c_uint32),
                               _p(states, C.c_uint32), _p(costs, C.c_double), int(threads))
        return states, costs

    def recognize(self, x, word_off, word_states, silence_word=0, word_penalty=10.0, am_threshold=20.0):
        """Recognizer::recognizeSequence_pruned over one utterance; returns (words, score of the best final word end)."""
        x = _f32(x).reshape(-1, self.dim)
        word_off = np.ascontiguousarray(word_off, np.uint32)
        word_states = np.ascontiguousarray(word_states, np.uint32)
        out = np.zeros(len(x) + 1, np.uint32)
        score = C.c_double(0.0)
        self.L.orc_recognize.restype = C.c_uint64
        n = self.L.orc_recognize(self.h, *self._tdp_args(), _p(x, C.c_float), C.c_uint64(len(x)),
                                 C.c_uint32(len(word_off) - 1), _p(word_off, C.c_uint32), _p(word_states, C.c_uint32),
                                 C.c_uint32(silence_word), C.c_double(word_penalty), C.c_double(am_threshold),
                                 _p(out, C.c_uint32), C.c_uint64(len(out)), C.byref(score))
        return out[:n].copy(), float(score.value)


class Reference:
    """The unmodified reference (oracle/_ref).  Same methods as ``Oracle``; indices are converted to the
    reference's uint16 StateIdx/DensityIdx (Types.hpp:16-17) at the boundary."""

    def __init__(self, dim, n_mix, pooling="none", max_approx=True, tdp=(0.0, 0.0, 0.0), silence=0,
                 mix_path=None, extra_config=None):
        if not have_ref():
            raise RuntimeError("oracle/_ref/libsietill_ref.so not built (make -C oracle ref; needs /root/reference)")
        L = self.L = C.CDLL(REF_SO)
        L.ref_model_create.restype = C.c_void_p
        for fn in ("ref_align_full", "ref_align_pruned", "ref_align_batch", "ref_am_score", "ref_tdp_score"):
            getattr(L, fn).restype = C.c_double
        L.ref_num_densities.restype = C.c_uint64
        L.ref_recognize.restype = C.c_uint64
        self.dim, self.n_mix = dim, n_mix
        cfg = {"verbosity": "noLog", "tdp-loop": float(tdp[0]), "tdp-forward": float(tdp[1]),
               "tdp-skip": float(tdp[2]), "action": "train"}
        if mix_path:
            cfg.update({"action": "recognize", "load-mixtures-from": mix_path})
        cfg.update(extra_config or {})
        fd, path = tempfile.mkstemp(suffix=".json")
        with os.fdopen(fd, "w") as f:
            json.dump(cfg, f)
        try:
            self.h = C.c_void_p(L.ref_model_create(path.encode(), C.c_uint64(dim), C.c_uint64(n_mix),
                                                   POOLING[pooling], int(max_approx), C.c_uint16(silence)))
        finally:
            os.unlink(path)
        self.max_approx = int(max_approx)

    def __del__(self):
        if getattr(self, "h", None):
            self.L.ref_model_destroy(self.h)
            self.h = None

    def sizes(self):
        out = np.zeros(5, np.uint64)
        self.L.ref_model_sizes(self.h, _p(out, C.c_uint64))
        return out

    def tables(self):
        t = _alloc_tables(self.sizes())
        self.L.ref_model_export(self.h, *_table_args(t))
        return t

    def write(self, path):
        if self.L.ref_model_write(self.h, path.encode()):
            raise IOError(path)

    @staticmethod
    def _pad(x):
        # the reference's _mm_loadu_ps reads 2 floats past the last frame (Mixtures.cpp:653)
        x = _f32(x).reshape(-1)
        return np.concatenate([x, np.zeros(4, np.float32)])

    def accumulate(self, states, x, first_pass=False, max_approx=None):
        st = np.ascontiguousarray(states, np.uint16)
        xp = self._pad(x)
        ma = self.max_approx if max_approx is None else int(max_approx)
        self.L.ref_accumulate(self.h, _p(st, C.c_uint16), _p(xp, C.c_float), C.c_uint64(len(st)),
                              int(first_pass), ma)

    def finalize(self):
        self.L.ref_finalize(self.h)

    def split(self, min_obs):
        self.L.ref_split(self.h, C.c_uint64(int(min_obs)))

    def eliminate(self, min_obs):
        self.L.ref_eliminate(self.h, C.c_double(float(min_obs)))

    def num_densities(self):
        return int(self.L.ref_num_densities(self.h))

    def min_score(self, x, states=None):
        F = _f32(x).reshape(-1, self.dim).shape[0]
        xp = self._pad(x)
        st = np.arange(self.n_mix, dtype=np.uint16) if states is None else np.ascontiguousarray(states, np.uint16)
        scores = np.zeros((F, len(st)), np.float64)
        argmin = np.zeros((F, len(st)), np.uint16)
        self.L.ref_min_score(self.h, _p(xp, C.c_float), C.c_uint64(F), _p(st, C.c_uint16),
                             C.c_uint64(len(st)), _p(scores, C.c_double), _p(argmin, C.c_uint16))
        return scores, argmin.astype(np.uint32)

    def score(self, x):
        """FeatureScorer::score of every (frame, state) (Mixtures.cpp:737-744)."""
        F = _f32(x).reshape(-1, self.dim).shape[0]
        xp = self._pad(x)
        st = np.arange(self.n_mix, dtype=np.uint16)
        scores = np.zeros((F, self.n_mix), np.float64)
        self.L.ref_score(self.h, _p(xp, C.c_float), C.c_uint64(F), _p(st, C.c_uint16), C.c_uint64(self.n_mix),
                         _p(scores, C.c_double))
        return scores

    def am_score(self, states, x):
        st = np.ascontiguousarray(states, np.uint16)
        xp = self._pad(x)
        return float(self.L.ref_am_score(self.h, _p(st, C.c_uint16), _p(xp, C.c_float), C.c_uint64(len(st))))

    def align_full(self, x, automaton):
        T = _f32(x).reshape(-1, self.dim).shape[0]
        xp = self._pad(x)
        aut = np.ascontiguousarray(automaton, np.uint16)
        states = np.zeros(T, np.uint16)
        cost = self.L.ref_align_full(self.h, _p(xp, C.c_float), C.c_uint64(T), _p(aut, C.c_uint16),
                                     C.c_uint64(len(aut)), _p(states, C.c_uint16))
        return float(cost), states.astype(np.uint32)

    def align_pruned(self, x, automaton, threshold):
        T = _f32(x).reshape(-1, self.dim).shape[0]
        xp = self._pad(x)
        aut = np.ascontiguousarray(automaton, np.uint16)
        states = np.zeros(T, np.uint16)
        cost = self.L.ref_align_pruned(self.h, _p(xp, C.c_float), C.c_uint64(T), _p(aut, C.c_uint16),
                                       C.c_uint64(len(aut)), C.c_double(threshold), _p(states, C.c_uint16))
        return float(cost), states.astype(np.uint32)

    def align_batch(self, x, frame_off, automata, aut_off):
        xp = self._pad(x)
        frame_off = np.ascontiguousarray(frame_off, np.uint64)
        automata = np.ascontiguousarray(automata, np.uint16)
        aut_off = np.ascontiguousarray(aut_off, np.uint32)
        U = len(frame_off) - 1
        states = np.zeros(int(frame_off[-1]), np.uint16)
        costs = np.zeros(U, np.float64)
        self.L.ref_align_batch(self.h, _p(xp, C.c_float), C.c_uint64(U), _p(frame_off, C.c_uint64),
                               _p(automata, C.c_uint16), _p(aut_off, C.c_uint32), _p(states, C.c_uint16),
                               _p(costs, C.c_double))
        return states.astype(np.uint32), costs

    def linear_segmentation(self, x, automaton, approx=True):
        x = _f32(x).reshape(-1, self.dim)
        xp = self._pad(x)
        aut = np.ascontiguousarray(automaton, np.uint16)
        states = np.zeros(len(x), np.uint16)
        bounds = np.zeros(2, np.uint64)
        self.L.ref_linear_segmentation(self.h, int(approx), _p(xp, C.c_float), C.c_uint64(len(x)), _p(aut, C.c_uint16),
                                       C.c_uint64(len(aut)), _p(states, C.c_uint16), _p(bounds, C.c_uint64))
        return states.astype(np.uint32), (int(bounds[0]), int(bounds[1]))

    def edit_distance(self, ref, hyp):
        r = np.ascontiguousarray(ref, np.uint16)
        h = np.ascontiguousarray(hyp, np.uint16)
        out = np.zeros(4, np.uint64)
        self.L.ref_edit_distance(self.h, _p(r, C.c_uint16), C.c_uint64(len(r)), _p(h, C.c_uint16), C.c_uint64(len(h)),
                                 _p(out, C.c_uint64))
        return tuple(int(v) for v in out)

    def recognize(self, x):
        T = _f32(x).reshape(-1, self.dim).shape[0]
        xp = self._pad(x)
        out = np.zeros(256, np.uint32)
        n = self.L.ref_recognize(self.h, _p(xp, C.c_float), C.c_uint64(T), _p(out, C.c_uint32), C.c_uint64(256))
        return out[:n].copy()


# ---- load-time feature processing: SignalAnalysis::process_features, one utterance at a time ----------
def oracle_process_features(raw, n_first=12, n_second=1, deriv_step=3, mean=None, stddev=None, energy_max_norm=True):
    if not os.path.exists(ORACLE_SO):
        build_port()
    L = C.CDLL(ORACLE_SO)
    raw = np.ascontiguousarray(raw, np.float32)
    T, nf = raw.shape
    out = np.zeros((T, nf + n_first + n_second), np.float32)
    m = None if mean is None else np.ascontiguousarray(mean, np.float64)
    sd = None if stddev is None else np.ascontiguousarray(stddev, np.float64)
    L.orc_process_features(_p(raw, C.c_float), C.c_uint64(T), C.c_uint32(nf), C.c_uint32(n_first), C.c_uint32(n_second),
                           C.c_uint32(deriv_step), _p(m, C.c_double), _p(sd, C.c_double), int(energy_max_norm),
                           _p(out, C.c_float))
    return out


def reference_process_features(raw, n_first=12, n_second=1, deriv_step=3, mean=None, stddev=None, energy_max_norm=True):
    """The unmodified SignalAnalysis (oracle/_ref)."""
    L = C.CDLL(REF_SO)
    L.ref_process_features.restype = C.c_uint64
    raw = np.ascontiguousarray(raw, np.float32)
    T, nf = raw.shape
    cfg = {"verbosity": "noLog", "n-features-file": int(nf), "n-features-first": int(n_first),
           "n-features-second": int(n_second), "deriv-step": int(deriv_step), "energy-max-norm": bool(energy_max_norm)}
    fd, path = tempfile.mkstemp(suffix=".json")
    with os.fdopen(fd, "w") as f:
        json.dump(cfg, f)
    out = np.zeros((T, nf + n_first + n_second), np.float32)
    m = None if mean is None else np.ascontiguousarray(mean, np.float64)
    sd = None if stddev is None else np.ascontiguousarray(stddev, np.float64)
    try:
        nt = L.ref_process_features(path.encode(), _p(m, C.c_double), _p(sd, C.c_double), _p(raw, C.c_float),
                                    C.c_uint64(T), _p(out, C.c_float))
    finally:
        os.unlink(path)
    assert nt == out.shape[1]
    return out
