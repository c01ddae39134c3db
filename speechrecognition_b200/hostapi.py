"""ctypes view of the C++ host classes (csrc/host/: MixtureModel, Aligner, Trainer, ...) through the flat shim
in csrc/host/host_capi.cpp.  Used by the parity tests so that they exercise the same C++ objects a C++ caller
of the drop-in would; the compute methods run on the GPU only (no CPU fallback — they raise without one).
"""
import ctypes as C

import numpy as np

from . import capi

POOLING = {"none": 0, "mixture": 1, "global": 2}
TABLES = ["means", "mean_acc", "mean_w", "mean_wlog", "mean_wacc", "mean_refs", "vars", "vars_inv",
          "var_acc", "var_wacc", "var_refs", "norm", "mix_offsets", "mean_idx", "var_idx"]
_CT = {np.dtype(np.float64): C.c_double, np.dtype(np.uint64): C.c_uint64, np.dtype(np.uint32): C.c_uint32}


class HostError(RuntimeError):
    pass


def _lib():
    L = capi.load()
    if not getattr(L, "_sthost_ready", False):
        L.sthost_last_error.restype = C.c_char_p
        L.sthost_model_create.restype = C.c_void_p
        L.sthost_model_create.argtypes = [C.c_uint32, C.c_uint32, C.c_int, C.c_int, C.c_int, C.c_int, C.c_char_p]
        L.sthost_model_destroy.argtypes = [C.c_void_p]
        L.sthost_model_num_densities.restype = C.c_uint64
        L.sthost_model_num_densities.argtypes = [C.c_void_p]
        L.sthost_tdp_score.restype = C.c_double
        L.sthost_segment_automaton.restype = C.c_uint64
        L._sthost_ready = True
    return L


def set_device_mstep(on):
    """Default of the "device-mstep" key for models created afterwards: True = MixtureModel::finalize runs on the GPU
    (sgmm_mstep) and the statistics stay there; False = statistics come home, finalize is host code."""
    _lib().sthost_set_device_mstep(int(bool(on)))


def _p(a, t):
    return None if a is None else a.ctypes.data_as(C.POINTER(t))


def _ck(L, rc):
    if rc:
        raise HostError(L.sthost_last_error().decode())


class HostModel:
    """sietill_b200::MixtureModel."""

    def __init__(self, dim, n_mix, pooling="none", max_approx=True, host_only=False, device=0, load_from=None, _handle=None):
        self.L = _lib()
        self.dim, self.n_mix = dim, n_mix
        if _handle is not None:
            self.h = _handle
            return
        h = self.L.sthost_model_create(dim, n_mix, POOLING[pooling], int(max_approx), int(host_only), device,
                                       (load_from or "").encode())
        if not h:
            raise HostError(self.L.sthost_last_error().decode())
        self.h = C.c_void_p(h)

    def close(self):
        if getattr(self, "h", None):
            self.L.sthost_model_destroy(self.h)
            self.h = None

    __del__ = close

    def sizes(self):
        out = np.zeros(5, np.uint64)
        _ck(self.L, self.L.sthost_model_sizes(self.h, _p(out, C.c_uint64)))
        return out

    def tables(self):
        d, n, M, V, D = [int(v) for v in self.sizes()]
        f8, u8, u4 = np.float64, np.uint64, np.uint32
        t = {"means": np.zeros(M * d, f8), "mean_acc": np.zeros(M * d, f8), "mean_w": np.zeros(M, f8),
             "mean_wlog": np.zeros(M, f8), "mean_wacc": np.zeros(M, f8), "mean_refs": np.zeros(M, u8),
             "vars": np.zeros(V * d, f8), "vars_inv": np.zeros(V * d, f8), "var_acc": np.zeros(V * d, f8),
             "var_wacc": np.zeros(V, f8), "var_refs": np.zeros(V, u8), "norm": np.zeros(V, f8),
             "mix_offsets": np.zeros(n + 1, u4), "mean_idx": np.zeros(D, u4), "var_idx": np.zeros(D, u4)}
        _ck(self.L, self.L.sthost_model_export(self.h, *[_p(t[k], _CT[t[k].dtype]) for k in TABLES]))
        return t

    def load_tables(self, t):
        want = {"mean_refs": np.uint64, "var_refs": np.uint64, "mix_offsets": np.uint32, "mean_idx": np.uint32, "var_idx": np.uint32}
        t = {k: np.ascontiguousarray(t[k], want.get(k, np.float64)).reshape(-1) for k in TABLES}
        _ck(self.L, self.L.sthost_model_import(self.h, C.c_uint32(len(t["mean_w"])), C.c_uint32(len(t["var_wacc"])),
                                               *[_p(t[k], _CT[t[k].dtype]) for k in TABLES]))

    def finalize(self):
        _ck(self.L, self.L.sthost_model_finalize(self.h))

    def split(self, min_obs):
        _ck(self.L, self.L.sthost_model_split(self.h, C.c_uint64(int(min_obs))))

    def eliminate(self, min_obs):
        _ck(self.L, self.L.sthost_model_eliminate(self.h, C.c_double(float(min_obs))))

    def num_densities(self):
        return int(self.L.sthost_model_num_densities(self.h))

    def write(self, path):
        _ck(self.L, self.L.sthost_model_write(self.h, str(path).encode()))

    def read(self, path):
        _ck(self.L, self.L.sthost_model_read(self.h, str(path).encode()))

    # ---- GPU
    def accumulate(self, ali, x, first_pass=False):
        """ali: capi.ALI_DTYPE array (or uint32 states).  Returns the AM-score numerator of the E-step."""
        ali = as_alignment(ali)
        x = np.ascontiguousarray(x, np.float32)
        s = C.c_double(0.0)
        _ck(self.L, self.L.sthost_model_accumulate(self.h, ali.ctypes.data_as(C.c_void_p), _p(x, C.c_float),
                                                   C.c_uint64(len(ali)), int(first_pass), C.byref(s)))
        return s.value

    def accumulate_soft(self, ali, x):
        """MixtureModel::accumulate(first_pass=false, max_approx=false).  Returns the sum of the frames' sum scores."""
        ali = as_alignment(ali)
        x = np.ascontiguousarray(x, np.float32)
        s = C.c_double(0.0)
        _ck(self.L, self.L.sthost_model_accumulate_soft(self.h, ali.ctypes.data_as(C.c_void_p), _p(x, C.c_float),
                                                        C.c_uint64(len(ali)), C.byref(s)))
        return s.value

    def scores(self, x, prepared=True):
        """FeatureScorer::score per (frame, state): sum scores for a model built with max_approx=False."""
        x = np.ascontiguousarray(x, np.float32).reshape(-1, self.dim)
        sc = np.zeros((len(x), self.n_mix), np.float64)
        _ck(self.L, self.L.sthost_model_scores(self.h, _p(x, C.c_float), C.c_uint64(len(x)), int(prepared),
                                               _p(sc, C.c_double)))
        return sc

    def process_features(self, raw, frame_off, n_first=12, n_second=1, deriv_step=3, energy_max_norm=True,
                         normalization_path=""):
        """SignalAnalysis::process_features for all segments on this model's GPU context; returns the processed matrix
        (which also stays resident as the context's feature matrix)."""
        raw = np.ascontiguousarray(raw, np.float32)
        fo = np.ascontiguousarray(frame_off, np.uint64)
        out = np.empty((raw.shape[0], raw.shape[1] + n_first + n_second), np.float32)
        _ck(self.L, self.L.sthost_process_features(self.h, C.c_uint32(raw.shape[1]), C.c_uint32(n_first), C.c_uint32(n_second),
                                                   C.c_uint32(deriv_step), int(energy_max_norm), str(normalization_path).encode(),
                                                   _p(raw, C.c_float), C.c_uint64(raw.shape[0]), C.c_uint64(len(fo) - 1),
                                                   _p(fo, C.c_uint64), _p(out, C.c_float)))
        return out

    def score_sequence(self, x, prepared=True):
        x = np.ascontiguousarray(x, np.float32).reshape(-1, self.dim)
        sc = np.zeros((len(x), self.n_mix), np.float64)
        am = np.zeros((len(x), self.n_mix), np.uint32)
        _ck(self.L, self.L.sthost_model_score_sequence(self.h, _p(x, C.c_float), C.c_uint64(len(x)), int(prepared),
                                                       _p(sc, C.c_double), _p(am, C.c_uint32)))
        return sc, am

    def align_sequence(self, tdp, silence, x, automaton, pruned=False, threshold=1e300):
        x = np.ascontiguousarray(x, np.float32).reshape(-1, self.dim)
        aut = np.ascontiguousarray(automaton, np.uint16)
        out = np.zeros(len(x), capi.ALI_DTYPE)
        cost = C.c_double(0.0)
        _ck(self.L, self.L.sthost_align_sequence(self.h, C.c_double(tdp[0]), C.c_double(tdp[1]), C.c_double(tdp[2]),
                                                 C.c_uint32(silence), _p(x, C.c_float), C.c_uint64(len(x)),
                                                 _p(aut, C.c_uint16), C.c_uint64(len(aut)), int(pruned),
                                                 C.c_double(threshold), out.ctypes.data_as(C.c_void_p), C.byref(cost)))
        return cost.value, out


def as_alignment(a):
    a = np.asarray(a)
    if a.dtype == capi.ALI_DTYPE:
        return np.ascontiguousarray(a)
    out = np.zeros(len(a), capi.ALI_DTYPE)
    out["count"], out["state"], out["weight"] = 1, a.astype(np.uint16), 1.0
    return out


def train(dim, pooling, x, frame_off, words, word_off, tdp, min_obs, num_splits, num_aligns, num_estimates,
          alignment_pruning=False, initial=None, mixture_path="", alignment_path="", device=0,
          pruning_threshold=float("inf"), max_approx=True):
    """Trainer::train on the sietill digits lexicon.  Returns (stats[n,4], final alignment, HostModel, timers)."""
    L = _lib()
    x = np.ascontiguousarray(x, np.float32)
    fo = np.ascontiguousarray(frame_off, np.uint64)
    w = np.ascontiguousarray(words, np.uint64)
    wo = np.ascontiguousarray(word_off, np.uint64)
    U = len(fo) - 1
    init = None if initial is None else as_alignment(initial)
    cap = 4 + (num_splits + 1) * (1 + num_aligns * max(1, num_estimates))
    stats = np.zeros((cap, 4), np.float64)
    n_stats = C.c_uint64(0)
    final = np.zeros(int(fo[-1]), capi.ALI_DTYPE)
    timers = np.zeros(4, np.float64)
    h = C.c_void_p()
    _ck(L, L.sthost_train(C.c_uint32(dim), POOLING[pooling], device, _p(x, C.c_float), C.c_uint64(U), _p(fo, C.c_uint64),
                          _p(w, C.c_uint64), _p(wo, C.c_uint64), C.c_double(tdp[0]), C.c_double(tdp[1]), C.c_double(tdp[2]),
                          C.c_uint32(min_obs), C.c_uint32(num_splits), C.c_uint32(num_aligns), C.c_uint32(num_estimates),
                          int(alignment_pruning), C.c_double(pruning_threshold), int(max_approx), None if init is None else init.ctypes.data_as(C.c_void_p),
                          mixture_path.encode(), alignment_path.encode(), _p(stats, C.c_double), C.c_uint64(cap),
                          C.byref(n_stats), final.ctypes.data_as(C.c_void_p), _p(timers, C.c_double), C.byref(h)))
    model = HostModel(dim, 106, _handle=h)
    return stats[:n_stats.value], final, model, dict(zip(["align", "estimate", "score", "io"], timers))


def shard_utterances(frame_off, world_size):
    L = _lib()
    fo = np.ascontiguousarray(frame_off, np.uint64)
    out = np.zeros(world_size + 1, np.uint64)
    _ck(L, L.sthost_shard_utterances(_p(fo, C.c_uint64), C.c_uint64(len(fo) - 1), int(world_size), _p(out, C.c_uint64)))
    return out


def tdp_score(tdp, silence, state, jump):
    return float(_lib().sthost_tdp_score(C.c_double(tdp[0]), C.c_double(tdp[1]), C.c_double(tdp[2]), C.c_uint32(silence),
                                         C.c_uint32(state), C.c_uint64(jump)))


def segment_automaton(word_ids):
    w = np.ascontiguousarray(word_ids, np.uint64)
    out = np.zeros(4096, np.uint16)
    n = int(_lib().sthost_segment_automaton(_p(w, C.c_uint64), C.c_uint64(len(w)), _p(out, C.c_uint16), C.c_uint64(len(out))))
    return out[:n].copy()


def edit_distance(ref, hyp):
    """Recognizer::editDistance -> (total, substitutions, insertions, deletions)."""
    r = np.ascontiguousarray(ref, np.uint64)
    h = np.ascontiguousarray(hyp, np.uint64)
    out = np.zeros(4, np.uint64)
    _lib().sthost_edit_distance(_p(r, C.c_uint64), C.c_uint64(len(r)), _p(h, C.c_uint64), C.c_uint64(len(h)),
                                _p(out, C.c_uint64))
    return tuple(int(v) for v in out)


def recognize(model, x, frame_off, words, word_off, tdp, word_penalty=10.0, am_threshold=20.0):
    """Recognizer::recognize with a HostModel over the sietill digits lexicon.
    Returns (list of recognised word-id arrays, dict of WER/SER/error counts)."""
    L = _lib()
    x = np.ascontiguousarray(x, np.float32)
    fo = np.ascontiguousarray(frame_off, np.uint64)
    w = np.ascontiguousarray(words, np.uint64)
    wo = np.ascontiguousarray(word_off, np.uint64)
    U = len(fo) - 1
    hyp = np.zeros(int(fo[-1]) + 1, np.uint64)
    off = np.zeros(U + 1, np.uint64)
    stats = np.zeros(8, np.float64)
    _ck(L, L.sthost_recognize(model.h, _p(x, C.c_float), C.c_uint64(U), _p(fo, C.c_uint64), _p(w, C.c_uint64),
                              _p(wo, C.c_uint64), C.c_double(tdp[0]), C.c_double(tdp[1]), C.c_double(tdp[2]),
                              C.c_double(word_penalty), C.c_double(am_threshold), _p(hyp, C.c_uint64),
                              C.c_uint64(len(hyp)), _p(off, C.c_uint64), _p(stats, C.c_double)))
    names = ["wer", "ser", "errors", "substitutions", "insertions", "deletions", "ref_words", "seconds"]
    return [hyp[int(off[u]):int(off[u + 1])].astype(np.uint32) for u in range(U)], dict(zip(names, stats))


def linear_segmentation(x, automaton, approx=True):
    """Trainer::linear_segmentation_approximation / _running_sums on one utterance -> (states[T], (b1, b2))."""
    L = _lib()
    x = np.ascontiguousarray(x, np.float32)
    aut = np.ascontiguousarray(automaton, np.uint16)
    states = np.zeros(len(x), np.uint16)
    bounds = np.zeros(2, np.uint64)
    _ck(L, L.sthost_linear_segmentation(int(approx), _p(x, C.c_float), C.c_uint64(x.shape[0]), C.c_uint32(x.shape[1]),
                                        _p(aut, C.c_uint16), C.c_uint64(len(aut)), _p(states, C.c_uint16),
                                        _p(bounds, C.c_uint64)))
    return states.astype(np.uint32), (int(bounds[0]), int(bounds[1]))
