"""Synthetic sietill-shaped corpora and the host-side automaton builder.

Shapes follow SURVEY.md §8.0 / §8(d): the German-digits lexicon of the reference
(src/sietill/Lexicon.cpp:70-85: silence = 1 state x1 repetition, 11 words of 9 or 12 states x2
repetitions -> 106 states), utterance automata ``sil w1 sil w2 ... sil`` (Training.cpp:239-253) and
the bundled corpus' utterance-length distribution.  Seeded with numpy's PCG64 (seed 20161107).
"""
import numpy as np

SEED = 20161107
# (states, repetitions) per word, in lexicon order; word 0 is silence (Lexicon.cpp:70-85)
SIETILL_WORDS = [(1, 1), (9, 2), (9, 2), (9, 2), (9, 2), (12, 2), (9, 2), (12, 2), (9, 2), (9, 2), (9, 2), (9, 2)]
# T distribution of data/new_features over corpus_train.json
SIETILL_T = ([279, 500, 701, 1399], [0.903, 0.072, 0.005, 0.020])


class Lexicon:
    """Linear whole-word HMMs with contiguous state ids (mirror of Lexicon::add_word, Lexicon.cpp:11-23)."""

    def __init__(self, words=SIETILL_WORDS, silence=0):
        self.words, self.silence = list(words), silence
        self.first_state, self.automata = [], []
        nxt = 0
        for n, rep in self.words:
            self.first_state.append(nxt)
            self.automata.append(np.repeat(np.arange(nxt, nxt + n, dtype=np.uint32), rep))
            nxt += n
        self.num_states = nxt
        self.silence_state = int(self.automata[silence][0])

    def segment_automaton(self, word_ids):
        """sil w1 sil w2 ... sil  (Trainer::build_segment_automaton, Training.cpp:239-253)."""
        parts = []
        for w in word_ids:
            parts.append(self.automata[self.silence])
            parts.append(self.automata[w])
        parts.append(self.automata[self.silence])
        return np.concatenate(parts).astype(np.uint32)


def chain_lexicon(n_states, phones_per_utt=50, states_per_phone=3):
    """rwth-shaped stress (config 5): utterance = random chain of tied triphone states, 3 per phone."""
    return {"n_states": n_states, "phones": phones_per_utt, "spp": states_per_phone}


def random_model(rng, dim, n_states, dens_per_state, pooling="none"):
    """Parameter tables of a GMM in the reference's SoA layout (Mixtures.hpp:72-87), state-major.

    mu ~ N(0, 1.5^2), sigma^2 ~ LogUniform[0.05, 2], weights ~ Dirichlet(1).
    Returns a dict usable by ``Oracle.load_tables`` / ``MixtureModel.load_tables``; norm_ and
    mean_weights_log_ are derived in f64 exactly as finalize() would (Mixtures.cpp:268-274,403).
    """
    D = n_states * dens_per_state
    V = {"none": D, "mixture": n_states, "global": 1}[pooling]
    means = rng.normal(0.0, 1.5, size=(D, dim))
    var = np.exp(rng.uniform(np.log(0.05), np.log(2.0), size=(V, dim)))
    w = rng.dirichlet(np.ones(dens_per_state), size=n_states).reshape(-1)
    w = np.maximum(w, 1e-6)
    norm = np.array([(dim * np.log(2 * np.pi) + np.sum(np.log(v))) / 2 for v in var])
    mean_idx = np.arange(D, dtype=np.uint32)
    if pooling == "none":
        var_idx = mean_idx.copy()
    elif pooling == "mixture":
        var_idx = np.repeat(np.arange(n_states, dtype=np.uint32), dens_per_state)
    else:
        var_idx = np.zeros(D, np.uint32)
    # sufficient statistics that finalize() (Mixtures.cpp:374-461) turns back into these parameters, so the
    # model can also travel as a .mix file: counts follow the weights, sum_x = mean * count,
    # sum_xx = (var + mean_ref^2) * count_v with mean_ref the own / mixture / global mean.
    cnt = w * 100.0 * dens_per_state
    macc = means * cnt[:, None]
    if pooling == "none":
        vcnt, mref = cnt.copy(), means
    elif pooling == "mixture":
        vcnt = cnt.reshape(n_states, dens_per_state).sum(1)
        mref = macc.reshape(n_states, dens_per_state, dim).sum(1) / vcnt[:, None]
    else:
        vcnt = np.array([cnt.sum()])
        mref = macc.sum(0, keepdims=True) / vcnt[:, None]
    vacc = (var + mref * mref) * vcnt[:, None]
    return {
        "means": means.reshape(-1).copy(), "mean_acc": macc.reshape(-1).copy(), "mean_w": w,
        "mean_wlog": np.log(w), "mean_wacc": cnt, "mean_refs": np.ones(D, np.uint64),
        "vars": var.reshape(-1).copy(), "vars_inv": (1.0 / var).reshape(-1).copy(),
        "var_acc": vacc.reshape(-1).copy(), "var_wacc": vcnt, "var_refs": np.ones(V, np.uint64),
        "norm": norm, "mix_offsets": (np.arange(n_states + 1) * dens_per_state).astype(np.uint32),
        "mean_idx": mean_idx, "var_idx": var_idx,
    }


def sample_path(rng, T, S):
    """Monotone 0-1 path over S positions in T frames, ending at S-1 (needs S <= T)."""
    cuts = np.sort(rng.choice(np.arange(1, T), size=S - 1, replace=False)) if S > 1 else np.array([], int)
    pos = np.zeros(T, np.int64)
    pos[cuts] = 1
    return np.cumsum(pos)


def emit_frames(rng, tables, dim, states):
    """x_f = mu + sigma * N(0,1) from a random density of the frame's state, cast to f32."""
    off = tables["mix_offsets"]
    n = (off[states + 1] - off[states]).astype(np.int64)
    k = off[states] + (rng.random(len(states)) * n).astype(np.int64)
    mu = tables["means"].reshape(-1, dim)[tables["mean_idx"][k]]
    sd = np.sqrt(tables["vars"].reshape(-1, dim)[tables["var_idx"][k]])
    return (mu + sd * rng.standard_normal((len(states), dim))).astype(np.float32)


def make_corpus(rng, lexicon, tables, dim, n_utts, t_choices=SIETILL_T, words_per_utt=(3, 7)):
    """Sietill-shaped corpus: returns dict(x[F,d] f32, frame_off[U+1] u64, automata u32, aut_off[U+1] u32,
    true_states[F] u32, words=list)."""
    xs, auts, f_off, a_off, truth, words = [], [], [0], [0], [], []
    Ts = rng.choice(t_choices[0], size=n_utts, p=np.array(t_choices[1]) / np.sum(t_choices[1]))
    for u in range(n_utts):
        nw = int(rng.integers(words_per_utt[0], words_per_utt[1] + 1))
        wid = rng.integers(1, len(lexicon.words), size=nw)
        aut = lexicon.segment_automaton(wid)
        T = int(max(Ts[u], len(aut)))
        st = aut[sample_path(rng, T, len(aut))]
        xs.append(emit_frames(rng, tables, dim, st))
        auts.append(aut); truth.append(st); words.append(wid)
        f_off.append(f_off[-1] + T); a_off.append(a_off[-1] + len(aut))
    return {"x": np.concatenate(xs), "frame_off": np.array(f_off, np.uint64),
            "automata": np.concatenate(auts).astype(np.uint32), "aut_off": np.array(a_off, np.uint32),
            "true_states": np.concatenate(truth).astype(np.uint32), "words": words}


def make_chain_corpus(rng, tables, dim, n_states, n_utts, T=500, phones=50, spp=3):
    """Config-5-shaped corpus: each utterance is a linear chain of `phones` x `spp` random tied states."""
    S = phones * spp
    xs, auts, truth = [], [], []
    for _ in range(n_utts):
        aut = rng.integers(0, n_states, size=S).astype(np.uint32)
        st = aut[sample_path(rng, T, S)]
        xs.append(emit_frames(rng, tables, dim, st)); auts.append(aut); truth.append(st)
    return {"x": np.concatenate(xs), "frame_off": (np.arange(n_utts + 1) * T).astype(np.uint64),
            "automata": np.concatenate(auts), "aut_off": (np.arange(n_utts + 1) * S).astype(np.uint32),
            "true_states": np.concatenate(truth).astype(np.uint32)}


def linear_alignment(frame_off, automata, aut_off):
    """Uniform initial segmentation: frame t of an utterance -> position floor(t*S/T).  (Stand-in for the
    reference's energy-based linear segmentation, Training.cpp:429-546, which is a 'next' row.)"""
    out = np.zeros(int(frame_off[-1]), np.uint32)
    for u in range(len(frame_off) - 1):
        f0, f1 = int(frame_off[u]), int(frame_off[u + 1])
        aut = automata[int(aut_off[u]):int(aut_off[u + 1])]
        T = f1 - f0
        out[f0:f1] = aut[(np.arange(T) * len(aut)) // T]
    return out
