"""CIDEr-D reward on integer token ids -- restatement of the scorer the reference's training fitness calls.

TEST INFRASTRUCTURE (see oracle/__init__.py).  PINNED against the imported reference by oracle/refcheck.py
(CiderD(df='corpus') on synthetic token strings) and by tests/golden/golden_cider_v1.npz.

Follows
  * array_to_str                              src/algorithm/tools/utils.py:34-40   (a caption = tokens up to AND INCLUDING the first 0)
  * CaptPolicy.compute_ciders                 src/captioning/policies.py:145-193   (res i is scored against the refs of image i // seq_per_img)
  * precook / cook_refs / cook_test           cider/pyciderevalcap/ciderD/ciderD_scorer.py:17-51
  * CiderScorer.compute_doc_freq              :107-118   (corpus mode: one "document" per scored caption, so each image counts seq_per_img times)
  * counts2vec / sim / compute_cider          :120-199   (tf-idf vectors, clipped cosine, Gaussian length penalty, x10)
  * CiderScorer.compute_score                 :201-213
Every floating-point operation is done in the reference's order (dict insertion order = n-gram length, then position), in float64,
so the result is bit-identical to the reference's.
"""
import numpy as np

N_GRAM = 4
SIGMA = 6.0


def caption_tokens(arr):
    """array_to_str: tokens up to and including the first 0 (all of them if there is none)"""
    out = []
    for t in arr:
        out.append(int(t))
        if int(t) == 0:
            break
    return out


def precook(words, n=N_GRAM):
    """n-gram -> count, insertion-ordered like the reference's defaultdict (k = 1..n outer, position inner)"""
    counts = {}
    for k in range(1, n + 1):
        for i in range(len(words) - k + 1):
            g = tuple(words[i:i + k])
            counts[g] = counts.get(g, 0) + 1
    return counts


def corpus_df(refs_per_caption):
    """compute_doc_freq over crefs (one entry per scored caption) -> (document_frequency dict, ref_len = log(#captions))"""
    df = {}
    for refs in refs_per_caption:
        for g in set(g for ref in refs for g in ref):
            df[g] = df.get(g, 0.0) + 1.0
    return df, np.log(float(len(refs_per_caption)))


def counts2vec(cnts, df, ref_len, n=N_GRAM):
    vec = [dict() for _ in range(n)]
    length = 0
    norm = [0.0 for _ in range(n)]
    for g, tf in cnts.items():
        d = np.log(max(1.0, df.get(g, 0.0)))
        k = len(g) - 1
        vec[k][g] = float(tf) * (ref_len - d)
        norm[k] += pow(vec[k][g], 2)
        if k == 1:
            length += tf
    return vec, [np.sqrt(x) for x in norm], length


def sim(vec_hyp, vec_ref, norm_hyp, norm_ref, length_hyp, length_ref, n=N_GRAM, sigma=SIGMA):
    delta = float(length_hyp - length_ref)
    val = np.array([0.0 for _ in range(n)])
    for k in range(n):
        for g in vec_hyp[k]:
            r = vec_ref[k].get(g, 0.0)
            val[k] += min(vec_hyp[k][g], r) * r
        if norm_hyp[k] != 0 and norm_ref[k] != 0:
            val[k] /= (norm_hyp[k] * norm_ref[k])
        val[k] *= np.e ** (-(delta ** 2) / (2 * sigma ** 2))
    return val


def cider_scores(seq, gts, seq_per_img=1, df=None, ref_len=None):
    """seq int [rows, T]; gts: list over images of int arrays [n_refs_i, T]; rows = len(gts) * seq_per_img.
    df None = corpus mode (the document frequencies come from the refs of this very batch).
    -> (mean score, scores float64 [rows]) exactly like CiderD.compute_score."""
    rows = len(seq)
    assert rows == len(gts) * seq_per_img
    cooked_refs = [[precook(caption_tokens(r)) for r in refs] for refs in gts]
    crefs = [cooked_refs[i // seq_per_img] for i in range(rows)]
    if df is None:
        df, ref_len = corpus_df(crefs)
    scores = []
    for i in range(rows):
        vec, norm, length = counts2vec(precook(caption_tokens(seq[i])), df, ref_len)
        score = np.array([0.0 for _ in range(N_GRAM)])
        for ref in crefs[i]:
            vec_ref, norm_ref, length_ref = counts2vec(ref, df, ref_len)
            score += sim(vec, vec_ref, norm, norm_ref, length, length_ref)
        s = np.mean(score)
        s /= len(crefs[i])
        s *= 10.0
        scores.append(s)
    return np.mean(np.array(scores)), np.array(scores)
