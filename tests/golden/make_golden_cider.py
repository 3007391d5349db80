"""Generate tests/golden/golden_cider_v1.npz by running the REAL reference scorer (cider/pyciderevalcap CiderD(df='corpus'),
fed exactly like CaptPolicy.compute_ciders does, src/captioning/policies.py:145-193, with the reference's own array_to_str)
on seeded synthetic token ids.  Run in the build container only:   python tests/golden/make_golden_cider.py"""
import os
import sys
from collections import OrderedDict
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import _refimport  # noqa: E402

CASES = {
    # name: images, seq_per_img, refs per image (max), vocab (small vocab => many shared n-grams), seed
    'c1': dict(n_img=6, spi=1, n_refs=5, vocab=12, seed=1),
    'c2': dict(n_img=16, spi=5, n_refs=5, vocab=40, seed=2),
    'c3': dict(n_img=4, spi=2, n_refs=3, vocab=5, seed=3),
}
T = 16


def case_inputs(c):
    """seq [n_img*spi, T], gts list of [n_refs_i, T] int arrays; captions end with 0 at a random length (some have no 0 at all,
    some are empty = start with 0); hypotheses are noisy copies of references so that scores are not all zero"""
    rs = np.random.RandomState(c['seed'])
    gts = []
    for i in range(c['n_img']):
        k = rs.randint(1, c['n_refs'] + 1)
        refs = np.zeros((k, T), dtype=np.int64)
        for j in range(k):
            ln = rs.randint(0, T + 1)
            refs[j, :ln] = rs.randint(1, c['vocab'], ln)
        gts.append(refs)
    rows = c['n_img'] * c['spi']
    seq = np.zeros((rows, T), dtype=np.int64)
    for r in range(rows):
        refs = gts[r // c['spi']]
        base = refs[rs.randint(len(refs))].copy()
        flip = rs.rand(T) < 0.25
        base[flip] = rs.randint(0, c['vocab'], flip.sum())
        seen0 = False
        for t in range(T):           # like the decoder: nothing but 0 after the first 0
            if seen0:
                base[t] = 0
            seen0 = seen0 or base[t] == 0
        seq[r] = base
    return seq, gts


def reference_scores(seq, gts, spi):
    _refimport.load()
    from algorithm.tools.utils import array_to_str
    from pyciderevalcap.ciderD.ciderD import CiderD
    batch = len(seq)
    res = OrderedDict((i, [array_to_str(seq[i])]) for i in range(batch))
    g = OrderedDict((i, [array_to_str(gts[i][j]) for j in range(len(gts[i]))]) for i in range(len(gts)))
    res_ = [{'image_id': i, 'caption': res[i]} for i in range(batch)]
    g2 = {i: g[i % batch // spi] for i in range(batch)}
    score, scores = CiderD(df='corpus').compute_score(g2, res_)
    return float(score), np.asarray(scores, dtype=np.float64)


if __name__ == '__main__':
    out = {}
    for name, c in CASES.items():
        seq, gts = case_inputs(c)
        score, scores = reference_scores(seq, gts, c['spi'])
        out[name + '_score'] = np.float64(score)
        out[name + '_scores'] = scores
        print(name, score, scores[:4])
    np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden_cider_v1.npz'), **out)
