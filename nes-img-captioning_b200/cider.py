"""CIDEr-D reward with the reference's interface (cider/pyciderevalcap/ciderD/ciderD.py: CiderD(df=...).compute_score)
and the batch-side preparation for the device scorer k_cider (csrc/cider.cu).

What depends only on the batch -- the reference captions' tf-idf vectors, norms and lengths, and the document
frequencies -- is computed here once per batch, in float64 and in the reference's operation order
(ciderD_scorer.py:107-145: dict insertion order = n-gram length, then position), so the uploaded numbers are the very
ones the reference would use.  The per-caption work (2P x rows hypotheses per NES iteration, the reference's
Python-dict hot spot) runs on the device: `nes_cider_scores`.
"""
import numpy as np
import torch

from . import _lib

N_GRAM = 4
PENALTY_LEN = 128


def caption_tokens(arr):
    """array_to_str (src/algorithm/tools/utils.py:34-40): tokens up to and including the first 0"""
    out = []
    for t in arr:
        out.append(int(t))
        if int(t) == 0:
            break
    return out


def pack_ngram(g):
    key = len(g)
    for j, t in enumerate(g):
        assert 0 <= int(t) < 32768, 'token ids must be below 32768'
        key |= int(t) << (3 + 15 * j)
    return key


def precook(words, n=N_GRAM):
    counts = {}
    for k in range(1, n + 1):
        for i in range(len(words) - k + 1):
            g = tuple(words[i:i + k])
            counts[g] = counts.get(g, 0) + 1
    return counts


class CiderD:
    """df='corpus': document frequencies from the references of the scored batch (one document per scored caption, so an
    image counts seq_per_img times -- ciderD_scorer.py:107-118 as fed by captioning/policies.py:145-193).
    Otherwise pass the content of the reference's data/<df>.p pickle: document_frequency = {tuple of token strings or
    ints: count}, ref_len = number of documents (ciderD_scorer.py:72-75)."""

    def __init__(self, ctx, n=4, sigma=6.0, df='corpus', document_frequency=None, ref_len=None):
        assert n == N_GRAM, 'the device scorer is built for 4-grams'
        self.ctx, self._n, self._sigma, self._df = ctx, n, float(sigma), df
        if df != 'corpus':
            assert document_frequency is not None and ref_len is not None, 'give the pickled document frequencies'
            self.document_frequency = {tuple(int(t) for t in g): float(c) for g, c in document_frequency.items()}
            self.ref_len = np.log(float(ref_len))
        self.rows_per_img = 1
        self.n_img = 0

    def method(self):
        return 'CIDEr-D'

    def set_refs(self, gts, seq_per_img=1, rows_per_img=1):
        """gts: list over images of int arrays [n_refs_i, T] (data['gts'] of the reference's loader).
        seq_per_img: how many scored captions each image has in the REFERENCE's batch (enters the corpus statistics);
        rows_per_img: how many consecutive rows of the device batch belong to one image (1 when unique rows are decoded)."""
        cooked = [[precook(caption_tokens(r)) for r in refs] for refs in gts]
        if self._df == 'corpus':
            df = {}
            for refs in cooked:
                for g in set(g for ref in refs for g in ref):
                    df[g] = df.get(g, 0.0) + float(seq_per_img)      # each image is a document seq_per_img times
            ref_len = np.log(float(len(cooked) * seq_per_img))
        else:
            df, ref_len = self.document_frequency, self.ref_len
        idf_keys = np.fromiter((pack_ngram(g) for g in df), dtype=np.uint64, count=len(df))
        idf_vals = np.array([ref_len - np.log(max(1.0, c)) for c in df.values()], dtype=np.float64)
        img_ptr, ent_ptr, keys, ws, norms, lens = [0], [0], [], [], [], []
        for refs in cooked:
            for ref in refs:
                norm = [0.0] * N_GRAM
                length = 0
                for g, tf in ref.items():
                    d = np.log(max(1.0, df.get(g, 0.0)))
                    k = len(g) - 1
                    v = float(tf) * (ref_len - d)
                    norm[k] += pow(v, 2)
                    if k == 1:
                        length += tf
                    keys.append(pack_ngram(g)); ws.append(v)
                ent_ptr.append(len(keys))
                norms.append([np.sqrt(x) for x in norm]); lens.append(length)
            img_ptr.append(len(lens))
        penalty = np.array([np.e ** (-(float(d) ** 2) / (2 * self._sigma ** 2)) for d in range(-64, PENALTY_LEN - 64)], dtype=np.float64)
        a = dict(idf_keys=idf_keys, idf_vals=idf_vals, img_ptr=np.asarray(img_ptr, dtype=np.int32), ent_ptr=np.asarray(ent_ptr, dtype=np.int32),
                 keys=np.asarray(keys, dtype=np.uint64), ws=np.asarray(ws, dtype=np.float64),
                 norms=np.ascontiguousarray(np.asarray(norms, dtype=np.float64).reshape(-1, 4)), lens=np.asarray(lens, dtype=np.int32), penalty=penalty)
        assert max(lens, default=0) <= 63, 'reference captions longer than 64 words'
        _lib.check(self.ctx.lib.nes_cider_set_tables(
            self.ctx._h, _lib.ptr(a['idf_keys']), _lib.ptr(a['idf_vals']), len(idf_keys), float(ref_len),
            _lib.ptr(a['img_ptr']), len(cooked), _lib.ptr(a['ent_ptr']), len(lens), _lib.ptr(a['keys']), _lib.ptr(a['ws']),
            _lib.ptr(a['norms']), _lib.ptr(a['lens']), _lib.ptr(a['penalty'])))
        self.n_img, self.rows_per_img, self.seq_per_img = len(cooked), int(rows_per_img), int(seq_per_img)
        return self

    def scores(self, seq=None, n_members=None, rows=None, device='cpu', rows_per_img=None):
        """per-caption scores float64 [n_members, rows]; seq int64 [n_members, rows, T] (torch / numpy, host or device) or
        None = the tokens of the last rollout, still resident on the device.  rows_per_img overrides set_refs' value for this call
        (sampled decode scores all batch_size * seq_per_img rows, greedy decode the unique ones)."""
        if seq is not None:
            seq = torch.as_tensor(seq).to(torch.int64).contiguous()
            if seq.dim() == 2:
                seq = seq.unsqueeze(0)
            n_members, rows, T = seq.shape
        else:
            T = self.ctx.dims['seq_length']
        out = torch.empty((int(n_members), int(rows)), dtype=torch.float64, device=device)
        _lib.check(self.ctx.lib.nes_cider_scores(self.ctx._h, _lib.ptr(seq), int(n_members), int(rows), int(T), int(rows_per_img or self.rows_per_img), _lib.ptr(out)))
        return out

    def compute_score(self, gts, res):
        """the reference's call (ciderD.py:28-52): gts {id: [ref strings]}, res [{'image_id': id, 'caption': [string]}];
        strings are space-separated token ids (array_to_str).  -> (mean, per-caption scores)"""
        def toks(s):
            return [int(w) for w in s.split()]
        cooked_gts = [[np.asarray(toks(r), dtype=np.int64) for r in gts[x['image_id']]] for x in res]
        T = max(1, max(len(toks(x['caption'][0])) for x in res))
        seq = np.zeros((len(res), T), dtype=np.int64)
        pad = np.full((len(res),), T, dtype=np.int64)
        for i, x in enumerate(res):
            t = toks(x['caption'][0])
            seq[i, :len(t)] = t
            pad[i] = len(t)
        # a caption that does not end with 0 and is shorter than T would be cut at a padding 0: keep such batches on T = own length
        assert all(pad[i] == T or seq[i, pad[i] - 1] == 0 for i in range(len(res))), 'captions must end with 0 or fill the row'
        self.set_refs(cooked_gts, seq_per_img=1, rows_per_img=1)
        s = self.scores(seq[None]).numpy()[0]
        return np.mean(s), s

    def fitness(self, n_members, rows, seq=None):
        """float(cider * 100) of captioning/policies.py:125 for every member: np.mean over the reference's batch (unique rows
        repeated seq_per_img times, like the replicated loader rows) in numpy, so the scalar is the reference's bit for bit"""
        s = self.scores(seq, n_members, rows).numpy()
        if self.seq_per_img > 1 and self.rows_per_img == 1:
            s = np.repeat(s, self.seq_per_img, axis=1)
        return np.array([float(np.mean(s[m]) * 100) for m in range(s.shape[0])], dtype=np.float64)
