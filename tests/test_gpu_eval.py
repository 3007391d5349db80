"""Validation decode with the reference's interface (src/captioning/eval_utils.py:60-107, src/captioning/policies.py:130-143):
eval_split over a loader that follows the reference's DataLoader protocol, against the oracle's greedy decode."""
from collections import namedtuple

import numpy as np
import pytest

from oracle import fc_caption

pytestmark = pytest.mark.gpu


class FakeLoader:
    """the slice of captioning/dataloader.py's interface eval_split uses: 5x replicated rows, bounds, infos, vocab"""

    def __init__(self, n_img, Fd, batch_size, seq_per_img, V):
        rs = np.random.RandomState(0)
        self.feats = rs.randn(n_img, Fd).astype(np.float32)
        self.batch_size, self.seq_per_img, self.n_img = batch_size, seq_per_img, n_img
        self.vocab = {str(i): 'w%d' % i for i in range(1, V + 1)}
        self.pos = 0

    def reset_iterator(self, split):
        self.pos = 0

    def get_vocab(self):
        return self.vocab

    def get_batch(self, split):
        ids = [(self.pos + k) % self.n_img for k in range(self.batch_size)]
        wrapped = self.pos + self.batch_size >= self.n_img
        self.pos = (self.pos + self.batch_size) % self.n_img
        return {'fc_feats': np.repeat(self.feats[ids], self.seq_per_img, axis=0), 'infos': [{'id': 1000 + i} for i in ids],
                'gts': [np.zeros((1, 16), dtype=np.int64) for _ in ids], 'bounds': {'it_max': self.n_img, 'wrapped': wrapped}}


def test_eval_split_and_accuracy_on_against_oracle(tmp_path):
    import nes_b200
    V, E, R, Fd = 99, 64, 64, 64
    exp = dict(policy_options=dict(net='fc_caption', fitness='greedy', model_options=dict(vocab_size=V, input_encoding_size=E, rnn_size=R, fc_feat_size=Fd)))
    pol = nes_b200.PolicyFactory.create(nes_b200.SuppDataset.MSCOCO, exp, seed=2, table_len=1 << 20, wave_members=2, rows=4)
    loader = FakeLoader(n_img=10, Fd=Fd, batch_size=4, seq_per_img=5, V=V)          # 3 batches, the last one wraps around
    stats, preds = nes_b200.eval_split(pol.policy_net, loader, str(tmp_path), do_eval=False)
    assert stats is None and [p['image_id'] for p in preds] == [1000 + i for i in range(10)]
    dims = fc_caption.Dims(V, E, R, Fd)
    seq, _ = fc_caption.decode_greedy(dims, pol.parameter_vector().numpy(), loader.feats)
    assert [p['caption'] for p in preds] == nes_b200.decode_sequence(loader.vocab, seq)
    assert any(p['caption'].startswith('w') for p in preds)
    _, some = nes_b200.eval_split(pol.policy_net, loader, str(tmp_path), num=6, do_eval=False)
    assert len(some) == 6
    # accuracy_on = the injected evaluator's CIDEr of those predictions (the reference's is pycocoevalcap / Java)
    Holder = namedtuple('Holder', ['loader'])
    Cfg = namedtuple('Cfg', ['num_val_items'])
    with pytest.raises(NotImplementedError):
        pol.accuracy_on(Holder(loader), Cfg(-1), str(tmp_path))
    seen = {}
    pol.language_eval = lambda predictions, directory, split: seen.update(n=len(predictions)) or {'CIDEr': 0.25 * len(predictions)}
    assert pol.accuracy_on(Holder(loader), Cfg(8), str(tmp_path)) == 2.0 and seen['n'] == 8
    pol.save(str(tmp_path / 'snap'), 'm.pth')
    import torch
    assert list(torch.load(str(tmp_path / 'snap' / 'm.pth')).keys())[0] == 'img_embed.weight'
    pol.ctx.close()
