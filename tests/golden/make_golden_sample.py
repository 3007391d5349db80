"""Generate tests/golden/golden_sample_v1.npz: the REAL reference's sampled decode -- FCModel._sample(fc, greedy=False),
src/captioning/nets.py:210-231 -- with the global numpy generator seeded, on the inputs of make_golden.py's small cases.

The uniforms the reference consumed are re-derivable: np.random.RandomState(seed).random_sample((T + 1, B)), row 0 being the
image step whose samples the reference throws away.  Run in the build container only:
    python tests/golden/make_golden_sample.py
"""
import os
import sys
import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
sys.path.insert(0, HERE)
from oracle import _refimport, noise  # noqa: E402
import make_golden as mg  # noqa: E402

SAMPLE_CASES = ('g1', 'g2', 'g5')
SEEDS = (301, 302)


def uniforms(seed, T, B):
    """[B, T]: the draw of row b at token step t = 1..T"""
    return np.random.RandomState(seed).random_sample((T + 1, B))[1:].T.copy()


def main():
    ref = _refimport.load()
    torch.set_grad_enabled(False)
    out = {}
    for name in SAMPLE_CASES:
        c = mg.CASES[name]
        dims, G, play, D, D_pad, L, theta, fc, offs = mg.case_inputs(c)
        opt = ref.ModelOptions(vocab_size=c['V'], input_encoding_size=c['E'], rnn_size=c['R'], fc_feat_size=c['Fd'],
                               layer_n=c.get('ln', False), layer_n_affine=c.get('affine', False))
        model = ref.FCModel(options=opt)
        eps = noise.noise_row(mg.TABLE_SEED, offs[0], play, D)
        seqs = np.zeros((len(SEEDS), 2, c['B'], 16), dtype=np.int16)
        lps = np.zeros((len(SEEDS), 2, c['B'], 16), dtype=np.float32)
        for k, seed in enumerate(SEEDS):
            for j, sign in enumerate((+1, -1)):
                th, _ = noise.member_theta(theta, eps, c['sigma'], sign)
                model.set_from_vector(torch.from_numpy(th.copy()))
                np.random.seed(seed)
                s, lp = model(torch.from_numpy(fc), greedy=False)
                seqs[k, j] = s.numpy(); lps[k, j] = lp.numpy()
        out[name + '_seq'] = seqs; out[name + '_logprobs'] = lps
        print(name, 'seq[0,0,0]', seqs[0, 0, 0])
    path = os.path.join(HERE, 'golden_sample_v1.npz')
    np.savez_compressed(path, **out)
    print('wrote', path, os.path.getsize(path), 'bytes')


if __name__ == '__main__':
    main()
