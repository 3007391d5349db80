"""Generate tests/golden/golden_vbn_v1.npz: the REAL reference's fc_caption with `vbn` (BatchNorm1d over the rows of the batch,
track_running_stats=False, src/captioning/nets.py:87-91,99-101,124-125), greedy decode of theta +- sigma*eps members.
The BN affine parameters (left at 1 / 0 by the reference's init) are moved by the perturbation like every other parameter.
Run in the build container only:  python tests/golden/make_golden_vbn.py"""
import os
import sys
import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from oracle import _refimport, fc_caption, noise  # noqa: E402

CASES = {
    'v1': dict(V=300, E=128, R=128, Fd=512, B=12, sigma=0.02, affine=True),
    'v2': dict(V=300, E=128, R=128, Fd=512, B=12, sigma=0.02, affine=False),
    'v3': dict(V=520, E=64, R=192, Fd=128, B=150, sigma=0.02, affine=True),      # more rows than one 128-row block
}
THETA_SEED, FC_SEED, TABLE_SEED, ITER, SLOTS = 13, 23, 4321, 3, 16


def case_inputs(c):
    dims = fc_caption.Dims(c['V'], c['E'], c['R'], c['Fd'], vbn=True, vbn_affine=c['affine'])
    G = noise.granularity(c['E'], c['R'], c['Fd'])
    play, D, D_pad = noise.padded_layout(dims.layout(), G)
    L = D_pad + (SLOTS - 1) * G
    theta = fc_caption.xavier_theta(dims, THETA_SEED)
    fc = noise.table_slice(FC_SEED, 0, c['B'] * c['Fd']).astype(np.float32).reshape(c['B'], c['Fd'])
    offs = noise.sample_offsets(TABLE_SEED, ITER, 1, L, D_pad, G)
    return dims, G, play, D, D_pad, L, theta, fc, offs


def main():
    ref = _refimport.load()
    torch.set_grad_enabled(False)
    out = {}
    for name, c in CASES.items():
        dims, G, play, D, D_pad, L, theta, fc, offs = case_inputs(c)
        opt = ref.ModelOptions(vocab_size=c['V'], input_encoding_size=c['E'], rnn_size=c['R'], fc_feat_size=c['Fd'],
                               layer_n=False, layer_n_affine=False, vbn=True, vbn_e=False, vbn_affine=c['affine'])
        model = ref.FCModel(options=opt, vbn=True)
        assert model.nb_learnable_params == D
        assert [n for n, _ in model.named_parameters()] == [n for n, _, _, _ in dims.layout()]
        eps = noise.noise_row(TABLE_SEED, offs[0], play, D)
        seqs = np.zeros((2, c['B'], 16), dtype=np.int16); lps = np.zeros((2, c['B'], 16), dtype=np.float32)
        for j, sign in enumerate((+1, -1)):
            th, _ = noise.member_theta(theta, eps, c['sigma'], sign)
            model.set_from_vector(torch.from_numpy(th.copy()))
            s, lp = model(torch.from_numpy(fc), greedy=True)
            seqs[j] = s.numpy(); lps[j] = lp.numpy()
        out[name + '_seq'] = seqs; out[name + '_logprobs'] = lps
        print(name, 'D', D, seqs[0, 0])
    path = os.path.join(HERE, 'golden_vbn_v1.npz')
    np.savez_compressed(path, **out)
    print('wrote', path, os.path.getsize(path), 'bytes')


if __name__ == '__main__':
    main()
