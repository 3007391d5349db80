"""Generate tests/golden/golden_vbne_v1.npz: the REAL reference's fc_caption with `vbn_e` -- img_embed and embed wrapped into
Sequential(layer, BatchNorm1d(track_running_stats=False, affine=vbn_affine)) (src/captioning/nets.py:156-165) -- alone, together with
the batch-norm cell (`vbn`) and together with the LayerNorm cell; greedy decode of theta +- sigma*eps members.  The affine parameters
(left at 1 / 0 by the reference's init, algorithm/nets.py:62-64) are moved off that point so that they carry information.
Run in the build container only:  python tests/golden/make_golden_vbne.py"""
import os
import sys
import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from oracle import _refimport, fc_caption, noise  # noqa: E402

CASES = {
    'e1': dict(V=300, E=128, R=128, Fd=512, B=12, sigma=0.02, vbn=False, affine=True),
    'e2': dict(V=300, E=128, R=128, Fd=512, B=12, sigma=0.02, vbn=True, affine=True),
    'e3': dict(V=520, E=64, R=192, Fd=128, B=150, sigma=0.02, vbn=False, affine=False),     # more rows than one 128-row block
    'e4': dict(V=300, E=64, R=64, Fd=128, B=9, sigma=0.02, vbn=False, affine=True, ln=True, ln_affine=True),
}
THETA_SEED, FC_SEED, TABLE_SEED, ITER, SLOTS = 17, 27, 4321, 5, 16


def make_dims(c):
    return fc_caption.Dims(c['V'], c['E'], c['R'], c['Fd'], layer_n=c.get('ln', False), layer_n_affine=c.get('ln_affine', False),
                           vbn=c['vbn'], vbn_affine=c['affine'], vbn_e=True)


def model_options(c):
    return dict(vocab_size=c['V'], input_encoding_size=c['E'], rnn_size=c['R'], fc_feat_size=c['Fd'], layer_n=c.get('ln', False),
                layer_n_affine=c.get('ln_affine', False), vbn=c['vbn'], vbn_e=True, vbn_affine=c['affine'])


def case_inputs(c):
    dims = make_dims(c)
    G = noise.granularity(c['E'], c['R'], c['Fd'])
    play, D, D_pad = noise.padded_layout(dims.layout(), G)
    L = D_pad + (SLOTS - 1) * G
    theta = fc_caption.xavier_theta(dims, THETA_SEED)
    for name, shape, start, cnt in dims.layout():
        if 'bn' in name or 'ln' in name or '.1.' in name:
            theta[start:start + cnt] += noise.table_slice(THETA_SEED + 100, start, cnt).astype(np.float32) * np.float32(0.15)
    fc = noise.table_slice(FC_SEED, 0, c['B'] * c['Fd']).astype(np.float32).reshape(c['B'], c['Fd'])
    offs = noise.sample_offsets(TABLE_SEED, ITER, 1, L, D_pad, G)
    return dims, G, play, D, D_pad, L, theta, fc, offs


def main():
    ref = _refimport.load()
    torch.set_grad_enabled(False)
    out = {}
    for name, c in CASES.items():
        dims, G, play, D, D_pad, L, theta, fc, offs = case_inputs(c)
        model = ref.FCModel(options=ref.ModelOptions(**model_options(c)), vbn=c['vbn'])
        assert model.nb_learnable_params == D
        assert [n for n, _ in model.named_parameters()] == [n for n, _, _, _ in dims.layout()], [n for n, _ in model.named_parameters()]
        eps = noise.noise_row(TABLE_SEED, offs[0], play, D)
        seqs = np.zeros((2, c['B'], 16), dtype=np.int16); lps = np.zeros((2, c['B'], 16), dtype=np.float32)
        for j, sign in enumerate((+1, -1)):
            th, _ = noise.member_theta(theta, eps, c['sigma'], sign)
            model.set_from_vector(torch.from_numpy(th.copy()))
            s, lp = model(torch.from_numpy(fc), greedy=True)
            seqs[j] = s.numpy(); lps[j] = lp.numpy()
            so, lo = fc_caption.decode_greedy(dims, th, fc)
            assert np.array_equal(so, seqs[j]) and np.array_equal(lo, lps[j]), 'oracle restatement differs from the reference'
        out[name + '_seq'] = seqs; out[name + '_logprobs'] = lps
        print(name, 'D', D, seqs[0, 0])
    path = os.path.join(HERE, 'golden_vbne_v1.npz')
    np.savez_compressed(path, **out)
    print('wrote', path, os.path.getsize(path), 'bytes')


if __name__ == '__main__':
    main()
