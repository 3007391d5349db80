"""Generate tests/golden/golden_v1.npz by running the REAL reference
(/root/reference, imported through oracle/_refimport.py) on inputs that are
fully determined by small integers, so the fixture stays tiny:

  theta  = oracle.fc_caption.xavier_theta(dims, theta_seed)
  fc     = float32(table_slice(fc_seed, 0, B*Fd))            (N(0,1) features)
  eps_i  = noise row at sample_offsets(table_seed, iteration, F, L, D_pad, G)[i]
  member = theta +/- float32(sigma) * eps_i                   (reference formula)

Outputs stored: the reference FCModel's seq / seq_logprobs for every member,
the reference NESMaster centred ranks / gradient estimate and the reference
Adam / SGD trajectories.  Run in the build container only:
    python tests/golden/make_golden.py
"""
import os
import sys
import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import _refimport, fc_caption, noise  # noqa: E402

CASES = {
    # name: V, E, R, Fd, B, F, sigma, ln, affine
    'g1': dict(V=199, E=64, R=64, Fd=128, B=8, F=3, sigma=0.01),
    'g2': dict(V=999, E=64, R=128, Fd=256, B=5, F=2, sigma=0.005),
    'g3': dict(V=9487, E=128, R=128, Fd=2048, B=16, F=2, sigma=0.01),
    'g4': dict(V=9487, E=512, R=512, Fd=2048, B=128, F=1, sigma=0.005),
    'g5': dict(V=300, E=128, R=128, Fd=512, B=12, F=2, sigma=0.01, ln=True, affine=True),
    'g6': dict(V=300, E=128, R=128, Fd=512, B=12, F=2, sigma=0.01, ln=True, affine=False),
}
THETA_SEED, FC_SEED, TABLE_SEED, ITER = 11, 22, 1234, 7
SLOTS = 64   # table_len = D_pad + (SLOTS-1)*G


def case_inputs(c):
    dims = fc_caption.Dims(c['V'], c['E'], c['R'], c['Fd'], layer_n=c.get('ln', False),
                           layer_n_affine=c.get('affine', False))
    G = noise.granularity(c['E'], c['R'], c['Fd'])
    play, D, D_pad = noise.padded_layout(dims.layout(), G)
    L = D_pad + (SLOTS - 1) * G
    theta = fc_caption.xavier_theta(dims, THETA_SEED)
    fc = noise.table_slice(FC_SEED, 0, c['B'] * c['Fd']).astype(np.float32).reshape(c['B'], c['Fd'])
    offs = noise.sample_offsets(TABLE_SEED, ITER, c['F'], L, D_pad, G)
    return dims, G, play, D, D_pad, L, theta, fc, offs


def main():
    ref = _refimport.load()
    torch.set_grad_enabled(False)
    out = {}
    for name, c in CASES.items():
        dims, G, play, D, D_pad, L, theta, fc, offs = case_inputs(c)
        opt = ref.ModelOptions(vocab_size=c['V'], input_encoding_size=c['E'], rnn_size=c['R'],
                               fc_feat_size=c['Fd'], layer_n=c.get('ln', False),
                               layer_n_affine=c.get('affine', False))
        model = ref.FCModel(options=opt)
        assert model.nb_learnable_params == D
        seqs = np.zeros((c['F'], 2, c['B'], 16), dtype=np.int16)
        lps = np.zeros((c['F'], 2, c['B'], 16), dtype=np.float32)
        for i in range(c['F']):
            eps = noise.noise_row(TABLE_SEED, offs[i], play, D)
            for j, sign in enumerate((+1, -1)):
                th, _ = noise.member_theta(theta, eps, c['sigma'], sign)
                model.set_from_vector(torch.from_numpy(th.copy()))
                s, lp = model(torch.from_numpy(fc), greedy=True)
                seqs[i, j] = s.numpy(); lps[i, j] = lp.numpy()
        model.set_from_vector(torch.from_numpy(theta.copy()))
        s0, lp0 = model(torch.from_numpy(fc), greedy=True)
        out[name + '_seq'] = seqs; out[name + '_logprobs'] = lps
        out[name + '_seq_theta'] = s0.numpy().astype(np.int16); out[name + '_logprobs_theta'] = lp0.numpy()
        out[name + '_offsets'] = offs
        out[name + '_theta_sum'] = np.float64(theta.astype(np.float64).sum())
        print(name, 'D', D, 'D_pad', D_pad, 'L', L, 'offs', offs, 'seq[0,0,0]', seqs[0, 0, 0])

    # ---- master math: reference NESMaster on a small noise stack built from table rows
    m = object.__new__(ref.NESMaster)
    c = CASES['g1']
    dims, G, play, D, D_pad, L, theta, fc, _ = case_inputs(c)
    F = 24
    offs = noise.sample_offsets(TABLE_SEED, ITER + 1, F, L, D_pad, G)
    rs = np.random.RandomState(5)
    fit = rs.randn(F, 2) * 3 + 50          # tie-free
    sigma = 0.01
    nz = np.stack([noise.member_theta(theta, noise.noise_row(TABLE_SEED, o, play, D), sigma, +1)[1] for o in offs])
    out['mm_fitness'] = fit
    out['mm_offsets'] = offs
    out['mm_centered_ranks'] = m.compute_centered_ranks(fit)
    g = m.gradient_estimate(fit, nz)
    out['mm_grad'] = g
    out['mm_kat_in'] = np.array([[101, 200], [2, 100]], dtype=np.float64)
    out['mm_kat_out'] = m.compute_centered_ranks(np.array([[101, 200], [2, 100]]))
    # ---- optimizer trajectories on that gradient, l2coeff as shipped (mscoco_nes.json: 1e-7)
    for oname, kw in (('adam', dict(stepsize=0.001)), ('sgd', dict(stepsize=0.02, momentum=0.9))):
        cls = {'adam': ref.optimizers.Adam, 'sgd': ref.optimizers.SGD}[oname]
        opt_r = cls(theta.copy(), **kw)
        th = theta.copy()
        ratios, thetas = [], []
        for step in range(3):
            gi = (g * np.float32(1.0 + 0.5 * step)).astype(np.float32)
            ratio, th = opt_r.update(-gi + 1e-7 * th)
            ratios.append(ratio); thetas.append(np.asarray(th, dtype=np.float64).copy())
        out['opt_%s_ratios' % oname] = np.array(ratios)
        # full float64 vectors would be 0.6 MB each: keep every 16th element plus two checksums
        for tag, v in (('theta1', thetas[0]), ('theta3', thetas[-1])):
            out['opt_%s_%s_s16' % (oname, tag)] = v[::16].copy()
            out['opt_%s_%s_sum' % (oname, tag)] = np.float64(v.sum())
            out['opt_%s_%s_norm' % (oname, tag)] = np.float64(np.linalg.norm(v))
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden_v1.npz')
    np.savez_compressed(path, **out)
    print('wrote', path, os.path.getsize(path), 'bytes')


if __name__ == '__main__':
    main()
