"""Generate tests/golden/golden_sens_v1.npz: the REAL reference's output-gradient sensitivity (SM-G-SUM,
src/algorithm/safe_mutations.py:34-115 over FCModel.forward_for_sensitivity, src/captioning/nets.py:22-69) for seeded
thetas / features, as `PolicyNet.calc_sensitivity` leaves it in memory and in the sens_t*_p*_*.pt file, plus the mutation
`evolve` then applies with torch's generator seeded.  SM-G-ABS has no golden: the reference's method raises AttributeError
(safe_mutations.py:125).  Run in the build container only:  python tests/golden/make_golden_sens.py"""
import os
import sys
import tempfile
import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from oracle import _refimport, fc_caption, noise  # noqa: E402

# rows = images in the batch (fc_feats carries 5 copies of each, :41); bs = orig_batch_size (only the first bs images are used, :43-44)
CASES = {
    's1': dict(V=230, E=32, R=32, Fd=64, rows=6, bs=4, underflow=1e-4, scale=1.0),
    's2': dict(V=299, E=32, R=64, Fd=128, rows=5, bs=8, underflow=0.01, scale=1.0, ln=True, affine=True),
    's3': dict(V=230, E=32, R=32, Fd=64, rows=6, bs=6, underflow=0.01, scale=1.0, vbn=True, affine=True),
    # the CUDA library wants E, R, Fd in multiples of 64: the two cases the device test uses
    's5': dict(V=99, E=64, R=64, Fd=64, rows=5, bs=4, underflow=1e-4, scale=1.5),
    's6': dict(V=130, E=64, R=64, Fd=128, rows=4, bs=4, underflow=0.01, scale=1.0, ln=True, affine=True),
    's4': dict(V=199, E=64, R=32, Fd=64, rows=3, bs=3, underflow=0.01, scale=4.0),     # V1 % 100 == 0: the all-zero padding group (:62-64)
}
THETA_SEED, FC_SEED = 17, 29


def model_options(c):
    return dict(vocab_size=c['V'], input_encoding_size=c['E'], rnn_size=c['R'], fc_feat_size=c['Fd'],
                layer_n=bool(c.get('ln')) and not c.get('vbn'), layer_n_affine=bool(c.get('affine')) and bool(c.get('ln')),
                vbn=bool(c.get('vbn')), vbn_e=False, vbn_affine=bool(c.get('affine')) and bool(c.get('vbn')),
                safe_mutations='SM-G-SUM', safe_mutation_underflow=c['underflow'])


def case_inputs(c):
    mo = model_options(c)
    dims = fc_caption.Dims(c['V'], c['E'], c['R'], c['Fd'], layer_n=mo['layer_n'], layer_n_affine=mo['layer_n_affine'],
                           vbn=mo['vbn'], vbn_affine=mo['vbn_affine'])
    theta = (fc_caption.xavier_theta(dims, THETA_SEED) * np.float32(c['scale'])).astype(np.float32)
    fc = noise.table_slice(FC_SEED, 0, c['rows'] * c['Fd']).astype(np.float32).reshape(c['rows'], c['Fd'])
    return dims, theta, np.repeat(fc, 5, axis=0)


def main():
    ref = _refimport.load()
    out = {}
    for name, c in CASES.items():
        dims, theta, fc5 = case_inputs(c)
        model = ref.FCModel(options=ref.ModelOptions(**model_options(c)), vbn=bool(c.get('vbn')))
        model.set_from_vector(torch.from_numpy(theta.copy()))
        with tempfile.TemporaryDirectory() as d:
            model.calc_sensitivity(3, 1, {'fc_feats': fc5}, c['bs'], d)
            files = os.listdir(d)
            assert len(files) == 1 and files[0] == 'sens_t3_p1_%d.pt' % os.getpid()
            sens = model.get_sensitivity_vector().numpy().copy()
            assert np.array_equal(torch.load(os.path.join(d, files[0])).numpy(), sens)
        torch.manual_seed(5)
        raw = torch.empty(theta.size).normal_(mean=0.0, std=0.05).numpy().copy()
        torch.manual_seed(5)
        got = model.evolve(0.05)
        assert np.array_equal(got, raw / sens)
        out[name + '_sens'] = sens
        print(name, 'D', theta.size, 'clamped %.3f' % (sens == 1).mean(), 'max %.1f' % sens.max(), 'nan', int(np.isnan(sens).sum()))
    path = os.path.join(HERE, 'golden_sens_v1.npz')
    np.savez_compressed(path, **out)
    print('wrote', path, os.path.getsize(path), 'bytes')


if __name__ == '__main__':
    main()
