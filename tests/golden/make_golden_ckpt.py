"""Generate tests/golden/ckpt_v1/: checkpoint files written by the REAL reference (imported from /root/reference) for a tiny
fc_caption model, so the loaders of this package can be tested against the reference's on-disk formats:

  model.pth       PolicyNet.serialize        (src/algorithm/nets.py:153-155: torch.save(self.state_dict()))
  adam.tar        Adam.save_to_file          (src/algorithm/nic_nes/optimizers.py:85-96) after two updates of a flat 1000-long theta
  sgd.tar         SGD.save_to_file           (optimizers.py:49-58) after two updates
  expect.npz      theta, the two gradients, theta after each update, and theta after one MORE update from the saved state

Run in the build container only:  python tests/golden/make_golden_ckpt.py
"""
import os
import sys
import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from oracle import _refimport  # noqa: E402

DIMS = dict(vocab_size=49, input_encoding_size=64, rnn_size=64, fc_feat_size=64)


def main():
    ref = _refimport.load()
    torch.set_grad_enabled(False)
    out_dir = os.path.join(HERE, 'ckpt_v1')
    os.makedirs(out_dir, exist_ok=True)
    torch.manual_seed(5)
    model = ref.FCModel(options=ref.ModelOptions(layer_n=False, layer_n_affine=False, **DIMS))
    model.serialize(path=os.path.join(out_dir, 'model.pth'))
    mtheta = model.parameter_vector().numpy().copy()
    # the model file is checked through a strided sample and two checksums (the full vector would double the fixture)
    exp = {'model_theta_s16': mtheta[::16].copy(), 'model_theta_sum': np.float64(mtheta.astype(np.float64).sum()),
           'model_theta_abs': np.float64(np.abs(mtheta.astype(np.float64)).sum()), 'model_D': np.int64(mtheta.size)}
    # optimizer files on a flat 1000-long theta (the optimizers only see a flat vector, optimizers.py:8-13)
    rs = np.random.RandomState(8)
    theta = (rs.randn(1000) * 0.1).astype(np.float32)
    grads = [(rs.randn(theta.size) * 0.01).astype(np.float32) for _ in range(3)]
    exp.update({'theta': theta, 'g0': grads[0], 'g1': grads[1], 'g2': grads[2]})
    for name, cls, kw in (('adam', ref.optimizers.Adam, dict(stepsize=0.002)), ('sgd', ref.optimizers.SGD, dict(stepsize=0.05, momentum=0.9))):
        opt = cls(theta.copy(), **kw)
        th = theta.copy()
        for i in range(2):
            _, th = opt.update(-grads[i] + 1e-7 * th)
        opt.save_to_file(os.path.join(out_dir, name + '.tar'))
        exp[name + '_theta2'] = np.asarray(th, dtype=np.float64)
        _, th3 = opt.update(-grads[2] + 1e-7 * th)
        exp[name + '_theta3'] = np.asarray(th3, dtype=np.float64)
    np.savez_compressed(os.path.join(out_dir, 'expect.npz'), **exp)
    for f in sorted(os.listdir(out_dir)):
        print(f, os.path.getsize(os.path.join(out_dir, f)))


if __name__ == '__main__':
    main()
