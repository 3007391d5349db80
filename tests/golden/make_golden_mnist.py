"""Generate tests/golden/golden_mnist_v1.npz: the REAL reference's MnistNet / ClfPolicy (src/classification/nets.py:8-37,
src/classification/policies.py:16-86) on seeded inputs: logits, the rollout fitness (-CrossEntropyLoss) and the number of correct
predictions for theta and for theta +- sigma*eps members built from the noise-table spec, plus the SM-G-SUM sensitivity of theta
(safe_mutations.py:34-115 over MnistNet.forward_for_sensitivity, as experiments/mnist_es.json configures it).
Run in the build container only:  python tests/golden/make_golden_mnist.py"""
import os
import sys
import tempfile
import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from oracle import _refimport, mnist, noise  # noqa: E402

THETA_SEED, IMG_SEED, TABLE_SEED, ITER, SLOTS, G = 41, 43, 777, 5, 32, 64
ROWS, SIGMA, F_PAIRS, UNDERFLOW, SENS_BS = 19, 0.05, 3, 0.2, 19          # 19 rows: not a multiple of the 8 images a block takes


def padded_layout():
    return noise.padded_layout(mnist.layout(), G)


def case_inputs():
    play, D, D_pad = padded_layout()
    L = D_pad + (SLOTS - 1) * G
    theta = mnist.seeded_theta(THETA_SEED, scale=2.0)
    images = noise.table_slice(IMG_SEED, 0, ROWS * 784).astype(np.float32).reshape(ROWS, 1, 28, 28)
    labels = np.random.RandomState(IMG_SEED).randint(0, 10, ROWS).astype(np.int64)
    offs = noise.sample_offsets(TABLE_SEED, ITER, F_PAIRS, L, D_pad, G)
    return play, D, D_pad, L, theta, images, labels, offs


def main():
    ref = _refimport.load()
    from classification.nets import MnistNet
    from classification.policies import ClfPolicy
    from algorithm.policies import PolicyOptions, SuppDataset, Net
    play, D, D_pad, L, theta, images, labels, offs = case_inputs()
    assert D == 8490
    opts = ref.ModelOptions(safe_mutations='SM-G-SUM', safe_mutation_underflow=UNDERFLOW)
    net = MnistNet(options=opts)
    assert [n for n, _ in net.named_parameters()] == [n for n, _, _, _ in mnist.layout()]
    policy = object.__new__(ClfPolicy)
    policy.policy_net, policy.vbn = net, False
    cfg = type('Cfg', (), {'cuda': False})()
    out = {}
    members = [theta]
    for o in offs:
        eps = noise.noise_row(TABLE_SEED, o, play, D)
        for sign in (+1, -1):
            members.append(noise.member_theta(theta, eps, SIGMA, sign)[0])
    logits, fits, cors = [], [], []
    for th in members:
        net.set_from_vector(torch.from_numpy(th.copy()))
        with torch.no_grad():
            lg = net(torch.from_numpy(images))
        fit = policy.rollout(torch.empty(images.shape), (torch.from_numpy(images), torch.from_numpy(labels)), cfg)
        logits.append(lg.numpy().copy()); fits.append(fit)
        cors.append(int(lg.argmax(dim=1).eq(torch.from_numpy(labels)).sum()))
    out['logits'] = np.stack(logits); out['fitness'] = np.array(fits, dtype=np.float64); out['correct'] = np.array(cors, dtype=np.int32)
    # SM-G-SUM sensitivity of theta on the first SENS_BS images
    net.set_from_vector(torch.from_numpy(theta.copy()))
    with tempfile.TemporaryDirectory() as d:
        net.calc_sensitivity(1, 0, (torch.from_numpy(images), torch.from_numpy(labels)), SENS_BS, d)
        out['sens'] = net.get_sensitivity_vector().numpy().copy()
    torch.set_grad_enabled(False)
    print('fitness', out['fitness'], 'correct', out['correct'], 'sens clamped %.3f max %.1f' % ((out['sens'] == 1).mean(), out['sens'].max()))
    path = os.path.join(HERE, 'golden_mnist_v1.npz')
    np.savez_compressed(path, **out)
    print('wrote', path, os.path.getsize(path), 'bytes')


if __name__ == '__main__':
    main()
