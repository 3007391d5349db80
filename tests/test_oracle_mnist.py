"""BASELINE configs[0] (experiments/mnist_es.json): the MNIST oracle (oracle/mnist.py) and the package's differentiable forward against
what the REAL reference's MnistNet / ClfPolicy / Sensitivity produced (tests/golden/make_golden_mnist.py)."""
import importlib
import os
import sys

import numpy as np
import torch

from oracle import mnist, noise

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, 'golden'))
import make_golden_mnist as mg  # noqa: E402

GOLD = np.load(os.path.join(HERE, 'golden', 'golden_mnist_v1.npz'))


def members():
    play, D, D_pad, L, theta, images, labels, offs = mg.case_inputs()
    out = [theta]
    for o in offs:
        eps = noise.noise_row(mg.TABLE_SEED, o, play, D)
        out += [noise.member_theta(theta, eps, mg.SIGMA, +1)[0], noise.member_theta(theta, eps, mg.SIGMA, -1)[0]]
    return out, images, labels


def test_layout_is_the_references():
    assert mnist.D == 8490 and [c for c in mnist.COUNTS] == [250, 10, 5000, 20, 3200, 10]
    lay = importlib.import_module('nes-img-captioning_b200.layout')
    tensors, D, D_pad = lay.build_layout(lay.mnist_tensors(), 64)
    assert D == 8490 and [(t.name, t.shape, t.start, t.count) for t in tensors] == mnist.layout()
    assert D_pad == noise.padded_layout(mnist.layout(), 64)[2] == 8704


def test_oracle_matches_reference_golden():
    ms, images, labels = members()
    assert len(ms) == GOLD['logits'].shape[0] == 7
    for k, th in enumerate(ms):
        assert np.allclose(mnist.forward(th, images), GOLD['logits'][k], rtol=1e-5, atol=1e-5)
        fit, cor = mnist.rollout(th, images, labels)
        assert abs(fit - GOLD['fitness'][k]) <= 1e-6 * abs(GOLD['fitness'][k]) and cor == GOLD['correct'][k]
    assert len(set(GOLD['fitness'].tolist())) == 7


def test_package_sum_sensitivity_matches_reference_golden():
    """SM-G-SUM as mnist_es.json configures it (underflow 0.2): the package's autograd path on the CPU against the reference's vector"""
    sm = importlib.import_module('nes-img-captioning_b200.sensitivity')
    lay = importlib.import_module('nes-img-captioning_b200.layout')
    play, D, D_pad, L, theta, images, labels, offs = mg.case_inputs()
    tensors, _, _ = lay.build_layout(lay.mnist_tensors(), 64)
    raw = sm.sum_sensitivity_of(torch.from_numpy(theta.copy()), lambda th, i: sm.mnist_forward(th, tensors, torch.from_numpy(images) if i < 0 else torch.from_numpy(images[i:i + 1])))
    got = raw.clone()
    got[got < mg.UNDERFLOW] = mg.UNDERFLOW
    got /= mg.UNDERFLOW
    want = GOLD['sens']
    assert 0.05 < (want == 1).mean() < 0.9
    assert np.allclose(got.numpy(), want, rtol=1e-5, atol=0)
