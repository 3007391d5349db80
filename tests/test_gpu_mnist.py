"""BASELINE configs[0] on the device: the MNIST population rollout (nes_mnist_rollout, csrc/mnist.cu) through the C ABI against the REAL
reference's goldens and the pinned oracle; ClfPolicy / MnistNet with the reference's interface; SM-G-SUM as mnist_es.json configures it."""
import os
import sys

import numpy as np
import pytest
import torch

from oracle import mnist, noise

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, 'golden'))
import make_golden_mnist as mg  # noqa: E402

GOLD = np.load(os.path.join(HERE, 'golden', 'golden_mnist_v1.npz'))


def _ctx(L, n_parents=1):
    import nes_b200
    return nes_b200.NesContext(tensors=nes_b200.layout.mnist_tensors(), flat_granule=mg.G, seed=mg.TABLE_SEED, table_len=L, n_parents=n_parents)


def test_population_rollout_matches_reference_golden():
    play, D, D_pad, L, theta, images, labels, offs = mg.case_inputs()
    ctx = _ctx(L)
    assert ctx.D == D and ctx.D_pad == D_pad
    ctx.set_theta(theta)
    assert np.array_equal(ctx.sample_offsets(mg.ITER, mg.F_PAIRS).numpy(), offs)
    offs2 = np.repeat(offs, 2)
    signs = np.tile(np.array([1.0, -1.0], dtype=np.float32), len(offs))
    n0 = ctx.gpu_launches
    fit, cor, lg = ctx.mnist_rollout(offs2, signs, images, labels, mg.SIGMA, want_logits=True)
    assert ctx.gpu_launches - n0 == 2
    assert fit.dtype == torch.float64 and cor.dtype == torch.int32 and tuple(lg.shape) == (6, mg.ROWS, 10)
    assert np.allclose(lg.numpy(), GOLD['logits'][1:], rtol=1e-4, atol=1e-4)
    assert np.allclose(fit.numpy(), GOLD['fitness'][1:], rtol=1e-5, atol=0)
    assert np.array_equal(cor.numpy(), GOLD['correct'][1:])
    # the members the kernel rebuilds in shared memory are nes_evolve's: same logits from the materialised vector at sigma = 0
    member, _ = ctx.evolve(int(offs[1]), mg.SIGMA, -1.0)
    ctx2 = _ctx(L)
    ctx2.set_theta(member)
    f2, c2, l2 = ctx2.mnist_rollout([0], [1.0], images, labels, 0.0, want_logits=True)
    assert np.array_equal(l2[0].numpy(), lg[3].numpy()) and f2[0].item() == fit[3].item()
    # sigma = 0: theta itself; device-resident inputs and outputs
    f0, c0 = ctx.mnist_rollout(torch.zeros(1, dtype=torch.int64, device=ctx.dev), torch.ones(1, device=ctx.dev), torch.from_numpy(images).to(ctx.dev),
                               torch.from_numpy(labels).to(ctx.dev), 0.0, out_device=ctx.dev)
    assert f0.is_cuda and abs(f0[0].item() - GOLD['fitness'][0]) <= 1e-5 * abs(GOLD['fitness'][0]) and c0[0].item() == GOLD['correct'][0]
    # empty population, bad offset
    fe, ce = ctx.mnist_rollout(np.zeros(0, dtype=np.int64), np.zeros(0, dtype=np.float32), images, labels, mg.SIGMA)
    assert fe.numel() == 0
    import nes_b200
    with pytest.raises(nes_b200.NesError):
        ctx.mnist_rollout([3], [1.0], images, labels, mg.SIGMA)
    ctx.close(); ctx2.close()


def test_per_member_batches_parents_and_many_rows():
    """every member on its own batch (mscoco_es.json style single_batch = false), two parents, 150 rows, against the oracle"""
    play, D, D_pad, L, theta, _, _, offs = mg.case_inputs()
    ctx = _ctx(L, n_parents=2)
    thetas = [theta, mnist.seeded_theta(7, scale=1.5)]
    ctx.set_theta(thetas[0], 0); ctx.set_theta(thetas[1], 1)
    P, rows = 5, 150
    rs = np.random.RandomState(3)
    images = rs.randn(P, rows, 1, 28, 28).astype(np.float32)
    labels = rs.randint(0, 10, (P, rows)).astype(np.int64)
    parents = np.array([0, 1, 1, 0, 1], dtype=np.int32)
    signs = np.array([1, -1, 1, -1, -1], dtype=np.float32)
    o5 = np.array([offs[0], offs[1], offs[2], offs[0], offs[0]], dtype=np.int64)
    fit, cor = ctx.mnist_rollout(o5, signs, images, labels, 0.1, parents=parents)
    for i in range(P):
        th, _ = noise.member_theta(thetas[parents[i]], noise.noise_row(mg.TABLE_SEED, o5[i], play, D), 0.1, int(signs[i]))
        f, c = mnist.rollout(th, images[i], labels[i])
        assert abs(fit[i].item() - f) <= 1e-5 * abs(f) and cor[i].item() == c, (i, fit[i].item(), f, cor[i].item(), c)
    ctx.close()


def test_clf_policy_interface_and_gradient_safe_mutation(tmp_path):
    """PolicyFactory.create(SuppDataset.MNIST, mnist_es.json's policy_options) -> ClfPolicy: rollout / accuracy_on / calc_sensitivity +
    evolve_model (nic_es_worker.py:172-175) with SM-G-SUM at underflow 0.2, and a population rollout under that mutation"""
    import nes_b200
    play, D, D_pad, L, theta, images, labels, offs = mg.case_inputs()
    exp = {'policy_options': {'net': 'mnist', 'vbn': False, 'model_options': {'safe_mutation_underflow': mg.UNDERFLOW, 'safe_mutations': 'SM-G-SUM'}}}
    policy = nes_b200.PolicyFactory.create(nes_b200.SuppDataset.MNIST, exp, flat_granule=mg.G, seed=mg.TABLE_SEED, table_len=L)
    assert isinstance(policy, nes_b200.ClfPolicy) and policy.nb_learnable_params() == 8490
    init = policy.parameter_vector().numpy()
    assert np.all(init[5260:5280] == 0) and np.abs(init[250:260]).max() > 0 and np.abs(init[:250]).max() <= 0.2 + 1e-6     # conv2.bias zero, conv1 default init
    policy.set_from_parameter_vector(theta)
    data = (torch.from_numpy(images), torch.from_numpy(labels))
    fit = policy.rollout(None, data, None)
    assert isinstance(fit, float) and abs(fit - GOLD['fitness'][0]) <= 1e-5 * abs(GOLD['fitness'][0])
    cfg = type('Cfg', (), {'num_val_batches': 0})()
    acc = policy.accuracy_on([data, (data[0][:7], data[1][:7])], cfg, None)
    c7 = mnist.rollout(theta, images[:7], labels[:7])[1]
    assert abs(acc - np.mean([GOLD['correct'][0] / mg.ROWS, c7 / 7])) < 1e-12
    assert np.allclose(policy.get_model()(data[0]).numpy(), GOLD['logits'][0], rtol=1e-4, atol=1e-4)
    policy.calc_sensitivity(1, 0, data, mg.SENS_BS, str(tmp_path))
    sens = policy.get_model().get_sensitivity_vector().numpy()
    assert np.allclose(sens, GOLD['sens'], rtol=1e-3, atol=1e-3)
    assert os.listdir(tmp_path) == ['sens_t1_p0_%d.pt' % os.getpid()]
    fitp, corp = policy.rollout_population(offs[:2], [1.0, -1.0], data, 0.3)
    for i, sg in enumerate((+1, -1)):
        th, _ = noise.member_theta(theta, noise.noise_row(mg.TABLE_SEED, offs[i], play, D), 0.3, sg, sens=sens)
        f, c = mnist.rollout(th, images, labels)
        assert abs(fitp[i].item() - f) <= 1e-5 * abs(f) and corp[i].item() == c
    nz = policy.evolve_model(0.3)
    off = policy.get_model().last_offset
    want_m, want_nz = noise.member_theta(theta, noise.noise_row(mg.TABLE_SEED, off, play, D), 0.3, +1, sens=sens)
    assert np.array_equal(nz, want_nz) and np.array_equal(policy.parameter_vector().numpy(), want_m)
    # checkpoint round trip with the reference's state_dict names
    path = policy.serialized(str(tmp_path / 'm.pth'))
    sd = torch.load(path)
    assert list(sd.keys()) == [n for n, _ in nes_b200.layout.mnist_tensors()] and tuple(sd['conv2.weight'].shape) == (20, 10, 5, 5)
    policy.ctx.close()
