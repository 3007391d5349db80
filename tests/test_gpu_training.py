"""GPU: the whole path as the reference uses it -- NIC-NES iterations (nic_nes_master.py:120-137 + nic_nes_worker.py:131-154) with the
CIDEr-D training fitness scored on the device -- must actually learn: on a fixed synthetic batch the population's mean reward rises."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def test_nes_with_device_cider_improves_the_reward():
    import nes_b200
    V, E, R, Fd, B, F, sigma = 59, 64, 64, 64, 16, 64, 0.05
    ctx = nes_b200.NesContext(dict(vocab_size=V, input_encoding_size=E, rnn_size=R, fc_feat_size=Fd), seed=5, table_len=1 << 24, wave_members=2 * F, rows=B)
    nes_b200.FCModel(ctx, parent=0, rng_state=3)
    opt = nes_b200.Adam(ctx, None, stepsize=0.02)
    rs = np.random.RandomState(0)
    fc = torch.from_numpy(rs.randn(B, Fd).astype(np.float32)).cuda()
    gts = []
    for i in range(B):                       # every image has 3 references sharing most words: a learnable target
        base = rs.randint(1, V + 1, 8)
        refs = np.zeros((3, 16), dtype=np.int64)
        for j in range(3):
            r = base.copy()
            r[rs.randint(8)] = rs.randint(1, V + 1)
            refs[j, :8] = r
        gts.append(refs)
    scorer = nes_b200.CiderD(ctx, df='corpus').set_refs(gts)

    def fitness_fn(seq, lp):
        return torch.from_numpy(scorer.fitness(seq.shape[0], B))          # scores the tokens the rollout left on the device

    means = []
    for it in range(30):
        _, fit = nes_b200.nes_iteration(ctx, opt, it, F, fc, sigma, 0.0, fitness_fn, precision=nes_b200.PREC_HYBRID)
        means.append(float(fit.mean()))
    first, last = np.mean(means[:3]), np.mean(means[-3:])
    print('mean CIDEr-D*100 of the population: first iterations %.3f, last %.3f' % (first, last))
    assert last > first + 1.0 and last > 2.0 * max(first, 0.1), means
    ctx.close()
