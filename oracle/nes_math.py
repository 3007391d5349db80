"""Rank shaping, NES gradient estimate, SGD / Adam -- numpy restatement.

TEST INFRASTRUCTURE (see oracle/__init__.py).  PINNED against the imported
reference (oracle/refcheck.py) and its one published known answer, the
docstring example at src/algorithm/nic_nes/nic_nes_master.py:189.

Follows
  * NESMaster.compute_ranks / compute_centered_ranks  nic_nes_master.py:184-205
  * NESMaster.gradient_estimate / batched_weighted_sum nic_nes_master.py:170-182,207-221
  * Optimizer.update, SGD._compute_step, Adam._compute_step  nic_nes/optimizers.py:15-22,44-47,78-83
  * the L2 term and sign convention of run_master          nic_nes_master.py:123-133
"""
import numpy as np


def compute_ranks(x, stable=True):
    """ranks in [0, len(x)).  The reference uses numpy's default (unstable)
    argsort; ties are therefore build dependent there.  `stable=True` is the
    documented rule of this build: ties broken by index."""
    x = np.asarray(x)
    assert x.ndim == 1
    order = np.argsort(x, kind='stable' if stable else None)
    ranks = np.empty(len(x), dtype=np.int64)
    ranks[order] = np.arange(len(x))
    return ranks


def compute_centered_ranks(x, stable=True):
    x = np.asarray(x)
    y = compute_ranks(x.ravel(), stable).reshape(x.shape).astype(np.float64)
    y /= (x.size - 1)
    y -= .5
    return y


def rank_weights(fitnesses, stable=True):
    """w_i = r(f_i+) - r(f_i-)  (float64 [F])."""
    r = compute_centered_ranks(np.asarray(fitnesses, dtype=np.float64), stable)
    return r[:, 0] - r[:, 1]


def gradient_estimate(fitnesses, noise_vecs, batch_size=500):
    """g = sum_i w_i * noise_i / (2F), float32 accumulation over chunks of
    `batch_size` rows exactly like the reference."""
    w = rank_weights(fitnesses)
    total = 0.
    for lo in range(0, len(w), batch_size):
        total = total + np.dot(np.asarray(w[lo:lo + batch_size], dtype=np.float32),
                               np.asarray(noise_vecs[lo:lo + batch_size], dtype=np.float32))
    total /= np.asarray(fitnesses).size
    return total


class Optimizer:
    def __init__(self, theta):
        self.theta = theta
        self.dim = len(theta)
        self.t = 0

    def update(self, globalg):
        self.t += 1
        step = self._compute_step(globalg)
        ratio = np.linalg.norm(step) / np.linalg.norm(self.theta)
        self.theta = self.theta + step
        return ratio, self.theta


class SGD(Optimizer):
    def __init__(self, theta, stepsize, momentum=0.9):
        super().__init__(theta)
        self.v = np.zeros(self.dim, dtype=np.float64)
        self.stepsize, self.momentum = stepsize, momentum

    def _compute_step(self, g):
        self.v = self.momentum * self.v + (1. - self.momentum) * g
        return -self.stepsize * self.v


class Adam(Optimizer):
    def __init__(self, theta, stepsize, beta1=0.9, beta2=0.999, epsilon=1e-08):
        super().__init__(theta)
        self.stepsize, self.beta1, self.beta2, self.epsilon = stepsize, beta1, beta2, epsilon
        self.m = np.zeros(self.dim, dtype=np.float64)
        self.v = np.zeros(self.dim, dtype=np.float64)

    def _compute_step(self, g):
        a = self.stepsize * np.sqrt(1 - self.beta2 ** self.t) / (1 - self.beta1 ** self.t)
        self.m = self.beta1 * self.m + (1 - self.beta1) * g
        self.v = self.beta2 * self.v + (1 - self.beta2) * (g * g)
        return -a * self.m / (np.sqrt(self.v) + self.epsilon)


def master_update(optimizer, theta_now, grad_estimate, l2coeff):
    """One master step: G = -g + l2*theta; optimizer.update(G)
    (nic_nes_master.py:123-133).  Returns (update_ratio, theta')."""
    reg = l2coeff * theta_now
    return optimizer.update(-grad_estimate + reg)
