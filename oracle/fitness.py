"""Log-prob-shaped fitness criteria -- restatement of src/captioning/fitness.py:43-132 as used by CaptPolicy.rollout
(src/captioning/policies.py:119-123: crit(sample_logprobs, gen_result, torch.from_numpy(ciders).float())).

TEST INFRASTRUCTURE (see oracle/__init__.py).  PINNED: the same torch float32 CPU ops as the reference, checked
bit-for-bit against the imported reference classes by oracle/refcheck.py.

kinds (Fitness enum, captioning/policies.py:21-62): 'greedy_logprob' -> AltLogFitnessCriterion, 'greedy_expprob' ->
ExpFitnessCriterion, 'greedy_linprob' -> LinFitnessCriterion, 'greedy_avgprob' -> AvgLogFitnessCriterion,
'sc_loss' -> LogFitnessCriterion.
"""
import math

import numpy as np
import torch

KINDS = ('greedy_logprob', 'greedy_expprob', 'greedy_linprob', 'greedy_avgprob', 'sc_loss')


def mask_of(seq):
    """a position counts if the PREVIOUS token was not the terminator; the first always counts (fitness.py:33-34)"""
    m = (seq > 0).float()
    return torch.cat([m.new(m.size(0), 1).fill_(1), m[:, :-1]], 1).contiguous().view(-1)


def shaped_fitness(kind, logprobs, seq, scores):
    """logprobs float32 [rows, T], seq int64 [rows, T], scores float64 [rows] (per-caption CIDEr-D, as compute_ciders returns
    them before np.repeat) -> python float"""
    logprobs = torch.as_tensor(logprobs, dtype=torch.float32)
    seq = torch.as_tensor(seq, dtype=torch.int64)
    rewards = np.repeat(np.asarray(scores, dtype=np.float64)[:, np.newaxis], seq.shape[1], 1)       # policies.py:191
    reward = torch.from_numpy(rewards.copy()).float().contiguous().view(-1)
    inp = logprobs.contiguous().view(-1)
    mask = mask_of(seq)
    if kind == 'sc_loss':
        out = - inp * reward * mask
    elif kind == 'greedy_logprob':
        pfact = torch.log10(torch.exp(inp) + 1 / (10 - 1)) + np.log10(10 - 1)
        out = pfact * reward * mask
    elif kind == 'greedy_avgprob':
        pfact = torch.log10(torch.exp(inp) + 1 / (10 - 1)) + np.log10(10 - 1)
        out = 0.5 * reward * mask + 0.5 * pfact * reward * mask
    elif kind == 'greedy_expprob':
        out = (torch.exp(torch.exp(inp)) - 1) / (math.e - 1) * reward * mask
    elif kind == 'greedy_linprob':
        out = torch.exp(inp) * reward * mask
    else:
        raise ValueError(kind)
    out = torch.sum(out) / torch.sum(mask)
    return float(out.item())
