"""fc_caption forward + greedy decode, restated on torch CPU float32.

TEST INFRASTRUCTURE (see oracle/__init__.py).  PINNED: bit-identical to the
imported reference on seeded inputs (oracle/refcheck.py, tests/golden/).

Follows, without copying, the reference's
  * LSTMCore.forward            src/captioning/nets.py:98-134  (5-chunk maxout LSTM)
  * FCModel._sample(greedy)     src/captioning/nets.py:183-245 (18-iteration loop)
  * PolicyNet.set_from_vector   src/algorithm/nets.py:163-165  (flat theta -> tensors)
torch's CPU kernels are used for the float32 linear algebra precisely because
they are what the reference itself calls, so results agree to the bit.
"""
import numpy as np
import torch
import torch.nn.functional as F

from .noise import flat_layout


class Dims:
    def __init__(self, vocab_size, input_encoding_size, rnn_size, fc_feat_size,
                 seq_length=16, layer_n=False, layer_n_affine=False, vbn=False, vbn_affine=False, vbn_e=False):
        self.vocab_size = vocab_size
        self.input_encoding_size = input_encoding_size
        self.rnn_size = rnn_size
        self.fc_feat_size = fc_feat_size
        self.seq_length = seq_length
        self.vbn = bool(vbn)                       # batch norm over the rows of the batch, no running statistics (captioning/nets.py:87-91)
        self.vbn_affine = bool(vbn_affine)
        self.vbn_e = bool(vbn_e)                   # batch norm on the outputs of img_embed and embed (captioning/nets.py:156-165)
        self.layer_n = bool(layer_n) and not self.vbn
        self.layer_n_affine = bool(layer_n_affine)

    def layout(self):
        return flat_layout(self.vocab_size, self.input_encoding_size, self.rnn_size,
                           self.fc_feat_size, self.layer_n, self.layer_n_affine, self.vbn, self.vbn_affine, self.vbn_e)

    def num_params(self):
        lay = self.layout()
        return lay[-1][2] + lay[-1][3]


def unflatten(dims, theta):
    """flat float32 theta -> {name: tensor view}; asserts the length like
    src/algorithm/nets.py:164."""
    theta = torch.as_tensor(np.asarray(theta))
    assert theta.numel() == dims.num_params()
    return {name: theta[start:start + cnt].view(*shape) for name, shape, start, cnt in dims.layout()}


def xavier_theta(dims, seed):
    """Reference-style init (src/algorithm/nets.py:62-69): Xavier-normal weights,
    zero biases, LayerNorm left at 1/0.  The normals come from this build's own
    Philox table generator (oracle/noise.py) so that fixtures can rebuild theta
    from a seed on any box without storing D floats."""
    from .noise import table_slice
    theta = np.zeros(dims.num_params(), dtype=np.float32)
    for name, shape, start, cnt in dims.layout():
        if 'ln' in name or 'bn' in name or '1' in name:              # algorithm/nets.py:62-64 skips them: torch's default 1 / 0
            if name.endswith('weight'):
                theta[start:start + cnt] = 1.0
        elif name.endswith('weight'):
            fan_out, fan_in = shape
            std = np.float32((2.0 / (fan_in + fan_out)) ** 0.5)
            theta[start:start + cnt] = table_slice(seed, start, cnt).astype(np.float32) * std
    return theta


def _ln(x, n, w, b):
    return F.layer_norm(x, (n,), w, b, 1e-5)


def _bn(x, w, b):
    """nn.BatchNorm1d(track_running_stats=False): the statistics of the batch at hand, in train and eval mode alike"""
    return F.batch_norm(x, None, None, w, b, True, 0.1, 1e-5)


def img_embed(dims, p, fc):
    """src/captioning/nets.py:150,156-165,195: Linear, followed by batch norm over the rows when vbn_e"""
    if dims.vbn_e:
        return _bn(F.linear(fc, p['img_embed.0.weight'], p['img_embed.0.bias']), p.get('img_embed.1.weight'), p.get('img_embed.1.bias'))
    return F.linear(fc, p['img_embed.weight'], p['img_embed.bias'])


def embed(dims, p, it):
    """src/captioning/nets.py:151,156-165,199: Embedding, followed by batch norm over the rows when vbn_e"""
    if dims.vbn_e:
        return _bn(F.embedding(it, p['embed.0.weight']), p.get('embed.1.weight'), p.get('embed.1.bias'))
    return F.embedding(it, p['embed.weight'])


def lstm_core(dims, p, xt, h, c):
    """src/captioning/nets.py:98-134."""
    R = dims.rnn_size
    a_i = F.linear(xt, p['core.i2h.weight'], p['core.i2h.bias'])
    a_h = F.linear(h, p['core.h2h.weight'], p['core.h2h.bias'])
    if dims.vbn:
        a_i = _bn(a_i, p.get('core.i2h_bn.weight'), p.get('core.i2h_bn.bias'))
        a_h = _bn(a_h, p.get('core.h2h_bn.weight'), p.get('core.h2h_bn.bias'))
    elif dims.layer_n:
        a_i = _ln(a_i, 5 * R, p.get('core.i2h_ln.weight'), p.get('core.i2h_ln.bias'))
        a_h = _ln(a_h, 5 * R, p.get('core.h2h_ln.weight'), p.get('core.h2h_ln.bias'))
    sums = a_i + a_h
    sig = torch.sigmoid(sums[:, :3 * R])
    in_gate, forget_gate, out_gate = sig[:, :R], sig[:, R:2 * R], sig[:, 2 * R:3 * R]
    in_transform = torch.max(sums[:, 3 * R:4 * R], sums[:, 4 * R:5 * R])
    next_c = forget_gate * c + in_gate * in_transform
    if dims.vbn:
        act_c = _bn(next_c, p.get('core.c_bn.weight'), p.get('core.c_bn.bias'))
    else:
        act_c = _ln(next_c, R, p.get('core.c_ln.weight'), p.get('core.c_ln.bias')) if dims.layer_n else next_c
    next_h = out_gate * torch.tanh(act_c)
    return next_h, next_c


def decode_greedy(dims, theta, fc_feats, trace=False):
    """FCModel._sample(fc, greedy=True) -- src/captioning/nets.py:183-245.

    Returns seq int64 [B, T], seq_logprobs float32 [B, T] (and, with trace, the
    per-step hidden states / logits the necessary schedule produces: 1 image
    step, T+1 core steps, T vocabulary projections).
    """
    with torch.no_grad():
        p = unflatten(dims, theta)
        fc = torch.as_tensor(np.asarray(fc_feats, dtype=np.float32))
        B, T, R = fc.shape[0], dims.seq_length, dims.rnn_size
        h = torch.zeros(B, R)
        c = torch.zeros(B, R)
        seq = torch.zeros(B, T, dtype=torch.long)
        lps = torch.zeros(B, T)
        tr = {'h': [], 'c': [], 'logits': [], 'x': []}
        unfinished = None
        it = None
        for t in range(T + 1):
            if t == 0:
                xt = img_embed(dims, p, fc)
            else:
                if t == 1:
                    it = torch.zeros(B, dtype=torch.long)   # <bos>
                xt = embed(dims, p, it)
            h, c = lstm_core(dims, p, xt, h, c)
            if trace:
                tr['x'].append(xt.numpy().copy()); tr['h'].append(h.numpy().copy()); tr['c'].append(c.numpy().copy())
            if t == 0:
                continue          # the reference computes and discards these logits (:202-206)
            logits = F.linear(h, p['logit.weight'], p['logit.bias'])
            logprobs = F.log_softmax(logits, dim=1)
            if trace:
                tr['logits'].append(logits.numpy().copy())
            lp, it = torch.max(logprobs, 1)
            unfinished = (it > 0) if t == 1 else unfinished * (it > 0)
            it = it * unfinished.type_as(it)
            seq[:, t - 1] = it
            lps[:, t - 1] = lp
            if unfinished.sum() == 0:
                break
        if trace:
            return seq.numpy(), lps.numpy(), tr
        return seq.numpy(), lps.numpy()


def choice_from_probs(prob_row, u):
    """what `np.random.choice(len(n_row), 1, p=n_row)` returns when the generator's next uniform is u
    (src/captioning/nets.py:217-222 + numpy's legacy RandomState.choice: float64 cdf = p.cumsum(), cdf /= cdf[-1],
    searchsorted(u, side='right')).  prob_row float32 = exp(log_softmax row).  -> (index, distance of u to the nearest
    edge of the chosen cdf interval, a measure of how close the draw was to flipping)"""
    row = np.asarray(prob_row, dtype=np.float32)
    n_row = row / np.linalg.norm(row, ord=1)
    cdf = n_row.astype(np.float64).cumsum()
    cdf /= cdf[-1]
    idx = int(cdf.searchsorted(u, side='right'))
    lo = cdf[idx - 1] if idx > 0 else 0.0
    return idx, min(u - lo, cdf[idx] - u)


def decode_sample(dims, theta, fc_feats, uniforms, forced=None):
    """FCModel._sample(fc, greedy=False) -- src/captioning/nets.py:183-245 -- with the uniforms the reference draws from the
    global numpy generator made an explicit input: uniforms float64 [B, T], [b, t-1] = the draw of row b at step t
    (the reference also draws B values at the image step t = 0 and throws the sampled tokens away, :197-198).
    forced int64 [B, T] (optional): feed these tokens forward instead of the sampled ones (per-step comparison).
    Returns seq int64 [B, T], seq_logprobs float32 [B, T], gap float64 [B, T] (see choice_from_probs; inf where no draw)."""
    with torch.no_grad():
        p = unflatten(dims, theta)
        fc = torch.as_tensor(np.asarray(fc_feats, dtype=np.float32))
        B, T, R = fc.shape[0], dims.seq_length, dims.rnn_size
        h = torch.zeros(B, R)
        c = torch.zeros(B, R)
        seq = torch.zeros(B, T, dtype=torch.long)
        lps = torch.zeros(B, T)
        gap = np.full((B, T), np.inf)
        unfinished = None
        it = None
        for t in range(T + 1):
            if t == 0:
                xt = img_embed(dims, p, fc)
            else:
                if t == 1:
                    it = torch.zeros(B, dtype=torch.long)
                xt = embed(dims, p, it)
            h, c = lstm_core(dims, p, xt, h, c)
            if t == 0:
                continue
            logprobs = F.log_softmax(F.linear(h, p['logit.weight'], p['logit.bias']), dim=1)
            prob_prev = torch.exp(logprobs).numpy()
            result = []
            for b in range(B):
                idx, g = choice_from_probs(prob_prev[b], float(uniforms[b, t - 1]))
                result.append(idx); gap[b, t - 1] = g
            it = torch.LongTensor(result)
            lp = logprobs.gather(1, it.unsqueeze(1)).view(-1)
            unfinished = (it > 0) if t == 1 else unfinished * (it > 0)
            it = it * unfinished.type_as(it)
            seq[:, t - 1] = it
            lps[:, t - 1] = lp
            if unfinished.sum() == 0:
                break
            if forced is not None:
                it = torch.as_tensor(np.asarray(forced)[:, t - 1], dtype=torch.long)
        return seq.numpy(), lps.numpy(), gap
