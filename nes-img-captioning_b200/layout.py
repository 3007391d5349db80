"""Flat-theta layout of the fc_caption model and the padded noise-row layout (product-side statement of
the spec in DESIGN.md; the C library computes the same thing internally and tests compare the two).

Parameter order = FCModel.named_parameters() of the reference (src/captioning/nets.py:150-153, LSTMCore
:81-96): img_embed, embed, logit, core.i2h, core.h2h [, core.i2h_ln, core.h2h_ln, core.c_ln]."""
import math
from collections import namedtuple

import numpy as np

TensorSpec = namedtuple('TensorSpec', ['name', 'shape', 'start', 'count', 'pstart'])


def fc_tensors(vocab_size, input_encoding_size, rnn_size, fc_feat_size, layer_n=False, layer_n_affine=False, vbn_e=False, vbn_e_affine=False):
    """layer_n: 0 none, 1 LayerNorm (core.*_ln), 2 batch norm over the rows = the reference's vbn (core.*_bn).  vbn_e: img_embed and
    embed are Sequential(layer, BatchNorm1d) in the reference (captioning/nets.py:156-165), which renames their parameters to
    `<layer>.0.*` and, with vbn_affine, puts `<layer>.1.weight / bias` right behind them"""
    V1, E, R, Fd = vocab_size + 1, input_encoding_size, rnn_size, fc_feat_size
    if vbn_e:
        t = [('img_embed.0.weight', (E, Fd)), ('img_embed.0.bias', (E,))]
        t += [('img_embed.1.weight', (E,)), ('img_embed.1.bias', (E,))] if vbn_e_affine else []
        t += [('embed.0.weight', (V1, E))]
        t += [('embed.1.weight', (E,)), ('embed.1.bias', (E,))] if vbn_e_affine else []
    else:
        t = [('img_embed.weight', (E, Fd)), ('img_embed.bias', (E,)), ('embed.weight', (V1, E))]
    t += [('logit.weight', (V1, R)), ('logit.bias', (V1,)),
          ('core.i2h.weight', (5 * R, E)), ('core.i2h.bias', (5 * R,)),
          ('core.h2h.weight', (5 * R, R)), ('core.h2h.bias', (5 * R,))]
    if layer_n and layer_n_affine:
        k = 'bn' if int(layer_n) == 2 else 'ln'
        t += [('core.i2h_%s.weight' % k, (5 * R,)), ('core.i2h_%s.bias' % k, (5 * R,)),
              ('core.h2h_%s.weight' % k, (5 * R,)), ('core.h2h_%s.bias' % k, (5 * R,)),
              ('core.c_%s.weight' % k, (R,)), ('core.c_%s.bias' % k, (R,))]
    return t


def mnist_tensors():
    """MnistNet.named_parameters() (src/classification/nets.py:13-15): D = 8490"""
    return [('conv1.weight', (10, 1, 5, 5)), ('conv1.bias', (10,)), ('conv2.weight', (20, 10, 5, 5)), ('conv2.bias', (20,)),
            ('fc1.weight', (10, 320)), ('fc1.bias', (10,))]


def granule(input_encoding_size, rnn_size, fc_feat_size):
    g = 1
    for x in (fc_feat_size, input_encoding_size, rnn_size):
        g = g * x // math.gcd(g, x)
    return g


def build_layout(tensors, G):
    """tensors: [(name, shape)] -> ([TensorSpec], D, D_pad)"""
    out, pos, p = [], 0, 0
    for name, shape in tensors:
        cnt = int(np.prod(shape))
        out.append(TensorSpec(name, tuple(shape), pos, cnt, p))
        pos += cnt
        p = -(-(p + cnt) // G) * G
    return out, pos, p


def fresh_model_vectors(tensors):
    """(blank theta, per-parameter std) such that blank + std * N(0, 1) is a freshly initialised model as
    PolicyNet.initialize_params leaves it (src/algorithm/nets.py:62-69): Xavier-normal on every '*weight*', zero on every '*bias*',
    parameters whose name contains 'bn', 'ln' or '1' skipped -- for the fc_caption tensors that is the LayerNorm / batch-norm
    affine pair, which stays at torch's default weight 1 / bias 0.  A tensor that keeps torch's default Kaiming-uniform init
    (MnistNet's conv1 / fc1) is not a Gaussian and raises."""
    D = tensors[-1].start + tensors[-1].count
    blank, std = np.zeros(D, dtype=np.float32), np.zeros(D, dtype=np.float32)
    for t in tensors:
        sl = slice(t.start, t.start + t.count)
        if 'bn' in t.name or 'ln' in t.name or t.name.startswith(('img_embed.1.', 'embed.1.')):
            if t.name.endswith('weight'):
                blank[sl] = 1.0
        elif '1' in t.name:
            raise NotImplementedError('%s keeps torch default init (not Gaussian): supply fresh models yourself' % t.name)
        elif 'weight' in t.name:
            fan_out, fan_in = t.shape[0], int(np.prod(t.shape[1:]))          # nn.init.xavier_normal_: std = sqrt(2 / (fan_in + fan_out))
            std[sl] = np.float32((2.0 / (fan_in + fan_out)) ** 0.5)
    return blank, std
