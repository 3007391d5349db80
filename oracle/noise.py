"""Noise table, padded row layout and offset sampler -- the build's own spec.

TEST INFRASTRUCTURE (see oracle/__init__.py).  PARITY UNPINNED against the
reference: the reference has no noise table (src/algorithm/nets.py:92-102 draws
`torch.empty_like(theta).normal_(0, sigma)` from an unseeded RNG and ships the
whole vector through Redis, src/algorithm/nic_nes/nic_nes_worker.py:142,156).
What IS pinned here is the generator: Philox4x32-10 (Salmon et al., SC'11,
Random123) against its published known-answer vectors (tests/test_oracle_noise.py).

Specification
-------------
table[4*b + j], b = 0 .. L/4-1, j = 0..3:
    (r0, r1, r2, r3) = philox4x32_10(counter=(lo32(b), hi32(b), 0, 0),
                                     key=(lo32(seed), hi32(seed)))
    u(r)   = (r + 0.5) * 2**-32                      (float64, in (0, 1))
    rad_a  = sqrt(-2 ln u(r0));  z0 = rad_a cos(2 pi u(r1)); z1 = rad_a sin(2 pi u(r1))
    rad_b  = sqrt(-2 ln u(r2));  z2 = rad_b cos(2 pi u(r3)); z3 = rad_b sin(2 pi u(r3))
    table[4b + j] = float16_rne(float32_rne(z_j))
The table is a float16 array: every eps value is exactly representable in the
tensor-core operand type, so the eps operand of the perturbation GEMMs carries
no quantisation error at all.

offset of noise row i in iteration `it`:
    r = philox4x32_10(counter=(i, lo32(it), hi32(it), 0x4F464653), key(seed))[0]
    off_i = G * ((r * n_slots) >> 32),   n_slots = (L - D_pad) / G + 1
G = lcm(fc_feat_size, input_encoding_size, rnn_size) elements, so that every
weight matrix's eps slice is a whole number of rows of the 2-D [L/K, K] view of
the table that the TMA descriptors use.

padded row layout: tensor k of the flat theta (registration order, see
`flat_layout`) starts at pstart_k = round_up(pstart_{k-1} + count_{k-1}, G);
D_pad = round_up(end, G).  eps for flat parameter p of tensor k in row i is
table[off_i + pstart_k + (p - start_k)].

member parameters (mirrors src/algorithm/nets.py:113 and
src/algorithm/nic_nes/nic_nes_worker.py:151):
    noise  = float32(sigma) * float32(eps)            (float32 multiply)
    theta+ = theta + noise ; theta- = theta - noise   (float32)
"""
import math
import numpy as np

M0 = np.uint64(0xD2511F53)
M1 = np.uint64(0xCD9E8D57)
W0 = 0x9E3779B9
W1 = 0xBB67AE85
MASK = np.uint64(0xFFFFFFFF)
OFFSET_DOMAIN = 0x4F464653  # "OFFS"
PARENT_DOMAIN = 0x50415245  # "PARE"
TOURN_DOMAIN = 0x544F5552   # "TOUR"


def philox4x32_10(c0, c1, c2, c3, k0, k1):
    """Vectorised Philox4x32-10.  Inputs: uint32-valued arrays (any int dtype)
    of a common shape (scalars broadcast); returns four uint32 arrays."""
    c0 = np.asarray(c0, dtype=np.uint64) & MASK
    c1 = np.asarray(c1, dtype=np.uint64) & MASK
    c2 = np.asarray(c2, dtype=np.uint64) & MASK
    c3 = np.asarray(c3, dtype=np.uint64) & MASK
    c0, c1, c2, c3 = np.broadcast_arrays(c0, c1, c2, c3)
    k0 = int(k0) & 0xFFFFFFFF
    k1 = int(k1) & 0xFFFFFFFF
    for _ in range(10):
        p0 = M0 * c0
        p1 = M1 * c2
        hi0, lo0 = p0 >> np.uint64(32), p0 & MASK
        hi1, lo1 = p1 >> np.uint64(32), p1 & MASK
        c0, c1, c2, c3 = (hi1 ^ c1 ^ np.uint64(k0), lo1, hi0 ^ c3 ^ np.uint64(k1), lo0)
        k0 = (k0 + W0) & 0xFFFFFFFF
        k1 = (k1 + W1) & 0xFFFFFFFF
    return (c0.astype(np.uint32), c1.astype(np.uint32), c2.astype(np.uint32), c3.astype(np.uint32))


def _key(seed):
    seed = int(seed) & 0xFFFFFFFFFFFFFFFF
    return seed & 0xFFFFFFFF, seed >> 32


def table_slice(seed, start, count):
    """float16 table[start : start+count] (start, count in elements)."""
    b0 = start // 4
    b1 = (start + count + 3) // 4
    out = np.empty((b1 - b0) * 4, dtype=np.float16)
    k0, k1 = _key(seed)
    CH = 1 << 20
    for lo in range(b0, b1, CH):
        hi = min(lo + CH, b1)
        b = np.arange(lo, hi, dtype=np.uint64)
        r0, r1, r2, r3 = philox4x32_10(b & MASK, b >> np.uint64(32), 0, 0, k0, k1)
        z = np.empty((hi - lo, 4), dtype=np.float64)
        for (ra, rb, col) in ((r0, r1, 0), (r2, r3, 2)):
            u1 = (ra.astype(np.float64) + 0.5) * 2.0 ** -32
            u2 = (rb.astype(np.float64) + 0.5) * 2.0 ** -32
            rad = np.sqrt(-2.0 * np.log(u1))
            ang = (2.0 * math.pi) * u2
            z[:, col] = rad * np.cos(ang)
            z[:, col + 1] = rad * np.sin(ang)
        out[(lo - b0) * 4:(hi - b0) * 4] = z.reshape(-1).astype(np.float32).astype(np.float16)
    s = start - b0 * 4
    return out[s:s + count]


def lcm(*xs):
    g = 1
    for x in xs:
        g = g * x // math.gcd(g, x)
    return g


def flat_layout(vocab_size, input_encoding_size, rnn_size, fc_feat_size,
                layer_n=False, layer_n_affine=False, vbn=False, vbn_affine=False, vbn_e=False):
    """[(name, shape, start, count)] in `named_parameters()` order of the
    reference FCModel (src/captioning/nets.py:150-153, LSTMCore :81-96).  vbn_e (:156-165) wraps img_embed and embed into
    Sequential(layer, BatchNorm1d): their parameters are renamed `<layer>.0.*` and, with vbn_affine, followed by `<layer>.1.weight / bias`."""
    V1, E, R, Fd = vocab_size + 1, input_encoding_size, rnn_size, fc_feat_size
    if vbn_e:
        tensors = [('img_embed.0.weight', (E, Fd)), ('img_embed.0.bias', (E,))]
        tensors += [('img_embed.1.weight', (E,)), ('img_embed.1.bias', (E,))] if vbn_affine else []
        tensors += [('embed.0.weight', (V1, E))]
        tensors += [('embed.1.weight', (E,)), ('embed.1.bias', (E,))] if vbn_affine else []
    else:
        tensors = [('img_embed.weight', (E, Fd)), ('img_embed.bias', (E,)), ('embed.weight', (V1, E))]
    tensors += [
        ('logit.weight', (V1, R)), ('logit.bias', (V1,)),
        ('core.i2h.weight', (5 * R, E)), ('core.i2h.bias', (5 * R,)),
        ('core.h2h.weight', (5 * R, R)), ('core.h2h.bias', (5 * R,)),
    ]
    if vbn:                                      # LSTMCore: `if self.vbn ... elif self.layer_n` (captioning/nets.py:87-96)
        layer_n = False
        if vbn_affine:
            tensors += [
                ('core.i2h_bn.weight', (5 * R,)), ('core.i2h_bn.bias', (5 * R,)),
                ('core.h2h_bn.weight', (5 * R,)), ('core.h2h_bn.bias', (5 * R,)),
                ('core.c_bn.weight', (R,)), ('core.c_bn.bias', (R,)),
            ]
    if layer_n and layer_n_affine:
        tensors += [
            ('core.i2h_ln.weight', (5 * R,)), ('core.i2h_ln.bias', (5 * R,)),
            ('core.h2h_ln.weight', (5 * R,)), ('core.h2h_ln.bias', (5 * R,)),
            ('core.c_ln.weight', (R,)), ('core.c_ln.bias', (R,)),
        ]
    out, pos = [], 0
    for name, shape in tensors:
        cnt = int(np.prod(shape))
        out.append((name, shape, pos, cnt))
        pos += cnt
    return out


def padded_layout(flat, G):
    """[(name, shape, start, count, pstart)], D, D_pad."""
    out, p = [], 0
    for name, shape, start, cnt in flat:
        out.append((name, shape, start, cnt, p))
        p = -(-(p + cnt) // G) * G
    D = flat[-1][2] + flat[-1][3]
    return out, D, p


def granularity(input_encoding_size, rnn_size, fc_feat_size):
    return lcm(fc_feat_size, input_encoding_size, rnn_size)


def sample_offsets(seed, iteration, F, table_len, D_pad, G):
    k0, k1 = _key(seed)
    n_slots = (table_len - D_pad) // G + 1
    assert n_slots >= 1 and n_slots < 2 ** 32
    i = np.arange(F, dtype=np.uint64)
    it = int(iteration) & 0xFFFFFFFFFFFFFFFF
    r0, _, _, _ = philox4x32_10(i, it & 0xFFFFFFFF, it >> 32, OFFSET_DOMAIN, k0, k1)
    slot = (r0.astype(np.uint64) * np.uint64(n_slots)) >> np.uint64(32)
    return (slot * np.uint64(G)).astype(np.int64)


def noise_row(seed, offset, playout, D):
    """eps row in FLAT order as float16[D] for one offset."""
    eps = np.empty(D, dtype=np.float16)
    for name, shape, start, cnt, pstart in playout:
        eps[start:start + cnt] = table_slice(seed, int(offset) + pstart, cnt)
    return eps


def proportional_scale(theta):
    """SM-PROPORTIONAL (src/algorithm/nets.py:107-111): params = theta.clone(); params[params == 0] = mean(|theta|);
    noise *= |params|.  The mean is taken in float64 and rounded to float32 (this build's rule; torch's float32 mean
    agrees to ~1 ulp and only matters for parameters that are exactly zero)."""
    theta = np.asarray(theta, dtype=np.float32)
    mean = np.float32(np.abs(theta).astype(np.float64).sum() / theta.size)
    return np.where(theta == 0, np.abs(mean), np.abs(theta)).astype(np.float32)


def sensitivity_vector(raw, underflow):
    """Sensitivity.set_sensitivity (src/algorithm/safe_mutations.py:27-32): clamp below at the underflow value, divide by the minimum"""
    s = np.asarray(raw, dtype=np.float32).copy()
    s[s < np.float32(underflow)] = np.float32(underflow)
    s /= s.min()
    return s


def member_theta(theta, eps, sigma, sign, scale=None, sens=None):
    """theta +/- float32(sigma)*eps in float32 (src/algorithm/nets.py:113,
    src/algorithm/nic_nes/nic_nes_worker.py:151); with a safe-mutation scale the noise is multiplied by it first (:107-111),
    with a sensitivity vector divided by it (:104-105)."""
    noise = np.float32(sigma) * eps.astype(np.float32)
    if scale is not None:
        noise = noise * scale.astype(np.float32)
    if sens is not None:
        noise = noise / sens.astype(np.float32)
    return (theta + noise) if sign > 0 else (theta - noise), noise


def sample_parents(seed, iteration, F, n_parents):
    """NIC-ES parent pick for offspring i (uniform, nic_es_worker.py:161 uses an unseeded random.choice;
    this build's rule): parent_i = (philox(i, it, PARENT_DOMAIN)[0] * n_parents) >> 32."""
    k0, k1 = _key(seed)
    i = np.arange(F, dtype=np.uint64)
    it = int(iteration) & 0xFFFFFFFFFFFFFFFF
    r0, _, _, _ = philox4x32_10(i, it & 0xFFFFFFFF, it >> 32, PARENT_DOMAIN, k0, k1)
    return ((r0.astype(np.uint64) * np.uint64(n_parents)) >> np.uint64(32)).astype(np.int32)


def sample_parents_tournament(seed, iteration, F, n_parents, tournament_size):
    """NIC-ES parent pick with selection == 'tournament' (nic_es_worker.py:150-158:
    `rs.choice(len(parents), min(len(parents), tournament_size), replace=False).min()` from an unseeded RandomState;
    parents are sorted best first).  This build's rule: a uniformly random t-subset by Floyd's algorithm,
        for d in 0..t-1: j = n - t + d;  r = (u_d * (j + 1)) >> 32;  pick = j if r is already in the set else r
    with u_d = word d % 4 of philox(counter = (i, lo32(it), hi32(it) + ((d // 4) << 16), TOURN_DOMAIN), key(seed));
    the parent is the lowest pick."""
    k0, k1 = _key(seed)
    n, t = int(n_parents), min(int(n_parents), int(tournament_size))
    assert 1 <= t and n <= 4096
    it = int(iteration) & 0xFFFFFFFFFFFFFFFF
    assert it >> 48 == 0
    i = np.arange(F, dtype=np.uint64)
    chosen = np.zeros((F, n), dtype=bool)
    best = np.full(F, n, dtype=np.int64)
    words = None
    for d in range(t):
        if d % 4 == 0:
            words = philox4x32_10(i, it & 0xFFFFFFFF, (it >> 32) + ((d // 4) << 16), TOURN_DOMAIN, k0, k1)
        j = n - t + d
        r = ((words[d % 4].astype(np.uint64) * np.uint64(j + 1)) >> np.uint64(32)).astype(np.int64)
        pick = np.where(chosen[np.arange(F), r], j, r)
        chosen[np.arange(F), pick] = True
        best = np.minimum(best, pick)
    return best.astype(np.int32)
