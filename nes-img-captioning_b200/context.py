"""NesContext: one per GPU.  Owns (through the C library) theta, the shared noise table and the decode
workspace; every method forwards to one extern "C" entry point of include/nes_b200.h."""
import ctypes

import numpy as np
import torch

from . import _lib
from .layout import build_layout, fc_tensors, granule

PREC_FAST, PREC_EXACT, PREC_HYBRID = 0, 1, 2
OPT_SGD, OPT_ADAM = 0, 1


class NesContext:
    def __init__(self, model_options=None, *, counts=None, tensors=None, flat_granule=64, seed=1234, table_len=None,
                 n_parents=1, wave_members=256, rows=128, device=0, fill=True):
        """model_options: dict / namedtuple with the fc_caption ModelOptions fields (vocab_size,
        input_encoding_size, rnn_size, fc_feat_size, layer_n, layer_n_affine[, seq_length]); or
        counts=[...] for a generic flat parameter vector (master math only), or tensors=[(name, shape), ...] for a flat vector
        with named tensors (the MNIST net, layout.mnist_tensors()).
        table_len: fp16 elements of the shared noise table; None = 2^28, or 256 padded noise rows rounded up to a power of two if
        that is more (2^32 = 8 GB for the E = R = 512 model: 2.1 M distinct rows, so a population of 5000 draws ~6 colliding pairs
        per iteration instead of ~100 and K7's rows do not overlap inside L2)."""
        self.lib = _lib.load()
        self.device = int(device)
        self._h = ctypes.c_void_p()
        self.seed, self.n_parents = int(seed), int(n_parents)
        self._evolve_counter = 0           # PolicyNet.evolve() without an explicit offset: one stream per context, not per net
        self.evolve_stream = 0             # ... offset by this (a worker / rank id) so that two processes on one seed differ
        if model_options is not None:
            mo = model_options._asdict() if hasattr(model_options, '_asdict') else dict(model_options)
            self.dims = dict(vocab_size=int(mo['vocab_size']), input_encoding_size=int(mo['input_encoding_size']),
                             rnn_size=int(mo['rnn_size']), fc_feat_size=int(mo['fc_feat_size']),
                             seq_length=int(mo.get('seq_length') or 16), layer_n=int(bool(mo.get('layer_n'))),
                             layer_n_affine=int(bool(mo.get('layer_n_affine'))))
            if mo.get('vbn'):                      # LSTMCore: `if self.vbn ... elif self.layer_n` (captioning/nets.py:87-96)
                self.dims.update(layer_n=2, layer_n_affine=int(bool(mo.get('vbn_affine'))))
            # batch norm on the embeddings as well (captioning/nets.py:156-165); its affine flag is vbn_affine whatever the cell uses
            self.dims.update(vbn_e=int(bool(mo.get('vbn_e'))), vbn_e_affine=int(bool(mo.get('vbn_e')) and bool(mo.get('vbn_affine'))))
            d = _lib.FcDims(**self.dims)
            self.G = granule(self.dims['input_encoding_size'], self.dims['rnn_size'], self.dims['fc_feat_size'])
            self.tensors, self.D, self.D_pad = build_layout(
                fc_tensors(self.dims['vocab_size'], self.dims['input_encoding_size'], self.dims['rnn_size'],
                           self.dims['fc_feat_size'], self.dims['layer_n'], self.dims['layer_n_affine'],
                           self.dims['vbn_e'], self.dims['vbn_e_affine']), self.G)
            self.table_len = self._default_table_len(table_len)
            _lib.check(self.lib.nes_ctx_create_fc(ctypes.byref(self._h), self.device, ctypes.byref(d), self.seed,
                                                  self.table_len, self.n_parents, int(wave_members), int(rows)))
            self.wave_members, self.rows_cap = int(wave_members), int(rows)
        else:
            assert counts is not None or tensors is not None, 'give model_options, counts or tensors'
            self.dims = None
            self.G = int(flat_granule)
            named = [(n, tuple(sh)) for n, sh in tensors] if tensors is not None else [('t%d' % i, (int(c),)) for i, c in enumerate(counts)]
            counts = [int(np.prod(sh)) for _, sh in named]
            self.tensors, self.D, self.D_pad = build_layout(named, self.G)
            self.table_len = self._default_table_len(table_len)
            arr = (ctypes.c_int64 * len(counts))(*[int(c) for c in counts])
            _lib.check(self.lib.nes_ctx_create_flat(ctypes.byref(self._h), self.device, arr, len(counts), self.G,
                                                    self.seed, self.table_len, self.n_parents))
        assert self.lib.nes_num_params(self._h) == self.D and self.lib.nes_padded_params(self._h) == self.D_pad \
            and self.lib.nes_granule(self._h) == self.G, 'host / library layout mismatch'
        self.dev = torch.device('cuda', self.device)
        self.use_stream(torch.cuda.current_stream(self.dev))
        if fill:
            self.fill_table()

    def _default_table_len(self, table_len):
        if table_len is not None:
            return int(table_len)
        return max(1 << 28, 1 << int(256 * self.D_pad - 1).bit_length())

    def next_evolve_iteration(self):
        """stream index of the next PolicyNet.evolve() that was given no offset: monotonically increasing per context (a fresh
        PolicyNet on the same context does not replay earlier noise), disjoint between evolve_stream ids"""
        it = (1 << 40) + (int(self.evolve_stream) << 24) + self._evolve_counter
        self._evolve_counter += 1
        return it

    # ---- plumbing
    def use_stream(self, stream):
        self._stream = stream
        _lib.check(self.lib.nes_set_stream(self._h, ctypes.c_void_p(stream.cuda_stream)))

    def sync(self):
        _lib.check(self.lib.nes_sync(self._h))

    def close(self):
        if self._h:
            self.lib.nes_ctx_destroy(self._h)
            self._h = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def gpu_launches(self):
        return int(self.lib.nes_gpu_launches(self._h))

    @property
    def n_slots(self):
        return (self.table_len - self.D_pad) // self.G + 1

    # ---- noise table
    def fill_table(self):
        _lib.check(self.lib.nes_fill_table(self._h))

    def table_slice(self, start, count):
        out = np.empty(int(count), dtype=np.float16)
        _lib.check(self.lib.nes_table_slice(self._h, int(start), int(count), _lib.ptr(out)))
        return out

    def sample_offsets(self, iteration, F, device=False):
        out = torch.empty(int(F), dtype=torch.int64, device=self.dev if device else 'cpu')
        _lib.check(self.lib.nes_sample_offsets(self._h, int(iteration), int(F), _lib.ptr(out)))
        return out

    def sample_parents(self, iteration, F, n_parents, device=False):
        out = torch.empty(int(F), dtype=torch.int32, device=self.dev if device else 'cpu')
        _lib.check(self.lib.nes_sample_parents(self._h, int(iteration), int(F), int(n_parents), _lib.ptr(out)))
        return out

    def sample_parents_tournament(self, iteration, F, n_parents, tournament_size, device=False):
        """parent index of every offspring under selection == 'tournament' (nic_es_worker.py:150-158): the lowest of
        min(n_parents, tournament_size) distinct indices"""
        out = torch.empty(int(F), dtype=torch.int32, device=self.dev if device else 'cpu')
        _lib.check(self.lib.nes_sample_parents_tournament(self._h, int(iteration), int(F), int(n_parents), int(tournament_size), _lib.ptr(out)))
        return out

    # ---- safe mutations (src/algorithm/nets.py:104-111)
    MUTATIONS = {'': 0, 'SM-PROPORTIONAL': 1, 'SM-VECTOR': 2, 'SM-G-SUM': 2, 'SM-G-ABS': 2, 'SCALE': 3}

    def set_mutation(self, kind):
        """'' (default), 'SM-PROPORTIONAL' (noise scaled by |theta_p|, zeros by mean|theta_p|), or one of the sensitivity mutations
        'SM-VECTOR' / 'SM-G-SUM' / 'SM-G-ABS' (noise divided by a sensitivity vector given through set_sensitivity; the
        gradient-based ones come from autograd, sensitivity.py / PolicyNet.calc_sensitivity)"""
        if kind not in self.MUTATIONS:
            raise NotImplementedError('mutation %r is not on the B200 path (SURVEY 8f-3)' % (kind,))
        _lib.check(self.lib.nes_set_mutation(self._h, self.MUTATIONS[kind]))
        self.mutation = kind

    def set_sensitivity(self, sens, parent=None, underflow=None):
        """sens float32 [D] (numpy / torch, host or device): noise /= sens for members of `parent` (None: every parent slot).
        underflow given: first clamp and normalise like Sensitivity.set_sensitivity (safe_mutations.py:27-32):
        sens[sens < underflow] = underflow; sens /= sens.min()"""
        sens = torch.as_tensor(sens).to(torch.float32).clone().contiguous()
        assert sens.numel() == self.D
        if underflow is not None:
            sens[sens < underflow] = underflow
            sens /= sens.min()
        _lib.check(self.lib.nes_set_sensitivity(self._h, -1 if parent is None else int(parent), _lib.ptr(sens)))

    # ---- theta
    def set_theta(self, vector, parent=0):
        """vector: numpy / torch, float32 or float64, host or device, length D (asserted like
        src/algorithm/nets.py:164)."""
        if isinstance(vector, np.ndarray):
            vector = torch.from_numpy(np.ascontiguousarray(vector))
        assert vector.numel() == self.D, 'len(vector) == nb_learnable_params'
        assert vector.dtype in (torch.float32, torch.float64)
        vector = vector.contiguous()
        _lib.check(self.lib.nes_set_theta(self._h, int(parent), _lib.ptr(vector), int(vector.dtype == torch.float64)))

    def get_theta(self, parent=0, dtype=torch.float32, device='cpu'):
        out = torch.empty(self.D, dtype=dtype, device=device)
        _lib.check(self.lib.nes_get_theta(self._h, int(parent), _lib.ptr(out), int(dtype == torch.float64)))
        return out

    def evolve(self, offset, sigma, sign=1.0, parent=0, want_member=True, want_noise=True, dst_parent=-1, device='cpu'):
        member = torch.empty(self.D, dtype=torch.float32, device=device) if want_member else None
        noise = torch.empty(self.D, dtype=torch.float32, device=device) if want_noise else None
        _lib.check(self.lib.nes_evolve(self._h, int(parent), int(offset), float(sign), float(sigma),
                                       _lib.ptr(member), _lib.ptr(noise), int(dst_parent)))
        return member, noise

    def evolve_batch(self, src_parents, offsets, sigmas, dst_parents):
        """dst_parents[i] <- theta(src_parents[i]) + sigmas[i] * eps(offsets[i]) for all i in one launch (host arrays)"""
        src = np.ascontiguousarray(src_parents, dtype=np.int32); dst = np.ascontiguousarray(dst_parents, dtype=np.int32)
        off = np.ascontiguousarray(offsets, dtype=np.int64); sg = np.ascontiguousarray(sigmas, dtype=np.float32)
        assert len(src) == len(dst) == len(off) == len(sg)
        _lib.check(self.lib.nes_evolve_batch(self._h, len(src), _lib.ptr(src), _lib.ptr(off), _lib.ptr(sg), _lib.ptr(dst)))

    def es_select_evolve(self, fitness, offsets, parents, n_keep, sigma, dst_base, want_order=True):
        """ESMaster.selection + the survivors written as parent slots dst_base .. dst_base + n_keep - 1, on the device
        (nes_es_select_evolve).  fitness float64 [n], offsets int64 [n], parents int32 [n] absolute slots (torch, any device).
        Returns the device tensor order int64 [n_keep] (offspring index at each rank) or None."""
        fitness = torch.as_tensor(fitness, dtype=torch.float64).contiguous()
        offsets = torch.as_tensor(offsets, dtype=torch.int64).contiguous()
        parents = None if parents is None else torch.as_tensor(parents, dtype=torch.int32).contiguous()
        n = fitness.numel()
        assert offsets.numel() == n and (parents is None or parents.numel() == n)
        order = torch.empty(int(n_keep), dtype=torch.int64, device=self.dev) if want_order else None
        _lib.check(self.lib.nes_es_select_evolve(self._h, _lib.ptr(fitness), _lib.ptr(offsets), _lib.ptr(parents), n, int(n_keep),
                                                 float(sigma), int(dst_base), _lib.ptr(order)))
        return order

    def copy_parent(self, src, dst):
        _lib.check(self.lib.nes_copy_parent(self._h, int(src), int(dst)))

    # ---- MNIST classifier (BASELINE configs[0]; classification/nets.py:25-37, classification/policies.py:16-86)
    def mnist_rollout(self, offsets, signs, images, labels, sigma, parents=None, shared_batch=None, out_device='cpu', want_logits=False):
        """offsets int64 [P], signs float32 [P]; images float32 [rows, 1, 28, 28] (shared) or [P, rows, 1, 28, 28]; labels int64 to match.
        Returns fitness float64 [P] (= -CrossEntropyLoss), correct int32 [P] (+ logits float32 [P, rows, 10] when want_logits)."""
        offsets = torch.as_tensor(offsets, dtype=torch.int64).contiguous()
        signs = torch.as_tensor(signs, dtype=torch.float32).contiguous()
        images = torch.as_tensor(images).to(torch.float32).contiguous()
        labels = torch.as_tensor(labels).to(torch.int64).contiguous()
        P = offsets.numel()
        assert signs.numel() == P
        if images.dim() in (2, 3) and images.shape[-1] == 28:          # [rows, 28, 28]
            images = images.unsqueeze(-3)
        if shared_batch is None:
            shared_batch = images.dim() == 4
        rows = images.shape[-4]
        assert tuple(images.shape[-3:]) == (1, 28, 28) and (shared_batch or images.shape[0] == P)
        assert labels.numel() == (1 if shared_batch else P) * rows
        assert images.device == labels.device
        if parents is not None:
            parents = torch.as_tensor(parents, dtype=torch.int32).contiguous()
            assert parents.numel() == P
        fit = torch.empty(P, dtype=torch.float64, device=out_device)
        cor = torch.empty(P, dtype=torch.int32, device=out_device)
        lg = torch.empty((P, rows, 10), dtype=torch.float32, device=out_device) if want_logits else None
        _lib.check(self.lib.nes_mnist_rollout(self._h, P, _lib.ptr(offsets), _lib.ptr(signs), _lib.ptr(parents), _lib.ptr(images), _lib.ptr(labels),
                                              int(rows), int(bool(shared_batch)), float(sigma), _lib.ptr(fit), _lib.ptr(cor), _lib.ptr(lg)))
        return (fit, cor, lg) if want_logits else (fit, cor)

    # ---- population rollout
    def rollout(self, offsets, signs, fc, sigma, parents=None, precision=PREC_EXACT, shared_batch=None,
                out_device='cpu', forced_tokens=None, trace=False, uniforms=None):
        """offsets int64 [P], signs float32 [P], fc float32 [rows, Fd] (shared) or [P, rows, Fd].
        uniforms float64 [P, rows, T] in [0, 1): sampled decode (FCModel._sample(greedy=False), captioning/nets.py:210-231),
        token = numpy's legacy choice evaluated at that draw; None = greedy.
        Returns seq int64 [P, rows, T], logprobs float32 [P, rows, T] (+ h_trace, logit_trace when trace)."""
        offsets = torch.as_tensor(offsets, dtype=torch.int64).contiguous()
        signs = torch.as_tensor(signs, dtype=torch.float32).contiguous()
        if isinstance(fc, np.ndarray):
            fc = torch.from_numpy(np.ascontiguousarray(fc, dtype=np.float32))
        fc = fc.to(torch.float32).contiguous()
        P = offsets.numel()
        assert signs.numel() == P
        if shared_batch is None:
            shared_batch = fc.dim() == 2
        rows = fc.shape[-2]
        assert fc.shape[-1] == self.dims['fc_feat_size']
        assert shared_batch or fc.shape[0] == P
        if parents is not None:
            parents = torch.as_tensor(parents, dtype=torch.int32).contiguous()
            assert parents.numel() == P
        T, R, V1 = self.dims['seq_length'], self.dims['rnn_size'], self.dims['vocab_size'] + 1
        seq = torch.empty((P, rows, T), dtype=torch.int64, device=out_device)
        lp = torch.empty((P, rows, T), dtype=torch.float32, device=out_device)
        if uniforms is not None:
            assert not trace and forced_tokens is None
            if isinstance(uniforms, np.ndarray):
                uniforms = torch.from_numpy(np.ascontiguousarray(uniforms, dtype=np.float64))
            uniforms = uniforms.to(torch.float64).contiguous()
            assert tuple(uniforms.shape) == (P, rows, T)
            _lib.check(self.lib.nes_rollout_sample(self._h, P, _lib.ptr(offsets), _lib.ptr(signs), _lib.ptr(parents), _lib.ptr(fc),
                                                   int(rows), int(bool(shared_batch)), float(sigma), _lib.ptr(uniforms), _lib.ptr(seq), _lib.ptr(lp)))
            return seq, lp
        if not trace and forced_tokens is None:
            _lib.check(self.lib.nes_rollout(self._h, P, _lib.ptr(offsets), _lib.ptr(signs), _lib.ptr(parents), _lib.ptr(fc),
                                            int(rows), int(bool(shared_batch)), float(sigma), int(precision),
                                            _lib.ptr(seq), _lib.ptr(lp)))
            return seq, lp
        h_tr = torch.zeros((T + 1, P, rows, R), dtype=torch.float32, device=self.dev) if trace else None
        l_tr = torch.zeros((T, P, rows, V1), dtype=torch.float32, device=self.dev) if trace else None
        if forced_tokens is not None:
            forced_tokens = torch.as_tensor(forced_tokens, dtype=torch.int64).contiguous()
            assert tuple(forced_tokens.shape) == (P, rows, T)
        _lib.check(self.lib.nes_rollout_trace(self._h, P, _lib.ptr(offsets), _lib.ptr(signs), _lib.ptr(parents), _lib.ptr(fc),
                                              int(rows), int(bool(shared_batch)), float(sigma), int(precision),
                                              _lib.ptr(forced_tokens), _lib.ptr(seq), _lib.ptr(lp), _lib.ptr(h_tr), _lib.ptr(l_tr)))
        self.sync()
        return seq, lp, h_tr, l_tr

    def logprob_fitness(self, n_members, rows, device='cpu'):
        out = torch.empty(int(n_members), dtype=torch.float64, device=device)
        _lib.check(self.lib.nes_logprob_fitness(self._h, int(n_members), int(rows), _lib.ptr(out)))
        return out

    FITNESS_KINDS = {'sc_loss': 0, 'greedy_logprob': 1, 'greedy_avgprob': 2, 'greedy_expprob': 3, 'greedy_linprob': 4}

    def shaped_fitness(self, kind, n_members, rows, scores=None, device='cpu'):
        """log-prob-shaped fitness (src/captioning/fitness.py) of the resident rollout: kind = the reference's Fitness value
        ('greedy_logprob', 'greedy_expprob', 'greedy_linprob', 'greedy_avgprob', 'sc_loss'); scores float64 [n_members, rows]
        per-caption CIDEr-D or None = the last CiderD.scores result still on the device.  -> float64 [n_members]"""
        if scores is not None:
            scores = torch.as_tensor(scores, dtype=torch.float64).contiguous()
            assert scores.numel() == int(n_members) * int(rows)
        out = torch.empty(int(n_members), dtype=torch.float64, device=device)
        _lib.check(self.lib.nes_shaped_fitness(self._h, int(self.FITNESS_KINDS[kind]), int(n_members), int(rows), _lib.ptr(scores), _lib.ptr(out)))
        return out

    # ---- master math
    def compute_ranks(self, x, device=False):
        x = torch.as_tensor(np.asarray(x, dtype=np.float64)).contiguous() if not torch.is_tensor(x) else x.to(torch.float64).contiguous()
        out = torch.empty(x.numel(), dtype=torch.int64, device=self.dev if device else 'cpu')
        _lib.check(self.lib.nes_compute_ranks(self._h, _lib.ptr(x), x.numel(), _lib.ptr(out)))
        return out

    def rank_weights(self, fitness, want_centered=True, want_weights=True):
        fit = fitness if torch.is_tensor(fitness) else torch.from_numpy(np.ascontiguousarray(fitness, dtype=np.float64))
        fit = fit.to(torch.float64).contiguous()
        F = fit.shape[0]
        assert fit.numel() == 2 * F
        cen = torch.empty((F, 2), dtype=torch.float64) if want_centered else None
        w = torch.empty(F, dtype=torch.float32) if want_weights else None      # None: the weights just stay on the device for K7
        _lib.check(self.lib.nes_rank_weights(self._h, _lib.ptr(fit), F, _lib.ptr(cen), _lib.ptr(w)))
        return cen, w

    STATIC_RANK_MAX_F = 32768

    def rank_weights_static(self, fitness, centered=None, weights=None):
        """nes_rank_weights_static: K6 for callers that cannot allocate.  fitness float64 [F, 2], centered float64 [F, 2] and weights
        float32 [F] are tensors ON THE CONTEXT'S DEVICE that the caller made beforehand (either output may be None); nothing is
        allocated, copied or synchronised here, the two kernels are enqueued on the context stream.  F <= STATIC_RANK_MAX_F."""
        assert torch.is_tensor(fitness) and fitness.dtype == torch.float64 and fitness.is_contiguous()
        F = fitness.shape[0]
        assert fitness.numel() == 2 * F
        assert centered is None or (centered.dtype == torch.float64 and centered.is_contiguous() and centered.numel() == 2 * F)
        assert weights is None or (weights.dtype == torch.float32 and weights.is_contiguous() and weights.numel() == F)
        _lib.check(self.lib.nes_rank_weights_static(self._h, _lib.ptr(fitness), F, _lib.ptr(centered), _lib.ptr(weights)))

    def gradient(self, offsets, weights, sigma, out_device='cpu', want=True, shard=(0, 1)):
        """K7.  weights None = the weights the last rank_weights call left on the device.  shard (r, n): only slice r of n of the
        parameter axis is summed, the rest of the vector is zero (all-reduce over the n ranks gives the whole)."""
        offsets = torch.as_tensor(offsets, dtype=torch.int64).contiguous()
        F = offsets.numel()
        if weights is not None:
            weights = torch.as_tensor(weights, dtype=torch.float32).contiguous()
            assert weights.numel() == F
        g = torch.empty(self.D, dtype=torch.float32, device=out_device) if want else None
        _lib.check(self.lib.nes_gradient_shard(self._h, _lib.ptr(offsets), _lib.ptr(weights), F, float(sigma), int(shard[0]), int(shard[1]), _lib.ptr(g)))
        return g

    def opt_init(self, kind, stepsize, momentum_or_beta1=0.9, beta2=0.999, epsilon=1e-8):
        _lib.check(self.lib.nes_opt_init(self._h, int(kind), float(stepsize), float(momentum_or_beta1), float(beta2), float(epsilon)))

    def opt_set_stepsize(self, stepsize):
        _lib.check(self.lib.nes_opt_set_stepsize(self._h, float(stepsize)))

    def opt_update(self, grad=None, l2coeff=0.0, want_ratio=True):
        if grad is not None:
            grad = torch.as_tensor(grad, dtype=torch.float32).contiguous()
            assert grad.numel() == self.D
        ratio = ctypes.c_double(0.0)
        _lib.check(self.lib.nes_opt_update(self._h, _lib.ptr(grad), float(l2coeff), ctypes.byref(ratio) if want_ratio else None))
        return ratio.value if want_ratio else None

    def opt_update_direct(self, globalg, want_ratio=True):
        """Optimizer.update(globalg): globalg float32 or float64 numpy / torch, length D"""
        if isinstance(globalg, np.ndarray):
            globalg = torch.from_numpy(np.ascontiguousarray(globalg))
        assert globalg.numel() == self.D and globalg.dtype in (torch.float32, torch.float64)
        globalg = globalg.contiguous()
        ratio = ctypes.c_double(0.0)
        _lib.check(self.lib.nes_opt_update_direct(self._h, _lib.ptr(globalg), int(globalg.dtype == torch.float64),
                                                  ctypes.byref(ratio) if want_ratio else None))
        return ratio.value if want_ratio else None

    def opt_get_state(self):
        m = torch.empty(self.D, dtype=torch.float64)
        v = torch.empty(self.D, dtype=torch.float64)
        t = ctypes.c_int64(0)
        _lib.check(self.lib.nes_opt_get_state(self._h, _lib.ptr(m), _lib.ptr(v), ctypes.byref(t)))
        return m, v, t.value

    def opt_set_state(self, m, v, t):
        m = None if m is None else torch.as_tensor(m, dtype=torch.float64).contiguous()
        v = None if v is None else torch.as_tensor(v, dtype=torch.float64).contiguous()
        _lib.check(self.lib.nes_opt_set_state(self._h, _lib.ptr(m), _lib.ptr(v), int(t)))

    def master_step(self, fitness, offsets, sigma, l2coeff=0.0, want_ratio=True):
        fit = fitness if torch.is_tensor(fitness) else torch.from_numpy(np.ascontiguousarray(fitness, dtype=np.float64))
        fit = fit.to(torch.float64).contiguous()
        offsets = torch.as_tensor(offsets, dtype=torch.int64).contiguous()
        F = offsets.numel()
        assert fit.numel() == 2 * F
        ratio = ctypes.c_double(0.0)
        _lib.check(self.lib.nes_master_step(self._h, _lib.ptr(fit), _lib.ptr(offsets), F, float(sigma), float(l2coeff),
                                            ctypes.byref(ratio) if want_ratio else None))
        return ratio.value if want_ratio else None

    def profile(self, on):
        _lib.check(self.lib.nes_profile_enable(self._h, int(bool(on))))

    def profile_read(self):
        ms = (ctypes.c_float * 6)()
        cnt = (ctypes.c_int32 * 6)()
        _lib.check(self.lib.nes_profile_read(self._h, ms, cnt))
        return {k: {'ms': float(ms[i]), 'count': int(cnt[i])} for i, k in enumerate(('plain', 'lstm', 'logit', 'finalize', 'fallback', 'materialise'))}

    def hybrid_stats(self, reset=True):
        out = (ctypes.c_int64 * 4)()
        _lib.check(self.lib.nes_hybrid_stats(self._h, out, int(reset)))
        return {'rows_refined': int(out[0]), 'candidates': int(out[1]), 'rows_overflowed': int(out[2])}

    def last_gradient_ms(self):
        ms = ctypes.c_float(0)
        _lib.check(self.lib.nes_last_gradient_ms(self._h, ctypes.byref(ms)))
        return ms.value

    def last_rollout_ms(self):
        ms = ctypes.c_float(0)
        _lib.check(self.lib.nes_last_rollout_ms(self._h, ctypes.byref(ms)))
        return ms.value


def gemm_selftest(A, W, E=None, sigma=0.0, precision=PREC_EXACT, device=0):
    lib = _lib.load()
    A = torch.as_tensor(A, dtype=torch.float32).contiguous()
    W = torch.as_tensor(W, dtype=torch.float32).contiguous()
    if E is not None:
        E = torch.as_tensor(E, dtype=torch.float32).contiguous()
    M, K = A.shape
    N = W.shape[0]
    C = torch.empty((M, N), dtype=torch.float32)
    _lib.check(lib.nes_gemm_selftest(int(device), M, N, K, _lib.ptr(A), _lib.ptr(W), _lib.ptr(E), float(sigma), int(precision), _lib.ptr(C)))
    return C
