"""PolicyNet / FCModel with the reference's interface (src/algorithm/nets.py:24-168,
src/captioning/nets.py:138-245), backed by a NesContext instead of nn.Module parameters.

Differences that follow from the design (documented in DESIGN.md):
  * evolve() draws its noise from the shared noise table (a fresh offset per call) instead of an
    unseeded torch.normal_ -- the returned vector is still float32 sigma*eps of length D;
  * set_from_vector() copies (the reference aliases the caller's memory, nets.py:163-165);
  * every Mutation runs on the CUDA path; for SM-G-SUM / SM-G-ABS the sensitivity vector itself comes from autograd
    (sensitivity.py here) and the kernels divide the noise by it.
"""
from collections import OrderedDict, namedtuple
from enum import Enum

import numpy as np
import torch

from .context import NesContext, PREC_EXACT
from . import sensitivity as _sm
from .layout import mnist_tensors
from .sensitivity import Sensitivity

_model_opt_fields = ['vocab_size', 'input_encoding_size', 'rnn_type', 'rnn_size', 'num_layers',
                     'drop_prob_lm', 'seq_length', 'fc_feat_size', 'vbn', 'vbn_e', 'vbn_affine', 'layer_n',
                     'layer_n_affine', 'safe_mutation_underflow', 'safe_mutations', 'safe_mutation_vector',
                     'safe_mutation_batch_size']
ModelOptions = namedtuple('ModelOptions', field_names=_model_opt_fields, defaults=(0,) * len(_model_opt_fields))


class Mutation(Enum):
    SAFE_GRAD_SUM = 'SM-G-SUM'
    SAFE_GRAD_ABS = 'SM-G-ABS'
    SAFE_VECTOR = 'SM-VECTOR'
    SAFE_PROPORTIONAL = 'SM-PROPORTIONAL'
    DEFAULT = ''


class PolicyNet:
    """Flat-theta get/set and Gaussian mutation for one individual (parent slot `parent` of `ctx`)."""

    def __init__(self, ctx: NesContext, parent=0, options=None, vbn=False):
        self.ctx, self.parent = ctx, int(parent)
        sm = options.get('safe_mutations') if isinstance(options, dict) else getattr(options, 'safe_mutations', None)
        self.mutations = Mutation(sm or '')
        ctx.set_mutation(self.mutations.value)        # context-wide: every parent slot mutates the same way
        self._underflow = (options.get('safe_mutation_underflow') if isinstance(options, dict) else getattr(options, 'safe_mutation_underflow', None)) or 0.01
        if self.mutations in (Mutation.SAFE_GRAD_SUM, Mutation.SAFE_GRAD_ABS):     # algorithm/nets.py:46-48
            self.sensitivity_wrapper = Sensitivity(self, self._underflow, self.mutations)
        if self.mutations == Mutation.SAFE_VECTOR:     # algorithm/nets.py:49-50
            vec = options.get('safe_mutation_vector') if isinstance(options, dict) else getattr(options, 'safe_mutation_vector', None)
            if vec is not None and not (isinstance(vec, (int, float)) and not vec):
                self.set_sensitivity_vector(vec)
        if vbn and ctx.dims is not None and ctx.dims['layer_n'] != 2:
            raise AssertionError('vbn=True needs a context created with model_options.vbn')
        self.nb_learnable_params = ctx.D
        self.nb_params = ctx.D
        self._theta_dtype = torch.float32

    # -- src/algorithm/nets.py:163-168
    def set_from_vector(self, vector):
        assert len(vector) == self.nb_learnable_params
        if isinstance(vector, np.ndarray):
            vector = torch.from_numpy(np.ascontiguousarray(vector))
        self._theta_dtype = vector.dtype if vector.dtype == torch.float64 else torch.float32
        self.ctx.set_theta(vector.to(self._theta_dtype), self.parent)

    def parameter_vector(self):
        return self.ctx.get_theta(self.parent, dtype=self._theta_dtype)

    def get_nb_learnable_params(self):
        return self.nb_learnable_params

    def count_parameters(self):
        return self.nb_params

    # -- src/algorithm/nets.py:83-119
    def evolve(self, sigma, rng_state=None, offset=None):
        """theta <- theta + sigma*eps(offset); returns the float32 noise vector (numpy, owned)."""
        if offset is None:
            it = self.ctx.next_evolve_iteration() if rng_state is None else (1 << 40) + int(rng_state)
            offset = int(self.ctx.sample_offsets(it, 1)[0])
        member, noise = self.ctx.evolve(offset, sigma, +1.0, self.parent, device=self.ctx.dev)
        self.ctx.set_theta(member, self.parent)
        self._theta_dtype = torch.float32
        self.last_offset = offset
        return noise.cpu().numpy().astype(np.float32)

    # -- state_dict compatibility with the reference's .pth files (named tensors in registration order)
    def state_dict(self):
        flat = self.ctx.get_theta(self.parent, dtype=torch.float32)
        return OrderedDict((t.name, flat[t.start:t.start + t.count].view(*t.shape).clone()) for t in self.ctx.tensors)

    def load_state_dict(self, sd):
        flat = torch.empty(self.ctx.D, dtype=torch.float32)
        for t in self.ctx.tensors:
            v = sd[t.name]
            assert tuple(v.shape) == tuple(t.shape), (t.name, tuple(v.shape), t.shape)
            flat[t.start:t.start + t.count] = v.reshape(-1).to(torch.float32)
        self.set_from_vector(flat)

    def serialize(self, path=''):
        torch.save(self.state_dict(), path)
        return path

    def from_serialized(self, serialized):
        self.load_state_dict(torch.load(serialized, map_location='cpu'))

    def calc_sensitivity(self, task_id, parent_id, experiences, batch_size, directory):
        """algorithm/nets.py:121-123: a no-op unless the mutation is gradient based.  SM-G-SUM / SM-G-ABS: the output-gradient
        sensitivity of this parent (safe_mutations.py:34-147; autograd on the device, cached per (task_id, parent_id) in memory and
        in sens_t*_p*_*.pt under `directory` like the reference) and handed to the device, where noise /= sensitivity"""
        if self.mutations in (Mutation.SAFE_GRAD_SUM, Mutation.SAFE_GRAD_ABS):
            if self.sensitivity_wrapper.calc_sensitivity(task_id, parent_id, experiences, batch_size, directory):
                self._sensitivity = self.sensitivity_wrapper.get_sensitivity().to(torch.float32)
                self.ctx.set_sensitivity(self._sensitivity, parent=self.parent)
        return None

    def set_sensitivity_vector(self, vector):
        """algorithm/nets.py:125-126 -> Sensitivity.set_sensitivity (safe_mutations.py:27-32): a path to a torch-saved vector, or
        the vector itself; clamped below at safe_mutation_underflow and divided by its minimum, then noise /= sensitivity"""
        if isinstance(vector, str):
            vector = torch.load(vector, map_location='cpu')
        self._sensitivity = torch.as_tensor(vector).to(torch.float32).clone()
        self._sensitivity[self._sensitivity < self._underflow] = self._underflow
        self._sensitivity /= self._sensitivity.min()
        self.ctx.set_sensitivity(self._sensitivity, parent=self.parent if self.mutations != Mutation.SAFE_VECTOR else None)

    def get_sensitivity_vector(self):
        return getattr(self, '_sensitivity', None)


class FCModel(PolicyNet):
    """fc_caption: img_embed -> maxout-LSTM core -> logit, greedy decode (src/captioning/nets.py:138-245)."""

    def __init__(self, ctx: NesContext = None, parent=0, options=None, rng_state=None, from_param_file=None,
                 vbn=False, **ctx_kwargs):
        if ctx is None:
            ctx = NesContext(options, **ctx_kwargs)
        super().__init__(ctx, parent, options, vbn)
        self.seq_length = ctx.dims['seq_length']
        self.precision = PREC_EXACT
        if from_param_file:
            self.from_serialized(from_param_file)
        else:
            self.initialize_params(rng_state)

    def initialize_params(self, rng_state=None):
        """Xavier-normal weights, zero biases, LayerNorm at 1/0 (src/algorithm/nets.py:62-69)."""
        gen = torch.Generator()
        gen.manual_seed(int(rng_state) if rng_state is not None else torch.seed() % (1 << 31))
        flat = torch.zeros(self.ctx.D, dtype=torch.float32)
        for t in self.ctx.tensors:
            if 'ln' in t.name or 'bn' in t.name or '1' in t.name:      # algorithm/nets.py:62-64: skipped, torch's default 1 / 0
                if t.name.endswith('weight'):
                    flat[t.start:t.start + t.count] = 1.0
            elif t.name.endswith('weight'):
                fan_out, fan_in = t.shape
                std = (2.0 / (fan_in + fan_out)) ** 0.5
                flat[t.start:t.start + t.count] = torch.randn(t.count, generator=gen) * std
        self.set_from_vector(flat)

    def sensitivity_forward(self, experiences, orig_bs):
        """(theta, i) -> FCModel.forward_for_sensitivity(data, orig_bs, i) (captioning/nets.py:22-69) as a function of the flat theta"""
        tensors, dims, fc = self.ctx.tensors, self.ctx.dims, experiences['fc_feats']
        return lambda theta, i: _sm.forward_for_sensitivity(theta, tensors, dims, fc, orig_bs, i=i)

    def forward(self, fc_feats, greedy=True):
        """-> seq int64 [B, T], seq_logprobs float32 [B, T]  (FCModel._sample, nets.py:183-245)"""
        if not greedy:
            raise NotImplementedError('sampled decode is not on the B200 path (SURVEY 8f-2)')
        off = torch.zeros(1, dtype=torch.int64)
        sg = torch.ones(1, dtype=torch.float32)
        par = torch.tensor([self.parent], dtype=torch.int32)
        seq, lp = self.ctx.rollout(off, sg, fc_feats, 0.0, parents=par, precision=self.precision, shared_batch=True)
        return seq[0], lp[0]

    __call__ = forward


class MnistNet(PolicyNet):
    """The reference's MNIST classifier (src/classification/nets.py:8-37; BASELINE configs[0], experiments/mnist_es.json): conv 1->10 5x5,
    relu, maxpool, conv 10->20 5x5, relu, maxpool, fc 320->10; D = 8490.  Backed by a flat NesContext with the six named tensors; the
    forward pass runs in k_mnist_rollout (csrc/mnist.cu).  vbn (BatchNorm2d over the batch) is not on the CUDA path."""

    def __init__(self, ctx: NesContext = None, parent=0, options=None, rng_state=None, from_param_file=None, vbn=False, **ctx_kwargs):
        assert not vbn, 'MnistNet with vbn is not on the B200 path'
        if ctx is None:
            ctx_kwargs.setdefault('table_len', 1 << 22)
            ctx = NesContext(tensors=mnist_tensors(), **ctx_kwargs)
        assert [t.name for t in ctx.tensors] == [n for n, _ in mnist_tensors()] and ctx.D == 8490
        super().__init__(ctx, parent, options, vbn)
        if from_param_file:
            self.from_serialized(from_param_file)
        else:
            self.initialize_params(rng_state)

    def initialize_params(self, rng_state=None):
        """src/algorithm/nets.py:55-69: Xavier-normal weights / zero biases, EXCEPT parameters whose name contains '1' (conv1, fc1), which
        keep torch's default nn.Conv2d / nn.Linear init: U(-1/sqrt(fan_in), 1/sqrt(fan_in)) for weight and bias"""
        gen = torch.Generator()
        gen.manual_seed(int(rng_state) if rng_state is not None else torch.seed() % (1 << 31))
        flat = torch.zeros(self.ctx.D, dtype=torch.float32)
        for t in self.ctx.tensors:
            layer = t.name.split('.')[0]
            w_shape = dict(mnist_tensors())[layer + '.weight']
            fan_in = int(np.prod(w_shape[1:]))
            fan_out = w_shape[0] * int(np.prod(w_shape[2:]))
            if '1' in t.name:
                bound = 1.0 / fan_in ** 0.5
                flat[t.start:t.start + t.count] = (torch.rand(t.count, generator=gen) * 2 - 1) * bound
            elif t.name.endswith('weight'):
                flat[t.start:t.start + t.count] = torch.randn(t.count, generator=gen) * (2.0 / (fan_in + fan_out)) ** 0.5
        self.set_from_vector(flat)

    def sensitivity_forward(self, experiences, orig_bs):
        """MnistNet.forward_for_sensitivity (classification/nets.py:39-49): the logits of the whole batch, or of sample i"""
        images = torch.as_tensor(experiences[0])
        tensors = self.ctx.tensors
        return lambda theta, i: _sm.mnist_forward(theta, tensors, images if i < 0 else images[i:i + 1])

    def forward(self, x):
        """logits float32 [B, 10] of this individual"""
        x = torch.as_tensor(x).to(torch.float32)
        labels = torch.zeros(x.shape[0], dtype=torch.int64, device=x.device)
        off = torch.zeros(1, dtype=torch.int64)
        _, _, lg = self.ctx.mnist_rollout(off, torch.ones(1), x, labels, 0.0, parents=torch.tensor([self.parent], dtype=torch.int32),
                                          shared_batch=True, out_device=x.device if x.is_cuda else 'cpu', want_logits=True)
        return lg[0]

    __call__ = forward
