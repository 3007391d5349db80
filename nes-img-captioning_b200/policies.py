"""Policy / CaptPolicy / PolicyFactory with the reference's interface (src/algorithm/policies.py:44-186,
src/captioning/policies.py:64-128).  `rollout` keeps its signature; the CIDEr-D scorer is off the CUDA
path (BASELINE.json north_star) and is injected as a callable.  `rollout_population` is the batched form
that the per-member evolve_model + rollout pair of NESWorker.fitness (nic_nes_worker.py:131-154)
degenerates to at F = 1."""
from collections import namedtuple
from enum import Enum

import numpy as np
import torch

from .cider import CiderD
from .context import NesContext, PREC_EXACT
from .nets import FCModel, MnistNet, ModelOptions, PolicyNet


class SuppDataset(Enum):
    MNIST = 'mnist'
    MSCOCO = 'mscoco'


class Net(Enum):
    MNIST = 'mnist'
    FC_CAPTION = 'fc_caption'


_opt_fields = ['net', 'safe_mutations', 'model_options', 'safe_mutation_underflow', 'fitness',
               'vbn', 'safe_mutation_batch_size', 'safe_mutation_vector']
PolicyOptions = namedtuple('PolicyOptions', field_names=_opt_fields, defaults=[None, '', {}, 0.01, None, False, 32, None])


class Policy:
    def __init__(self, dataset, options: PolicyOptions, ctx: NesContext = None, **ctx_kwargs):
        assert isinstance(dataset, SuppDataset)
        self.dataset, self.options = dataset, options
        self.net = Net(options.net)
        self.vbn = bool(options.vbn)
        self.ref_batch = None
        mo = dict(options.model_options)
        mo['vbn'] = self.vbn
        self.model_options = ModelOptions(**mo)
        self.policy_net: PolicyNet = None
        self._ctx, self._ctx_kwargs = ctx, ctx_kwargs
        self.init_model(self.generate_model())

    # ---- src/algorithm/policies.py:102-111
    def parameter_vector(self):
        assert self.policy_net is not None, 'set model first!'
        return self.policy_net.parameter_vector()

    def set_from_parameter_vector(self, vector):
        assert self.policy_net is not None, 'set model first!'
        assert isinstance(vector, np.ndarray) or isinstance(vector, torch.Tensor)
        if isinstance(vector, np.ndarray):
            vector = torch.from_numpy(vector)
        self.policy_net.set_from_vector(vector)

    def nb_learnable_params(self):
        assert self.policy_net is not None, 'set model first!'
        return self.policy_net.get_nb_learnable_params()

    def serialized(self, path=''):
        return self.policy_net.serialize(path=path)

    def save(self, path, filename):
        """src/algorithm/policies.py:93-97"""
        import os
        assert self.policy_net is not None, 'set model first!'
        os.makedirs(path, exist_ok=True)
        assert not os.path.exists(os.path.join(path, filename))
        torch.save(self.policy_net.state_dict(), os.path.join(path, filename))

    def init_model(self, model=None):
        assert self.policy_net is None
        assert isinstance(model, PolicyNet), '{}'.format(type(model))
        self.policy_net = model

    def set_model(self, model):
        assert self.policy_net is not None
        assert isinstance(model, PolicyNet) or isinstance(model, dict) or isinstance(model, str), '{}'.format(type(model))
        if isinstance(model, PolicyNet):
            self.policy_net.load_state_dict(model.state_dict())
        elif isinstance(model, dict):
            self.policy_net.load_state_dict(model)
        else:
            self.policy_net.from_serialized(model)

    def generate_model(self, from_param_file=None, start_rng=None):
        cls = MnistNet if self.net == Net.MNIST else FCModel              # src/algorithm/policies.py:71-74
        model = cls(self._ctx, 0, self.model_options, rng_state=start_rng, from_param_file=from_param_file,
                    vbn=self.vbn, **self._ctx_kwargs)
        self._ctx = model.ctx
        return model

    def evolve_model(self, sigma):
        assert self.policy_net is not None, 'set model first!'
        return self.policy_net.evolve(sigma)

    def get_model(self):
        return self.policy_net

    def calc_sensitivity(self, task_id, parent_id, experiences, batch_size, directory):
        assert self.policy_net is not None, 'set model first!'
        self.policy_net.calc_sensitivity(task_id, parent_id, experiences, batch_size, directory)

    def set_ref_batch(self, ref_batch):
        self.ref_batch = ref_batch

    def rollout(self, placeholder, data, config):
        raise NotImplementedError

    def accuracy_on(self, dataloader, config, directory) -> float:
        raise NotImplementedError('validation needs the COCO/Java scorers, which are outside the hot path (SURVEY 2.1)')


class CaptPolicy(Policy):
    """scorer(seq int64 [rows, T] numpy, data dict) -> float fitness (the reference's CIDEr-D * 100,
    src/captioning/policies.py:117-125,145-193).  The reference forwards batch_size*seq_per_img rows that
    are 5x replicated images (dataloader.py:175); greedy decode of a replica is identical, so unique rows
    are decoded once and token rows repeated (`seq_per_img`)."""

    def __init__(self, dataset, options, ctx=None, scorer=None, seq_per_img=1, **ctx_kwargs):
        super().__init__(dataset, options, ctx, **ctx_kwargs)
        self.scorer, self.seq_per_img = scorer, int(seq_per_img)
        self.precision = PREC_EXACT
        self.validation_fn = None        # callable(policy_net, dataloader, config, directory) -> float: the validation run of accuracy_on

    @property
    def ctx(self):
        return self.policy_net.ctx

    def _unique_rows(self, fc):
        fc = np.asarray(fc, dtype=np.float32)
        return fc[::self.seq_per_img] if self.seq_per_img > 1 else fc

    def accuracy_on(self, dataloader, config, directory) -> float:
        """validation score (captioning/policies.py:130-143).  The reference decodes the validation split and scores it with the COCO
        caption evaluators (Java); that run is outside the hot path, so it is whatever the caller plugs in."""
        assert directory is not None
        if self.validation_fn is None:
            raise NotImplementedError('validation is outside the hot path: set policy.validation_fn = callable(policy_net, dataloader, config, directory) -> float')
        return float(self.validation_fn(self.policy_net, dataloader, config, directory))

    def rollout_tokens(self, data):
        seq, lp = self.policy_net(torch.from_numpy(np.ascontiguousarray(self._unique_rows(data['fc_feats']))), greedy=True)
        if self.seq_per_img > 1:
            seq, lp = seq.repeat_interleave(self.seq_per_img, 0), lp.repeat_interleave(self.seq_per_img, 0)
        return seq, lp

    def rollout(self, placeholder, data, config):
        """fitness of this individual on the batch (captioning/policies.py:86-128, `greedy` fitness = CIDEr-D * 100).
        scorer = nes_b200.CiderD(ctx, ...): scored on the device from the resident tokens; any other callable
        scorer(seq [rows, T] numpy, data) -> float runs on the host like the reference's."""
        assert self.scorer is not None, 'CaptPolicy.rollout needs a scorer (nes_b200.CiderD or a callable)'
        if isinstance(self.scorer, CiderD):
            return float(self.rollout_population_fitness(np.zeros(1, dtype=np.int64), np.ones(1, dtype=np.float32), data, 0.0,
                                                         parents=np.array([self.policy_net.parent], dtype=np.int32))[0])
        seq, _ = self.rollout_tokens(data)
        return float(self.scorer(seq.numpy(), data))

    def rollout_population_fitness(self, offsets, signs, data, sigma, parents=None, uniforms=None):
        """CIDEr-D * 100 of every member (float64 numpy [P]): decode on the device, score the resident tokens on the device
        against data['gts'] (list over images of int arrays [n_refs, T]); only P x rows float64 scores travel to the host."""
        assert isinstance(self.scorer, CiderD), 'needs scorer = nes_b200.CiderD(ctx, ...)'
        fc = torch.from_numpy(np.ascontiguousarray(self._unique_rows(data['fc_feats'])))
        self.scorer.set_refs(data['gts'], seq_per_img=self.seq_per_img, rows_per_img=1)
        P = len(offsets)
        self.ctx.rollout(offsets, signs, fc, sigma, parents=parents, precision=self.precision, out_device=self.ctx.dev)
        kind = self.options.fitness or 'greedy'
        if kind in ('sample', 'self_critical', 'sc_loss'):
            return self._sampled_population_fitness(kind, offsets, signs, data, sigma, parents, uniforms)
        if kind == 'greedy':
            return self.scorer.fitness(P, fc.shape[0])                      # float(cider * 100), policies.py:125
        assert kind in self.ctx.FITNESS_KINDS, 'unknown fitness %r' % kind
        self.scorer.scores(None, P, fc.shape[0], device=self.ctx.dev)       # per-caption CIDEr-D stays on the device
        return self.ctx.shaped_fitness(kind, P, fc.shape[0]).numpy()        # crit(logprobs, seq, ciders), policies.py:119-123

    def _sampled_population_fitness(self, kind, offsets, signs, data, sigma, parents, uniforms):
        """the fitnesses that decode by sampling (Fitness.SAMPLE / SELF_CRITICAL / SC_LOSS, captioning/policies.py:27-30,113-125,164-186).
        Replicated rows draw independently, so all batch_size * seq_per_img rows are decoded; the self-critical baseline is the
        greedy caption of the image (unique rows, repeated).  uniforms float64 [P, rows, T] or None = torch.rand on the device
        (the reference: the unseeded global numpy generator)."""
        ctx, spi = self.ctx, self.seq_per_img
        fc_all = torch.from_numpy(np.ascontiguousarray(np.asarray(data['fc_feats'], dtype=np.float32)))
        P, rows, T = len(offsets), fc_all.shape[0], ctx.dims['seq_length']
        sc = kind in ('self_critical', 'sc_loss')
        # compute_ciders scores 2 * batch captions in self-critical mode, so every image is a corpus document 2 * seq_per_img times (:175-177)
        self.scorer.set_refs(data['gts'], seq_per_img=(2 * spi if sc else spi), rows_per_img=1)
        base = None
        if sc:
            fc_u = torch.from_numpy(np.ascontiguousarray(self._unique_rows(data['fc_feats'])))
            ctx.rollout(offsets, signs, fc_u, sigma, parents=parents, precision=self.precision, out_device=ctx.dev)
            base = self.scorer.scores(None, P, fc_u.shape[0], device=ctx.dev).repeat_interleave(spi, 1)
        if uniforms is None:
            uniforms = torch.rand((P, rows, T), dtype=torch.float64, device=ctx.dev)
        ctx.rollout(offsets, signs, fc_all, sigma, parents=parents, out_device=ctx.dev, uniforms=uniforms)
        scores = self.scorer.scores(None, P, rows, device=ctx.dev, rows_per_img=spi)
        if sc:
            scores = scores - base                                          # policies.py:183
        if kind == 'sc_loss':
            return ctx.shaped_fitness(kind, P, rows, scores=scores).numpy() # LogFitnessCriterion(sample_logprobs, gen_result, rewards)
        s = scores.cpu().numpy()
        return np.array([float(np.mean(s[m]) * 100) for m in range(P)], dtype=np.float64)

    def rollout_population(self, offsets, signs, data, sigma, parents=None, out_device='cpu'):
        """all members at once -> seq int64 [P, rows, T], logprobs float32 [P, rows, T] (unique rows)."""
        fc = torch.from_numpy(np.ascontiguousarray(self._unique_rows(data['fc_feats'])))
        return self.ctx.rollout(offsets, signs, fc, sigma, parents=parents, precision=self.precision, out_device=out_device)


class ClfPolicy(Policy):
    """src/classification/policies.py:14-86 for the MNIST net: fitness = -CrossEntropyLoss on a batch, accuracy over a loader"""

    @property
    def ctx(self):
        return self.policy_net.ctx

    def rollout(self, placeholder, data, config):
        """:16-48 -- `placeholder` is the reference's reusable input tensor; unused here (the batch goes to the device as it is)"""
        assert self.policy_net is not None, 'Set model first!'
        assert isinstance(self.policy_net, PolicyNet), '{}'.format(type(self.policy_net))
        inputs, labels = data
        fit, _ = self.ctx.mnist_rollout(torch.zeros(1, dtype=torch.int64), torch.ones(1), inputs, labels, 0.0,
                                        parents=torch.tensor([self.policy_net.parent], dtype=torch.int32), shared_batch=True)
        return float(np.float32(fit[0].item()))

    def rollout_population(self, offsets, signs, data, sigma, parents=None, out_device='cpu'):
        """batched form of ESWorker.fitness's evolve_model + rollout (nic_es_worker.py:142-192): fitness float64 [P], correct int32 [P]"""
        inputs, labels = data
        return self.ctx.mnist_rollout(offsets, signs, inputs, labels, sigma, parents=parents, out_device=out_device)

    def accuracy_on(self, dataloader, config, directory) -> float:
        """:50-86 -- the mean of the per-batch accuracies (the reference's own caveat about a smaller last batch applies)"""
        assert self.policy_net is not None, 'Set model first!'
        accuracies = []
        end = config.num_val_batches if getattr(config, 'num_val_batches', None) else len(dataloader)
        for i, (inputs, labels) in enumerate(dataloader):
            if i >= end:
                break
            _, cor = self.ctx.mnist_rollout(torch.zeros(1, dtype=torch.int64), torch.ones(1), inputs, labels, 0.0,
                                            parents=torch.tensor([self.policy_net.parent], dtype=torch.int32), shared_batch=True)
            accuracies.append(float(cor[0].item()) / torch.as_tensor(labels).numel())
        return np.mean(accuracies).item()


class PolicyFactory:
    @staticmethod
    def create(dataset, exp, **kw):
        """src/algorithm/policies.py:175-186"""
        options = PolicyOptions(**exp['policy_options'])
        if dataset == SuppDataset.MNIST:
            return ClfPolicy(dataset, options, **kw)
        if dataset == SuppDataset.MSCOCO:
            return CaptPolicy(dataset, options, **kw)
        raise NotImplementedError
