"""What the reference's masters carry from one iteration to the next: the annealing schedule, the patience
counter, the podium of the best elites, the running statistics and the `z_info_e*_i*-*.json` snapshot that ties them together,
so that a run of the reference can be resumed on this path and a run of this path by the reference.

  Config        the "config" block of an experiment JSON          src/algorithm/tools/utils.py:14-20
  Podium        E best (file, score) pairs, good / bad generation  src/algorithm/tools/podium.py:10-74
  Iteration     counters + curriculum (sigma /= stdev_divisor, batch_size = int(batch_size * bs_multiplier)) driven either by
                a fixed schedule or by patience                   src/algorithm/tools/iteration.py:11-235
  ESIteration   + parents / elite candidates                      src/algorithm/nic_es/iteration.py:9-153
  NESIteration  + the current model file                          src/algorithm/nic_nes/iteration.py:10-71
  Statistics    the lists a snapshot stores                       src/algorithm/tools/statistics.py:13-71,121-175
  save_infos / find_infos / load_infos                            src/algorithm/tools/snapshot.py:23-38, utils.py:149-153

Host-side state only: nothing here touches the device.  The population itself lives in HBM (parent slots of a NesContext); the
file names kept here are the checkpoint files written through PolicyNet.serialize, which is all the reference's snapshot holds.
The Redis task / result plumbing around these classes is out of scope."""
import json
import logging
import os
import re
import shutil
import time
from collections import namedtuple

import numpy as np

logger = logging.getLogger(__name__)

CONFIG_FIELDS = ('l2coeff', 'noise_stdev', 'stdev_divisor', 'eval_prob', 'snapshot_freq', 'log_dir', 'batch_size', 'patience',
                 'val_batch_size', 'num_val_batches', 'num_val_items', 'cuda', 'max_nb_iterations', 'ref_batch_size',
                 'bs_multiplier', 'stepsize_divisor', 'single_batch', 'schedule_limit', 'schedule_start')
Config = namedtuple('Config', CONFIG_FIELDS, defaults=(None,) * len(CONFIG_FIELDS))

INFOS_PATTERN = r'z_info_e[0-9]*?_i[0-9]*?-[0-9]*?.json'


class FileStore:
    """Where the models that the bookkeeping moves around live.  The reference keeps every model as a `.pth` file and copies / renames /
    removes those (tools/utils.py:108-141); this is that behaviour.  `SlotStore` below keeps them as parent slots of a NesContext instead,
    so that the same bookkeeping (pinned against the reference's classes) drives a population that never leaves the device."""

    def makedirs(self, directory): os.makedirs(directory, exist_ok=True)
    def isfile(self, path): return os.path.isfile(path)
    def copy(self, src, dst): shutil.copy(src, dst)

    def rename(self, src, dst):
        if os.path.exists(dst):
            os.remove(dst)
        os.rename(src, dst)

    def only_keep(self, directory, keep):
        for name in os.listdir(directory):
            path = os.path.join(directory, name)
            if os.path.isfile(path) and path not in keep:
                os.remove(path)


class SlotStore:
    """Models as parent slots of a NesContext: a path is a name for a slot, copy = nes_copy_parent in stream order (so the podium's
    sequential copies alias exactly as the reference's file copies do), rename = the name moves to the other slot, remove = the slot
    returns to the free pool.  `tags[path]` says where the slot's content came from (tests compare it with the content of the
    reference's files).  Slots handed in through `adopt` (the survivors nes_es_select_evolve wrote) belong to the store from then on."""

    def __init__(self, ctx, free_slots):
        self.ctx, self.free = ctx, list(free_slots)
        self.pool = set(self.free)          # only these return to the free list; adopted slots belong to whoever handed them in
        self.slot_of, self.tags = {}, {}

    def makedirs(self, directory): pass
    def isfile(self, path): return path in self.slot_of
    def slot(self, path): return self.slot_of[path]

    def adopt(self, path, slot, tag=None):
        assert path not in self.slot_of and slot not in self.slot_of.values()
        if slot in self.free:
            self.free.remove(slot)
        self.slot_of[path], self.tags[path] = slot, tag

    def _release(self, path):
        slot = self.slot_of.pop(path)
        if slot in self.pool:
            self.free.append(slot)
        self.tags.pop(path, None)

    def copy(self, src, dst):
        if src == dst:
            return
        if dst not in self.slot_of:
            if not self.free:
                raise OSError('no free parent slot for %s' % dst)
            self.slot_of[dst] = self.free.pop(0)
        self.ctx.copy_parent(self.slot_of[src], self.slot_of[dst])
        self.tags[dst] = self.tags.get(src)

    def rename(self, src, dst):
        if dst in self.slot_of:
            self._release(dst)
        self.slot_of[dst], self.tags[dst] = self.slot_of.pop(src), self.tags.pop(src, None)

    def only_keep(self, directory, keep):
        prefix = directory.rstrip(os.sep) + os.sep
        for path in [p for p in self.slot_of if p.startswith(prefix) and os.sep not in p[len(prefix):] and p not in keep]:
            self._release(path)


_FILES = FileStore()


class Podium:
    """The `num_elites` best (file, score) pairs seen so far, files copied to <directory>/best_elite/0_<rank>_elite.pth.  A
    generation is good when at least one file moved onto (or within) the podium; reading the flag resets it."""

    def __init__(self, patience, directory, num_elites, store=_FILES):
        self._n = num_elites
        self._patience = patience
        self._store = store
        self._dir = os.path.join(directory, 'best_elite')
        store.makedirs(self._dir)
        self._slots = [('', float('-inf'))] * num_elites
        self._bad = True

    def _slot_path(self, rank):
        return os.path.join(self._dir, '0_{}_elite.pth'.format(rank))

    def init_from_infos(self, infos):
        slots = []
        for rank, (path, score) in enumerate(infos['best_elites']):
            self._store.copy(path, self._slot_path(rank))
            slots.append((self._slot_path(rank), score))
        self._slots = slots

    def record_elites(self, elites_and_scores):
        # stable descending sort: on equal scores the incumbent (listed first) stays ahead, as sorted(..., reverse=True) keeps it
        ranked = sorted(list(self._slots) + list(elites_and_scores), key=lambda c: c[1], reverse=True)[:self._n]
        slots = []
        for rank, (path, score) in enumerate(ranked):
            if not path:
                continue                                    # an empty initial slot: dropped, ranks are NOT renumbered
            dst = self._slot_path(rank)
            slots.append((dst, score))
            if path != dst:
                try:
                    self._store.copy(path, dst)
                    self._bad = False
                except OSError:
                    logging.error('[Podium]: tried to copy non existing elite')
        self._slots = slots
        self._store.only_keep(self._dir, [p for p, _ in slots])

    def is_bad_generation(self):
        bad, self._bad = self._bad, True
        logger.info('BAD GENERATION' if bad else 'GOOD GENERATION')
        return bad

    def best_elites(self):
        return self._slots


class Iteration:
    """Counters of a run and the two ways the curriculum advances: every `schedule_limit` iterations from `schedule_start` on,
    or after more than `patience` generations without a new podium entry."""

    _INFO_KEYS = (('epoch', '_epoch', -1), ('iter', '_iteration', -1), ('bad_generations', '_bad_generations', 0),
                  ('noise_stdev', '_noise_stdev', 0), ('batch_size', '_batch_size', 0), ('times_orig_bs', '_times_orig_bs', 0),
                  ('nb_samples_used', '_nb_samples_used', 0))

    def __init__(self, config, exp, store=_FILES):
        self._store = store
        self._noise_stdev, self._batch_size = config.noise_stdev, config.batch_size
        self._times_orig_bs, self._nb_samples_used = 1, 0
        self._bad_generations, self._epoch, self._iteration = 0, 0, 0
        self._patience_reached = self._schedule_reached = False
        self._schedule_limit = config.schedule_limit
        self._schedule_start = config.schedule_start or 0
        self._stdev_divisor, self._bs_multiplier, self._patience = config.stdev_divisor, config.bs_multiplier, config.patience
        self._nb_offspring = exp['nb_offspring']
        self._nb_models_to_evaluate = 0
        self._task_results, self._eval_results, self._worker_ids = [], {}, []
        self._waiting_for_eval_run = False
        self._log_dir = exp['log_dir']
        self._models_dir = os.path.join(self._log_dir, 'models')
        self._podium = Podium(config.patience, os.path.join(self._models_dir, 'best'), num_elites=exp['num_elites'], store=store)

    # ---- snapshot
    def to_dict(self):
        return {'iter': self._iteration, 'epoch': self._epoch, 'noise_stdev': self._noise_stdev, 'batch_size': self._batch_size,
                'bad_generations': self._bad_generations, 'times_orig_bs': self._times_orig_bs,
                'nb_samples_used': self._nb_samples_used, 'best_elites': self.best_elites()}

    def init_from_infos(self, infos):
        # epoch and iteration restart one below the stored value: the loop increments before it works
        for key, attr, delta in self._INFO_KEYS:
            if key in infos:
                setattr(self, attr, infos[key] + delta)
        self._podium.init_from_infos(infos)

    # ---- curriculum
    def next_curriculum_step(self):
        self._noise_stdev /= self._stdev_divisor
        self._batch_size = int(self._batch_size * self._bs_multiplier)
        self._times_orig_bs *= self._bs_multiplier

    def check_schedule_limit(self):
        return self._schedule_limit and self._iteration >= self._schedule_start and \
            (self._iteration - self._schedule_start) % self._schedule_limit == 0

    def incr_iteration(self):
        self._task_results, self._eval_results, self._worker_ids = [], {}, []
        self._nb_models_to_evaluate = self._nb_offspring
        self._waiting_for_eval_run = False
        self._patience_reached = self._schedule_reached = False
        self._iteration += 1
        self._nb_samples_used += self._batch_size
        if self.check_schedule_limit():
            self._schedule_reached = True
            old = (self._noise_stdev, self._batch_size)
            self.next_curriculum_step()
            logging.warning('Next curriculum step reached; std %s -> %s, bs %s -> %s', old[0], self._noise_stdev, old[1], self._batch_size)

    def process_evaluated_elites(self):
        """Hand this iteration's evaluated elite candidates (those whose file exists) to the podium; count a bad generation when
        patience is on and nothing new reached the podium.  Returns the best (score, file) of the candidates."""
        cands = [(ind, sc) for ind, sc in self._eval_results.values() if self._store.isfile(ind)]
        best_sc, best_ind = float('-inf'), None
        for ind, sc in cands:
            if sc > best_sc:
                best_sc, best_ind = sc, ind
        self._podium.record_elites(cands)
        if self._patience and self._podium.is_bad_generation():
            self._bad_generations += 1
            if self._bad_generations > self._patience:
                self.next_curriculum_step()
                self._patience_reached = True
                self._bad_generations = 0
                logging.warning('Max patience reached; new std %s, bs %s', self._noise_stdev, self._batch_size)
        else:
            self._bad_generations = 0
        return best_sc, best_ind

    def anneal_optimizer(self, optimizer, config):
        """`optimizer.stepsize /= stepsize_divisor` when this iteration advanced the curriculum (nic_nes_master.py:139-141)."""
        if (self._patience_reached or self._schedule_reached) and config.stepsize_divisor:
            optimizer.stepsize = optimizer.stepsize / config.stepsize_divisor

    # ---- within one iteration
    def record_task_result(self, result):
        self._task_results.append(result)
        self._nb_models_to_evaluate -= 1

    def record_worker_id(self, worker_id):
        self._worker_ids = list(set(self._worker_ids + [worker_id]))

    def record_eval_result(self, result):
        raise NotImplementedError

    def models_left_to_eval(self):
        raise NotImplementedError

    def models_left_to_evolve(self):
        return self._nb_models_to_evaluate > 0

    def warn_waiting_for_evaluations(self):
        if not self.models_left_to_evolve() and self.models_left_to_eval() and not self._waiting_for_eval_run:
            logging.warning('Waiting for evaluations')
            self._waiting_for_eval_run = True

    def set_waiting_for_elite_ev(self, value): self._waiting_for_eval_run = value
    def set_nb_models_to_evaluate(self, n): self._nb_models_to_evaluate = n
    def set_batch_size(self, v): self._batch_size = v
    def set_noise_stdev(self, v): self._noise_stdev = v
    def incr_bad_gens(self): self._bad_generations += 1
    def incr_epoch(self): self._epoch += 1

    # ---- readers (the reference's names)
    def patience_reached(self): return self._patience_reached
    def schedule_reached(self): return self._schedule_reached
    def epoch(self): return self._epoch
    def iteration(self): return self._iteration
    def noise_stdev(self): return self._noise_stdev
    def get_noise_stdev(self): return self._noise_stdev
    def batch_size(self): return self._batch_size
    def times_orig_bs(self): return self._times_orig_bs
    def nb_samples_used(self): return self._nb_samples_used
    def eval_returns(self): return self._eval_results
    def task_results(self): return self._task_results
    def best_elites(self): return self._podium.best_elites()
    def best_elite(self): return self._podium.best_elites()[0][0]

    def schedule_progress(self):
        """(part, full) as the reference logs it under 'Schedule' / 'BadGen' (tools/iteration.py:92-104); None without either."""
        if self._patience:
            return self._bad_generations, self._patience
        if self._schedule_limit:
            if self._iteration <= self._schedule_start:
                return self._iteration, self._schedule_start
            return (self._iteration - self._schedule_start) % self._schedule_limit, self._schedule_limit
        return None


class ESIteration(Iteration):
    """NIC-ES: parents live beside the offspring in models/offspring (0_<i>_parent_params.pth), elite candidates in models/elite
    (0_<i>_elite_params.pth)."""

    def __init__(self, config, exp, store=_FILES):
        super().__init__(config, exp, store)
        self._offspring_dir = os.path.join(self._models_dir, 'offspring')
        self._elite_dir = os.path.join(self._models_dir, 'elite')
        self._pop_size = exp.get('population_size', self._nb_offspring)
        self._num_elite_cands = exp['num_elite_cands']
        self._parents, self._elites_to_evaluate = [], []

    def _parent_path(self, i): return os.path.join(self._offspring_dir, '0_{}_parent_params.pth'.format(i))
    def _elite_path(self, i): return os.path.join(self._elite_dir, '0_{}_elite_params.pth'.format(i))

    def to_dict(self):
        return {**super().to_dict(), 'elites_to_evaluate': self._elites_to_evaluate, 'parents': self._parents}

    def init_from_infos(self, infos):
        super().init_from_infos(infos)
        # files are placed under the id stored with them, the lists are renumbered by position (nic_es/iteration.py:41-48)
        for kind, key in ((self._elite_path, 'elites_to_evaluate'), (self._parent_path, 'parents')):
            self._store.makedirs(os.path.dirname(kind(0)))
            for i, path in infos[key]:
                self._store.copy(path, kind(i))
        self._elites_to_evaluate = [(n, self._elite_path(n)) for n in range(len(infos['elites_to_evaluate']))]
        self._parents = [(n, self._parent_path(n)) for n in range(len(infos['parents']))]

    def init_from_zero(self, exp, policy):
        # parent None = "draw a fresh random model": the first generation is pop_size random individuals
        self._parents = [(i, None) for i in range(self._pop_size)]
        self._store.makedirs(self._elite_dir)
        self._elites_to_evaluate = [(i, policy.generate_model().serialize(path=self._elite_path(i))) for i in range(self._num_elite_cands)]

    def init_from_singles(self, param_file_names, exp, policy):
        names = [param_file_names] if isinstance(param_file_names, str) else list(param_file_names)
        self._store.makedirs(self._elite_dir)
        self._store.makedirs(self._offspring_dir)
        self._parents, self._elites_to_evaluate = [], []
        for i, name in enumerate(names):
            self._parents.append((i, policy.generate_model(from_param_file=name).serialize(path=self._parent_path(i))))
            self._elites_to_evaluate.append((i, policy.generate_model(from_param_file=name).serialize(path=self._elite_path(i))))
        del self._elites_to_evaluate[self._num_elite_cands:]

    def record_parents(self, parents):
        """`parents`: files in the offspring directory, best first.  They are renamed to the parent names, the podium's elites are
        put in front of them, and every other file of the offspring directory is removed."""
        renamed = []
        for i, src in enumerate(parents):
            dst = self._parent_path(i)
            self._store.rename(src, dst)
            renamed.append(dst)
        self._parents = list(enumerate([e for e, _ in self.best_elites()] + renamed))
        self._store.only_keep(self._offspring_dir, [p for _, p in self._parents])

    def set_next_elites_to_evaluate(self, best_individuals):
        self._store.makedirs(self._elite_dir)
        cands = []
        for i, src in enumerate(best_individuals):
            self._store.copy(src, self._elite_path(i))
            cands.append((i, self._elite_path(i)))
        self._store.only_keep(self._elite_dir, [p for _, p in cands])
        self._elites_to_evaluate = cands

    def record_eval_result(self, eval_return):
        prev = self._eval_results.get(eval_return.evaluated_cand_id, ('', None))[1] or float('-inf')
        self._eval_results[eval_return.evaluated_cand_id] = (eval_return.evaluated_cand, max(eval_return.score, prev))

    def models_left_to_eval(self):
        return len(self._eval_results) < len(self._elites_to_evaluate)

    def offspring_dir(self): return self._offspring_dir
    def elite_dir(self): return self._elite_dir
    def parents(self): return self._parents
    def elites_to_evaluate(self): return self._elites_to_evaluate


class NESIteration(Iteration):
    """NIC-NES: one current model, models/current/0_current_params.pth."""

    def __init__(self, config, exp):
        super().__init__(config, exp)
        self._current_dir = os.path.join(self._models_dir, 'current')
        os.makedirs(self._current_dir, exist_ok=True)
        self._current_path = os.path.join(self._current_dir, '0_current_params.pth')
        self._model = None

    def to_dict(self):
        return {**super().to_dict(), 'current_model': self._model}

    def init_from_infos(self, infos):
        super().init_from_infos(infos)
        shutil.copy(infos['current_model'], self._current_path)
        self._model = self._current_path

    def init_from_zero(self, exp, policy):
        self._model = policy.generate_model().serialize(path=self._current_path)

    def init_from_single(self, param_file_name, exp, policy):
        self._model = policy.generate_model(from_param_file=param_file_name).serialize(path=self._current_path)

    def set_model(self, model):
        _FILES.only_keep(self._current_dir, [])
        self._model = model.serialize(path=self._current_path)

    def record_eval_result(self, result):
        prev = self._eval_results.get(0, ('', None))[1] or float('-inf')
        self._eval_results[0] = (self._model, max(result.eval_score, prev))

    def models_left_to_eval(self):
        return not self._eval_results

    def current_model(self): return self._model
    def score(self): return self._eval_results.get(0, ('', None))[1] or float(0)
    def fitnesses(self): return np.stack([r.fitness for r in self._task_results])
    def flat_fitnesses(self): return np.concatenate([r.fitness for r in self._task_results])


class IterationFactory:
    @staticmethod
    def create(config, exp):
        return {'nic_es': ESIteration, 'nic_nes': NESIteration}[exp['algorithm']](config, exp)


class Statistics:
    """The per-iteration lists a snapshot stores (plots and host-memory probes of the reference are not reproduced; their lists
    are carried through a resume unchanged)."""

    _KEYS = ('score_stats', 'score_stds', 'time_stats', 'acc_stats', 'norm_stats', 'noise_std_stats', 'bs_stats', 'mem_stats',
             'update_ratio_stats', 'time_elapsed', 'best_acc_so_far_stats')

    def __init__(self):
        self._d = {k: [] for k in self._KEYS}
        self._d['score_stats'] = [[], [], []]              # min, mean, max
        self._d['mem_stats'] = [[], [], []]
        self._d['time_elapsed'] = 0
        self._step_tstart = 0

    def init_from_infos(self, infos):
        for k in self._KEYS:
            if k in infos:
                self._d[k] = infos[k]

    def to_dict(self):
        return dict(self._d)

    def set_step_tstart(self): self._step_tstart = time.time()

    def record_step_time_stats(self):
        dt = time.time() - self._step_tstart
        self._d['time_elapsed'] += dt
        self._d['time_stats'].append(dt)

    def record_score_stats(self, scores):
        scores = np.asarray(scores)
        for lst, v in zip(self._d['score_stats'], (scores.min(), scores.mean(), scores.max())):
            lst.append(float(v))
        self._d['score_stds'].append(float(scores.std()))

    def record_norm_stats(self, param_vector):
        # mean |theta| (statistics.py:139-141); accepts a torch tensor or an array
        v = param_vector.detach().cpu().numpy() if hasattr(param_vector, 'detach') else np.asarray(param_vector)
        self._d['norm_stats'].append(float(np.abs(v).sum() / v.size))

    def record_time_stats(self, v): self._d['time_stats'].append(v)
    def record_acc_stats(self, v): self._d['acc_stats'].append(v)
    def record_best_acc_stats(self, v): self._d['best_acc_so_far_stats'].append(v)
    def record_std_stats(self, v): self._d['noise_std_stats'].append(v)
    def record_bs_stats(self, v): self._d['bs_stats'].append(v)
    def record_update_ratio(self, v): self._d['update_ratio_stats'].append(v)


def experiment_dict(algorithm, orig_bs, trainloader_lth=0):
    """The experiment's share of a snapshot (tools/experiment.py:45-50)."""
    return {'trainloader_lth': trainloader_lth, 'algorithm': algorithm, 'orig_bs': orig_bs}


def save_infos(directory, stats, it, experiment):
    """Write <directory>/z_info_e<epoch>_i<iter>-<trainloader_lth>.json, replacing any earlier snapshot file there.
    `experiment`: the dict of experiment_dict()."""
    os.makedirs(directory, exist_ok=True)
    for name in os.listdir(directory):
        if re.search(INFOS_PATTERN, name):
            os.remove(os.path.join(directory, name))
    name = 'z_info_e{}_i{}-{}.json'.format(it.epoch(), it.iteration(), experiment['trainloader_lth'])
    with open(os.path.join(directory, name), 'w') as f:
        json.dump({**stats.to_dict(), **it.to_dict(), **experiment}, f)
    return name


def find_infos(directory):
    for name in os.listdir(directory):
        if re.search(INFOS_PATTERN, name):
            return os.path.join(directory, name)
    return None


def load_infos(path):
    with open(path) as f:
        return json.load(f)


def setup_master(exp, policy, log_dir):
    """The bookkeeping half of tools/setup.py:21-45: Config, Iteration and Statistics from an experiment dict, resumed from
    `from_infos`, started from `from_single`, or from zero."""
    exp = dict(exp, log_dir=log_dir)
    os.makedirs(log_dir, exist_ok=True)
    config = Config(**exp['config'])
    it = IterationFactory.create(config, exp)
    stats = Statistics()
    if exp.get('from_infos') is not None:
        infos = load_infos(exp['from_infos'])
        stats.init_from_infos(infos)
        it.init_from_infos(infos)
    elif exp.get('from_single') is not None:
        (it.init_from_singles if exp['algorithm'] == 'nic_es' else it.init_from_single)(exp['from_single'], exp, policy)
    else:
        it.init_from_zero(exp, policy)
    return config, it, stats
