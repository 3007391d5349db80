"""One scripted run of the iteration bookkeeping (schedule, patience, podium, parents / elites, snapshot, resume), written against
the method names the reference's `Iteration` classes and this package's `schedule` module share.  tests/golden/make_golden_schedule.py
drives the REAL reference through it and stores every `to_dict()` on the way; tests/test_host_schedule.py drives this package through it
and compares.  File contents are irrelevant here (small marker files); paths are stored relative to the scenario root as '<ROOT>/...'."""
import json
import os
from collections import namedtuple

EvalES = namedtuple('EvalES', ['evaluated_cand_id', 'evaluated_cand', 'score'])
EvalNES = namedtuple('EvalNES', ['eval_score'])

NES_EXP = {'algorithm': 'nic_nes', 'nb_offspring': 8, 'num_elites': 2,
           'config': {'noise_stdev': 0.01, 'batch_size': 64, 'schedule_start': 3, 'schedule_limit': 4, 'stdev_divisor': 1.5,
                      'bs_multiplier': 1.5, 'patience': 0, 'stepsize_divisor': 2.0, 'snapshot_freq': 5, 'l2coeff': 0.002}}
NES_SCORES = [0.31, 0.29, 0.35, 0.35, 0.40, 0.12, 0.41, 0.39, 0.44, 0.44, 0.10, 0.50]

ES_EXP = {'algorithm': 'nic_es', 'nb_offspring': 6, 'population_size': 4, 'num_elites': 3, 'num_elite_cands': 2,
          'config': {'noise_stdev': 0.005, 'batch_size': 100, 'schedule_start': 0, 'schedule_limit': 0, 'stdev_divisor': 2.0,
                     'bs_multiplier': 2, 'patience': 2, 'snapshot_freq': 3}}
# per generation: the scores of the two elite candidates (second one evaluated twice in generation 1: the max counts)
ES_SCORES = [(0.50, 0.40), (0.45, 0.48), (0.30, 0.20), (0.10, 0.15), (0.05, 0.02), (0.60, 0.01), (0.55, 0.58), (0.01, 0.02),
             (0.02, 0.01), (0.03, 0.02), (0.70, 0.04)]


def rel(obj, root):
    """paths -> '<ROOT>/...' throughout a JSON-like structure (tuples become lists, as json would store them)"""
    if isinstance(obj, str):
        return obj.replace(root, '<ROOT>')
    if isinstance(obj, (list, tuple)):
        return [rel(o, root) for o in obj]
    if isinstance(obj, dict):
        return {k: rel(v, root) for k, v in obj.items()}
    if obj is None or isinstance(obj, (bool, int)):
        return obj
    return float(obj)


def listing(root):
    out = []
    for d, _, files in os.walk(root):
        out += [os.path.join(d, f).replace(root, '<ROOT>') for f in files]
    return sorted(out)


class FakeModel:
    """stands in for a PolicyNet: serialize(path) writes a marker file and returns the path (algorithm/nets.py:153-155)"""
    count = 0

    def serialize(self, path=''):
        FakeModel.count += 1
        with open(path, 'w') as f:
            f.write('model %d\n' % FakeModel.count)
        return path


class FakePolicy:
    def generate_model(self, from_param_file=None, start_rng=None):
        return FakeModel()


def state(it, root):
    d = rel(it.to_dict(), root)
    d['_patience_reached'] = bool(it.patience_reached())
    d['_schedule_reached'] = bool(it.schedule_reached())
    return d


def run_nes(ns, root, make_model=FakeModel, trace=None):
    """ns: namespace with Config, NESIteration, Statistics, save(directory, stats, it, lth) -> file name, load(path).
    Returns the list of states."""
    trace = [] if trace is None else trace
    log_dir = os.path.join(root, 'nes_a')
    os.makedirs(os.path.join(log_dir, 'models'), exist_ok=True)
    exp = dict(NES_EXP, log_dir=log_dir)
    config = ns.Config(**exp['config'])
    it = ns.NESIteration(config, exp)
    stats = ns.Statistics()
    it.init_from_zero(exp, FakePolicy())
    trace.append(('init', state(it, root)))
    stepsize = 0.01
    snap = None
    for n, score in enumerate(NES_SCORES):
        it.incr_iteration()
        if n % 3 == 2:
            it.incr_epoch()
        it.record_eval_result(EvalNES(eval_score=score))
        if n == 4:
            it.record_eval_result(EvalNES(eval_score=score - 0.2))      # a second, worse evaluation: the max is kept
        it.process_evaluated_elites()
        it.set_model(make_model())
        if it.patience_reached() or it.schedule_reached():
            stepsize /= config.stepsize_divisor
        stats.record_bs_stats(it.batch_size())
        stats.record_std_stats(it.noise_stdev())
        stats.record_acc_stats(it.score())
        stats.record_best_acc_stats(it.best_elites()[0][1])
        stats.record_update_ratio(0.001 * (n + 1))
        s = state(it, root)
        s['_stepsize'] = stepsize
        s['_score'] = it.score()
        trace.append(('it%d' % n, s))
        if config.snapshot_freq != 0 and it.iteration() % config.snapshot_freq == 0:
            snap = ns.save(os.path.join(log_dir, 'snapshot'), stats, it, 313, 'nic_nes', 64)
            trace.append(('snap%d' % n, snap))
    # resume from the last snapshot (iteration 10) in a fresh log dir and run two more iterations
    infos_path = os.path.join(log_dir, 'snapshot', snap)
    with open(infos_path) as f:
        trace.append(('infos', rel(json.load(f), root)))
    log_dir_b = os.path.join(root, 'nes_b')
    os.makedirs(os.path.join(log_dir_b, 'models'), exist_ok=True)
    exp_b = dict(NES_EXP, log_dir=log_dir_b)
    it_b = ns.NESIteration(config, exp_b)
    stats_b = ns.Statistics()
    infos = ns.load(infos_path)
    stats_b.init_from_infos(infos)
    it_b.init_from_infos(infos)
    trace.append(('resumed', state(it_b, root)))
    for n, score in enumerate([0.47, 0.52]):
        it_b.incr_iteration()
        it_b.record_eval_result(EvalNES(eval_score=score))
        it_b.process_evaluated_elites()
        trace.append(('res_it%d' % n, state(it_b, root)))
    trace.append(('stats_b', rel(stats_b.to_dict(), root)))
    trace.append(('files', listing(root)))
    return trace


def run_es(ns, root, trace=None):
    trace = [] if trace is None else trace
    log_dir = os.path.join(root, 'es_a')
    exp = dict(ES_EXP, log_dir=log_dir)
    for sub in ('offspring', 'elite'):
        os.makedirs(os.path.join(log_dir, 'models', sub), exist_ok=True)
    config = ns.Config(**exp['config'])
    it = ns.ESIteration(config, exp)
    stats = ns.Statistics()
    it.init_from_zero(exp, FakePolicy())
    trace.append(('init', state(it, root)))
    snap = None
    for g, scores in enumerate(ES_SCORES):
        it.incr_iteration()
        # elite candidates of this generation get evaluated
        cands = it.elites_to_evaluate()
        for (cid, path), sc in zip(cands, scores):
            it.record_eval_result(EvalES(evaluated_cand_id=cid, evaluated_cand=path, score=sc))
        if g == 1:
            it.record_eval_result(EvalES(evaluated_cand_id=cands[1][0], evaluated_cand=cands[1][1], score=scores[1] - 0.3))
        trace.append(('g%d_left' % g, bool(it.models_left_to_eval())))
        # offspring files of this generation; the best 4 become parents, the best 2 the next elite candidates
        offspring = []
        for k in range(exp['nb_offspring']):
            p = os.path.join(it.offspring_dir(), '%d_%d_offspring_params.pth' % (g + 1, k))
            FakeModel().serialize(path=p)
            offspring.append(p)
        it.process_evaluated_elites()
        it.set_next_elites_to_evaluate(offspring[:2])
        it.record_parents(offspring[:4])
        stats.record_bs_stats(it.batch_size())
        stats.record_std_stats(it.noise_stdev())
        trace.append(('g%d' % g, state(it, root)))
        if config.snapshot_freq != 0 and it.iteration() % config.snapshot_freq == 0:
            snap = ns.save(os.path.join(log_dir, 'snapshot'), stats, it, 0, 'nic_es', 100)
            trace.append(('snap%d' % g, snap))
    trace.append(('files_a', listing(log_dir)))
    infos_path = os.path.join(log_dir, 'snapshot', snap)
    with open(infos_path) as f:
        trace.append(('infos', rel(json.load(f), root)))
    log_dir_b = os.path.join(root, 'es_b')
    for sub in ('offspring', 'elite'):
        os.makedirs(os.path.join(log_dir_b, 'models', sub), exist_ok=True)
    it_b = ns.ESIteration(config, dict(ES_EXP, log_dir=log_dir_b))
    infos = ns.load(infos_path)
    it_b.init_from_infos(infos)
    trace.append(('resumed', state(it_b, root)))
    it_b.incr_iteration()
    for (cid, path), sc in zip(it_b.elites_to_evaluate(), (0.9, 0.65)):
        it_b.record_eval_result(EvalES(evaluated_cand_id=cid, evaluated_cand=path, score=sc))
    it_b.process_evaluated_elites()
    trace.append(('res_g0', state(it_b, root)))
    trace.append(('files_b', listing(log_dir_b)))
    return trace
