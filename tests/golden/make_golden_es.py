"""Generate tests/golden/golden_es_v1.json: the REAL reference's NIC-ES master bookkeeping -- ESMaster.selection
(algorithm/nic_es/nic_es_master.py:155-167), ESIteration (algorithm/nic_es/iteration.py), Iteration.process_evaluated_elites
(algorithm/tools/iteration.py:117-148) and Podium (algorithm/tools/podium.py) imported from /root/reference -- driven through
tests/es_scenario.py in the order ESMaster.run_master uses (nic_es_master.py:116-125).  Models are files whose CONTENT says where they
came from; after every generation the content of every parent, elite candidate and podium file is recorded.

Run in the build container only:  python tests/golden/make_golden_es.py
"""
import json
import os
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
sys.path.insert(0, os.path.dirname(HERE))
from oracle import _refimport  # noqa: E402
import es_scenario as sc  # noqa: E402


def content(path):
    with open(path) as f:
        return f.read()


class TaggedModel:
    n = 0

    def serialize(self, path=''):
        with open(path, 'w') as f:
            f.write('init_elite:%d' % TaggedModel.n)
        TaggedModel.n += 1
        return path


class TaggedPolicy:
    def generate_model(self, from_param_file=None, start_rng=None):
        return TaggedModel()


def main():
    _refimport.load()
    from algorithm.nic_es.iteration import ESIteration
    from algorithm.nic_es.nic_es_master import ESMaster, ESResult
    from algorithm.tools.utils import Config, mkdir_p

    out = []
    with tempfile.TemporaryDirectory() as root:
        exp = dict(sc.ES_EXP, log_dir=root)
        config = Config(**exp['config'])
        for d in ('models/offspring', 'models/elite', 'models/parents'):
            mkdir_p(os.path.join(root, d))
        it = ESIteration(config, exp)
        TaggedModel.n = 0
        it.init_from_zero(exp, TaggedPolicy())
        for gen in range(sc.N_GENERATIONS):
            it.incr_iteration()
            parents_seen = it.parents()
            fit = sc.fitnesses(gen)
            for k, f in enumerate(fit):
                # the file a worker writes for its offspring (nic_es_worker.py:184-190); which parent it descends from is the worker's
                # own random draw and does not enter the master's bookkeeping
                path = os.path.join(it.offspring_dir(), '%d_%d_offspring_params.pth' % (k, gen))
                with open(path, 'w') as fh:
                    fh.write('g%d:o%d' % (gen, k))
                it.record_task_result(ESResult(worker_id=k, evaluated_model_id=parents_seen[k % len(parents_seen)][0], evaluated_model=path,
                                               fitness=np.array([f], dtype=float)))
            for (cid, cpath), score in zip(it.elites_to_evaluate(), sc.elite_scores(gen)):
                it.record_eval_result(ESResult(worker_id=0, evaluated_cand_id=cid, evaluated_cand=cpath, score=score))
            best_sc, best_ind = it.process_evaluated_elites()
            best_content = content(best_ind)
            parents, scores = ESMaster.selection(it.task_results(), exp['population_size'], exp['num_elites'])
            it.set_next_elites_to_evaluate(parents[:exp['num_elite_cands']])
            it.record_parents(parents)
            out.append({
                'gen': gen,
                'sorted_scores': [float(s) for s in scores],
                'best_evaluated': [best_content, float(best_sc)],
                'parents': [content(p) for _, p in it.parents()],
                'elites_to_evaluate': [content(p) for _, p in it.elites_to_evaluate()],
                'podium': [[content(p), float(s)] for p, s in it.best_elites()],
                'noise_stdev': float(it.noise_stdev()), 'batch_size': int(it.batch_size()),
                'bad_generations': int(it.to_dict()['bad_generations']), 'patience_reached': bool(it.patience_reached()),
            })
            print(out[-1])
    path = os.path.join(HERE, 'golden_es_v1.json')
    with open(path, 'w') as f:
        json.dump({'exp': sc.ES_EXP, 'generations': out}, f, indent=1)
    print('wrote', path)


if __name__ == '__main__':
    main()
