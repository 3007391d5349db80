"""One scripted NIC-ES run -- which models are parents, elite candidates and podium elites after every generation -- written once and
driven twice: tests/golden/make_golden_es.py runs the REAL reference (ESMaster.selection, ESIteration, Podium from /root/reference,
models as files whose content names where they came from) and stores the outcome; tests/test_host_es.py drives
nes_b200.ESPopulation (models as device parent slots; here on a stub context that only tracks where every slot's content came from)
through the same script and compares generation by generation.

The script fixes everything the reference leaves to its workers' unseeded generators and to the data: the fitness of offspring k in
generation g (with ties, so the stable-sort rule matters) and the validation score of every elite candidate.  Content tags:
  'init_elite:j'  candidate j of ESIteration.init_from_zero (nic_es/iteration.py:50-61)
  'g<g>:o<k>'     offspring k of generation g (the file the worker wrote, nic_es_worker.py:188)
"""
import numpy as np

ES_EXP = {'algorithm': 'nic_es', 'nb_offspring': 14, 'population_size': 6, 'num_elites': 3, 'num_elite_cands': 2,
          'selection': 'uniform', 'tournament_size': 0,
          'config': {'noise_stdev': 0.005, 'batch_size': 64, 'schedule_start': 0, 'schedule_limit': 0, 'stdev_divisor': 2.0,
                     'bs_multiplier': 2, 'patience': 2, 'snapshot_freq': 0}}
N_GENERATIONS = 9


def fitnesses(gen):
    """fitness of the nb_offspring offspring of generation gen, in submission order; quantised so that ties occur"""
    rs = np.random.RandomState(1000 + gen)
    return np.round(rs.randn(ES_EXP['nb_offspring']) * 2.0, 0) + 10.0 * (gen % 3)


def elite_scores(gen):
    """validation score of elite candidate j in generation gen (ESWorker.accuracy, nic_es_worker.py:114-141); a candidate that equals a
    podium score exactly and a generation without any improvement are in there on purpose"""
    table = [(0.50, 0.40), (0.45, 0.62), (0.30, 0.62), (0.10, 0.15), (0.05, 0.02), (0.07, 0.01), (0.70, 0.58), (0.01, 0.02), (0.66, 0.71)]
    return table[gen % len(table)]
