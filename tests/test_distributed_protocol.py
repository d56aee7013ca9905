"""Master / worker transport against an in-process fake broker: wire format of the reference
(gentun/master.py:118, gentun/worker.py:38-50), batching on the worker side, credentials surviving the GA's
re-instantiation of the population class."""
import contextlib
import io
import json
import math
import random
import sys
import threading

import pytest

from tests import fake_pika
from gentun_b200.genetics import XgboostIndividual


def stub_fitness(genes):
    return sum(math.sin(float(v) * (i + 1)) for i, (_, v) in enumerate(sorted(genes.items())))


class StubIndividual(XgboostIndividual):
    def evaluate_fitness(self):
        self.fitness = stub_fitness(self.genes)


@pytest.fixture()
def broker(monkeypatch):
    monkeypatch.setitem(sys.modules, "pika", fake_pika)
    fake_pika.BROKER.reset()
    import gentun_b200 as gentun
    saved = dict(gentun.DistributedPopulation.default_credentials)
    yield fake_pika.BROKER
    gentun.DistributedPopulation.default_credentials = saved


def start_worker(gentun, batches, **kw):
    worker = gentun.GentunWorker(StubIndividual, "x", "y", user="wk", password="pw", **kw)
    t = threading.Thread(target=lambda: worker.work(max_batches=batches), daemon=True)
    t.start()
    return worker, t


def test_master_worker_round_trip_and_wire_format(broker):
    import gentun_b200 as gentun
    random.seed(5)
    with contextlib.redirect_stdout(io.StringIO()):
        pop = gentun.DistributedPopulation(StubIndividual, size=7, maximize=False, additional_parameters={"nfold": 3},
                                           host="broker.example", user="alice", password="s3cret", rabbit_queue="jobs")
        worker, thread = start_worker(gentun, batches=1, rabbit_queue="jobs", batch_size=16, batch_wait=0.2)
        best = pop.get_fittest()
        thread.join(timeout=10)
    assert not thread.is_alive()
    assert all(pop[i].get_fitness_status() for i in range(7))
    assert [pop[i].get_fitness() for i in range(7)] == [stub_fitness(pop[i].get_genes()) for i in range(7)]
    assert best.get_fitness() == min(pop[i].get_fitness() for i in range(7))
    jobs = [json.loads(body) for q, body, user in broker.log if q == "jobs"]
    assert [j[0] for j in jobs] == list(range(7))
    assert jobs[0][1] == pop[0].get_genes()
    assert jobs[0][2] == {"booster": "gbtree", "objective": "reg:linear", "eval_metric": "rmse", "nfold": 3,
                          "num_boost_round": 5000, "early_stopping_rounds": 100}
    replies = [json.loads(body) for q, body, user in broker.log if q != "jobs"]
    assert sorted(r[0] for r in replies) == list(range(7)) and all(isinstance(r[1], float) for r in replies)
    assert {user for q, body, user in broker.log if q == "jobs"} == {"alice"}


def test_credentials_survive_the_genetic_algorithm(broker):
    import gentun_b200 as gentun
    random.seed(6)
    with contextlib.redirect_stdout(io.StringIO()):
        pop = gentun.DistributedPopulation(StubIndividual, size=6, maximize=True, user="bob", password="pw", rabbit_queue="q2")
        worker, thread = start_worker(gentun, batches=2, rabbit_queue="q2", batch_size=16, batch_wait=0.2)
        ga = gentun.GeneticAlgorithm(pop, tournament_size=3)
        ga.run(2)          # generation 2's population was created by the GA without any credentials
        thread.join(timeout=10)
    assert type(ga.population) is gentun.DistributedPopulation
    assert ga.population.credentials["user"] == "bob" and ga.population.credentials["rabbit_queue"] == "q2"
    assert {user for q, body, user in broker.log if q == "q2"} == {"bob"}


def test_worker_batches_jobs_and_handles_tuples(broker):
    import gentun_b200 as gentun
    seen = []

    class Batched(XgboostIndividual):
        @classmethod
        def evaluate_batch(cls, individuals):
            seen.append(len(individuals))
            for ind in individuals:
                ind.set_fitness(stub_fitness(ind.get_genes()))

    random.seed(7)
    with contextlib.redirect_stdout(io.StringIO()):
        pop = gentun.DistributedPopulation(Batched, size=5, rabbit_queue="q3")
        worker = gentun.GentunWorker(Batched, "x", "y", rabbit_queue="q3", batch_size=3, batch_wait=0.2)
        t = threading.Thread(target=lambda: worker.work(max_batches=2), daemon=True)
        t.start()
        pop.get_fittest()
        t.join(timeout=10)
    assert seen == [3, 2]
    i, ind = gentun.GentunWorker._rebuild(StubIndividual, None, None, json.dumps([4, pop[0].get_genes(), {"nfold": 2}]))
    assert i == 4 and ind.nfold == 2


def test_worker_serves_heartbeats_while_a_batch_trains(broker):
    """A long batch must not starve the connection: the worker keeps calling process_data_events while the engine
    call runs (the reference's worker has a heartbeat thread, worker.py:26-33), and a failure in the training thread
    surfaces on the worker's own thread."""
    import time
    import gentun_b200 as gentun

    class Slow(XgboostIndividual):
        @classmethod
        def evaluate_batch(cls, individuals):
            time.sleep(0.35)
            for ind in individuals:
                ind.set_fitness(stub_fitness(ind.get_genes()))

    random.seed(9)
    with contextlib.redirect_stdout(io.StringIO()):
        pop = gentun.DistributedPopulation(Slow, size=3, rabbit_queue="q5")
        worker = gentun.GentunWorker(Slow, "x", "y", rabbit_queue="q5", batch_size=3, batch_wait=0.2, heartbeat_every=0.05)
        t = threading.Thread(target=lambda: worker.work(max_batches=1), daemon=True)
        t.start()
        pop.get_fittest()
        t.join(timeout=10)
    assert worker.heartbeats_served >= 3

    class Broken(XgboostIndividual):
        @classmethod
        def evaluate_batch(cls, individuals):
            raise RuntimeError("engine failed")

    with contextlib.redirect_stdout(io.StringIO()):
        w2 = gentun.GentunWorker(Broken, "x", "y", rabbit_queue="q6", batch_size=2, batch_wait=0.1, heartbeat_every=0.05)
        with pytest.raises(RuntimeError, match="engine failed"):
            w2._evaluate([Broken(None, None)])


def test_distributed_grid_population(broker):
    import gentun_b200 as gentun
    random.seed(8)
    with contextlib.redirect_stdout(io.StringIO()):
        pop = gentun.DistributedGridPopulation(StubIndividual, genes_grid={"eta": [0.1, 0.2], "max_depth": [3, 4, 5]},
                                               additional_parameters={"nfold": 3}, maximize=False, rabbit_queue="q4")
        worker, thread = start_worker(gentun, batches=1, rabbit_queue="q4", batch_size=16, batch_wait=0.2)
        pop.get_fittest()
        thread.join(timeout=10)
    assert pop.get_size() == 6 and all(pop[i].get_fitness_status() for i in range(6))
