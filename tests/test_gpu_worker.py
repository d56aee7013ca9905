"""The RabbitMQ worker protocol with the B200 engine behind it: a DistributedPopulation master that holds no
data ships `[i, genes, additional_parameters]` jobs (gentun/master.py:116-123), a GentunWorker with the data pulls them in a batch,
trains the batch in one engine call on the GPU and replies `[i, fitness]` (gentun/worker.py:37-55).  Transport: the in-process stand-in
for pika (tests/fake_pika.py); no broker exists on the box."""
import contextlib
import io
import random
import sys
import threading

import numpy as np
import pytest

from oracle import orc
from tests import fake_pika

pytestmark = pytest.mark.gpu


def test_worker_trains_a_queued_batch_on_the_gpu(wine, monkeypatch):
    monkeypatch.setitem(sys.modules, "pika", fake_pika)
    fake_pika.BROKER.reset()
    import gentun_b200 as gentun
    saved = dict(gentun.DistributedPopulation.default_credentials)
    x, y = wine
    params = {"nfold": 3, "num_boost_round": 8}
    random.seed(21)
    try:
        with contextlib.redirect_stdout(io.StringIO()):
            pop = gentun.DistributedPopulation(gentun.XgboostIndividual, size=6, maximize=False, additional_parameters=params,
                                               rabbit_queue="gpu_jobs")
            worker = gentun.GentunWorker(gentun.XgboostIndividual, x, y, rabbit_queue="gpu_jobs", batch_size=8, batch_wait=0.3)
            t = threading.Thread(target=lambda: worker.work(max_batches=1), daemon=True)
            t.start()
            best = pop.get_fittest()
            t.join(timeout=60)
        assert not t.is_alive()
        ctx = orc.Context(np.asarray(x.values, dtype=np.float32), np.asarray(y.values, dtype=np.float32), 3)
        for i in range(6):
            want = ctx.cv(pop[i].get_genes(), orc.make_opts(**params))["fitness"]
            assert abs(pop[i].get_fitness() - want) <= 1e-4 * abs(want)          # north_star: 1e-4 relative
        ctx.close()
        assert best.get_fitness() == min(pop[i].get_fitness() for i in range(6))
    finally:
        gentun.DistributedPopulation.default_credentials = saved
        gentun.clear_engines()
