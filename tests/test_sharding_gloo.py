"""Population sharding + fitness all-gather with two CPU ranks over gloo (the N > 1 path's host logic)."""
import os
import socket
import subprocess
import sys
import textwrap

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = textwrap.dedent("""
    import json, os, sys
    sys.path.insert(0, %r)
    import numpy as np
    import torch.distributed as dist
    from gentun_b200 import sharding
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%%s" %% os.environ["PORT"],
                            rank=int(os.environ["RANK"]), world_size=int(os.environ["WORLD_SIZE"]))
    genes = [{"max_depth": 3 + (i * 7) %% 8, "eta": 0.01 * (i + 1)} for i in range(11)]
    seen = []
    def local_eval(sub):
        seen.extend(g["eta"] for g in sub)
        return np.array([g["max_depth"] * 10 + g["eta"] for g in sub])
    out = sharding.evaluate_sharded(None, genes, {}, local_eval=local_eval)
    print(json.dumps({"rank": dist.get_rank(), "out": out.tolist(), "mine": len(seen)}))
    dist.destroy_process_group()
""") % ROOT


def free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def test_two_rank_allgather_reassembles_the_population():
    import json
    port = str(free_port())
    procs = []
    for rank in range(2):
        env = dict(os.environ, RANK=str(rank), WORLD_SIZE="2", PORT=port, MASTER_ADDR="127.0.0.1")
        procs.append(subprocess.Popen([sys.executable, "-c", WORKER], env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True))
    outs = []
    for p in procs:
        so, se = p.communicate(timeout=180)
        assert p.returncode == 0, se[-2000:]
        outs.append(json.loads(so.strip().splitlines()[-1]))
    want = [(3 + (i * 7) % 8) * 10 + 0.01 * (i + 1) for i in range(11)]
    for o in outs:
        np.testing.assert_allclose(o["out"], want)
    assert sorted(o["mine"] for o in outs) == [5, 6]      # every individual trained on exactly one rank


FAILING_WORKER = WORKER.replace(
    'return np.array([g["max_depth"] * 10 + g["eta"] for g in sub])',
    'if dist.get_rank() == 1: raise ValueError("boom on rank 1")\n    return np.array([g["max_depth"] * 10 + g["eta"] for g in sub])'
).replace('out = sharding.evaluate_sharded(None, genes, {}, local_eval=local_eval)',
          'try:\n    out = sharding.evaluate_sharded(None, genes, {}, local_eval=local_eval)\n'
          'except Exception as exc:\n    print(json.dumps({"rank": dist.get_rank(), "error": type(exc).__name__ + ": " + str(exc)}))\n'
          '    dist.destroy_process_group()\n    sys.exit(0)')


def test_a_failing_rank_makes_every_rank_raise():
    """One rank's evaluation throws: the others must not wait in the all-gather for ever; all of them raise."""
    import json
    port = str(free_port())
    procs = []
    for rank in range(2):
        env = dict(os.environ, RANK=str(rank), WORLD_SIZE="2", PORT=port, MASTER_ADDR="127.0.0.1")
        procs.append(subprocess.Popen([sys.executable, "-c", FAILING_WORKER], env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True))
    errs = {}
    for p in procs:
        so, se = p.communicate(timeout=180)
        assert p.returncode == 0, se[-2000:]
        o = json.loads(so.strip().splitlines()[-1])
        errs[o["rank"]] = o.get("error")
    assert errs[1] == "ValueError: boom on rank 1"
    assert errs[0] is not None and "rank(s) [1]" in errs[0]


def test_non_finite_fitness_raises_after_the_batch():
    from gentun_b200 import sharding
    import pytest

    class Eng(object):
        def evaluate(self, genes, **params):
            return np.array([1.0, float("nan"), 3.0]), None

    with pytest.raises(RuntimeError, match=r"individual\(s\) \[1\]"):
        sharding.evaluate_sharded(Eng(), [{"max_depth": 3}] * 3, {})


def test_shard_plan_balances_depths():
    from gentun_b200.sharding import shard_plan
    depths = [10, 3, 10, 4, 9, 3, 8, 5]
    plan = shard_plan(depths, 2)
    assert sorted(np.concatenate(plan).tolist()) == list(range(8))
    sums = [sum(depths[i] for i in p) for p in plan]
    assert abs(sums[0] - sums[1]) <= 2


def test_shard_plan_by_fit_cost_is_balanced_and_deterministic():
    from gentun_b200.sharding import fit_cost, shard_plan
    from tests.synth import seeded_population_genes
    genes = seeded_population_genes(256)
    costs = [fit_cost(g) for g in genes]
    assert min(costs) > 0
    for world in (2, 4, 8):
        plan = shard_plan(costs, world)
        assert sorted(np.concatenate(plan).tolist()) == list(range(256))
        assert {len(p) for p in plan} == {256 // world}                      # equal counts
        loads = [sum(costs[i] for i in p) for p in plan]
        assert max(loads) / (sum(loads) / world) < 1.01                      # within 1 % under the model
        again = shard_plan(costs, world)
        assert all(np.array_equal(a, b) for a, b in zip(plan, again))
    # deeper trees and smaller learning rates cost more
    base = dict(genes[0])
    assert fit_cost(dict(base, max_depth=10)) > fit_cost(dict(base, max_depth=4))
    assert fit_cost(dict(base, max_depth=10, eta=0.01)) > fit_cost(dict(base, max_depth=10, eta=0.5))


def test_single_process_path_calls_the_engine_once():
    from gentun_b200 import sharding

    class Eng(object):
        calls = 0

        def evaluate(self, genes, **params):
            Eng.calls += 1
            return np.arange(len(genes), dtype=np.float64), None

    out = sharding.evaluate_sharded(Eng(), [{"max_depth": 3}] * 4, {})
    assert out.tolist() == [0, 1, 2, 3] and Eng.calls == 1
