"""Population sharding over the GPUs of one box.

The reference farms individuals out one per AMQP message (gentun/master.py:116-123) and collects
`[i, fitness]` replies (master.py:126-129).  Here every rank (one process per GPU, torch.distributed)
holds a replica of the binned data, evaluates a slice of the individuals, and the only collective is
an all-gather of the fitness scalars.  Every rank runs the same seeded GA, so no genes are broadcast.
"""
import numpy as np


def _dist():
    try:
        import torch.distributed as dist
    except Exception:  # torch not importable: single-process use
        return None
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        return dist
    return None


def fit_cost(genes):
    """Relative device time of one individual's CV fit.  Measured on a B200 with tools/cost_model.py (random
    batches of the seeded 1M x 50 population at 100 rounds, least squares, 4 % rms error per batch of 16): the
    time is set by the tree depth -- a term per level and a term per node of the deepest levels -- hardly at all
    by the sampling genes (the levels with many small nodes are bound by the split scan, not by the rows), and
    it falls with the learning rate: a large eta shrinks the residuals quickly, later trees find fewer valid
    splits and stay smaller."""
    d = float(genes["max_depth"])
    e = min(float(genes.get("eta", 0.3)), 0.2)
    return 11.5 + 8.1 * d + 61.1 * 2.0 ** (d - 10.0) - e * (159.5 * 2.0 ** (d - 10.0) + 10.5 * d)


def shard_plan(costs, world_size):
    """Deal individuals to ranks: most expensive first, each to the rank with the least work so far among the
    ranks that still have room (every rank takes at most ceil(n / world_size) individuals, so the waves of
    per-individual device state stay the same size everywhere).  `costs` is any per-individual weight
    (fit_cost, or plain max_depth).  Deterministic: ties go to the earlier individual and the lower rank.
    Returns a list (per rank) of index arrays into the population."""
    costs = np.asarray(costs, dtype=np.float64)
    order = np.argsort(-costs, kind="stable")
    room = -(-len(costs) // world_size)
    load = np.zeros(world_size)
    picks = [[] for _ in range(world_size)]
    for i in order:
        open_ranks = [r for r in range(world_size) if len(picks[r]) < room]
        r = min(open_ranks, key=lambda q: (load[q], q))
        picks[r].append(int(i))
        load[r] += costs[i]
    return [np.asarray(p, dtype=np.int64) for p in picks]


def _raise_for_bad(values, where):
    bad = [i for i, v in enumerate(values) if not np.isfinite(v)]
    if bad:
        raise RuntimeError("gentun_b200: individual(s) %s produced a non-finite fitness (%s); the reference would hand "
                           "nan/inf to max()/min(), whose result then depends on the order of the population" % (bad, where))


def evaluate_sharded(engine, genes_list, params, local_eval=None):
    """Fitness vector for genes_list, identical on every rank.  A rank whose evaluation fails still takes part in the
    all-gather and reports the failure there, so every rank raises together instead of the others waiting for ever."""
    n = len(genes_list)
    if local_eval is None:
        def local_eval(sub):
            return engine.evaluate(sub, **params)[0]
    dist = _dist()
    if dist is None:
        out = np.asarray(local_eval(list(genes_list)), dtype=np.float64)
        _raise_for_bad(out, "single process")
        return out
    import torch
    world, rank = dist.get_world_size(), dist.get_rank()
    plan = shard_plan([fit_cost(g) for g in genes_list], world)
    mine = plan[rank]
    failure = None
    local = np.zeros(0)
    if len(mine):
        try:
            local = np.asarray(local_eval([genes_list[i] for i in mine]), dtype=np.float64)
        except Exception as exc:   # reported to every rank below
            failure = exc
            local = np.zeros(len(mine))
    width = max(len(p) for p in plan)
    backend = dist.get_backend()
    dev = torch.device("cuda", torch.cuda.current_device()) if backend == "nccl" else torch.device("cpu")
    # row 0: fitness values; row 1: 1.0 where row 0 holds a value this rank evaluated; last column of row 1: failure flag
    send = torch.zeros((2, width + 1), dtype=torch.float64, device=dev)
    if len(mine):
        send[0, :len(mine)] = torch.from_numpy(local).to(dev)
        send[1, :len(mine)] = 1.0
    send[1, width] = 1.0 if failure is not None else 0.0
    recv = torch.empty((world * 2 * (width + 1),), dtype=torch.float64, device=dev)
    dist.all_gather_into_tensor(recv, send.reshape(-1))
    table = recv.cpu().numpy().reshape(world, 2, width + 1)
    failed = [r for r in range(world) if table[r, 1, width] != 0.0]
    if failed:
        if failure is not None:
            raise failure
        raise RuntimeError("gentun_b200: fitness evaluation failed on rank(s) %s" % failed)
    out = np.empty(n)
    done = np.zeros(n, dtype=bool)
    for r in range(world):
        k = len(plan[r])
        out[plan[r]] = table[r, 0, :k]
        done[plan[r]] = table[r, 1, :k] == 1.0
    if not done.all():
        raise RuntimeError("gentun_b200: fitness all-gather left %d individuals unevaluated" % int((~done).sum()))
    _raise_for_bad(out, "gathered from %d ranks" % world)
    return out


def init_distributed(backend="nccl"):
    """Convenience for scripts launched with torchrun (one process per GPU): binds this rank to its GPU and
    joins the process group.  After this, Population.get_fittest() shards each generation's pending individuals
    over the ranks; run the same seeded script on every rank."""
    import os
    import torch
    import torch.distributed as dist
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if backend == "nccl":
        torch.cuda.set_device(local_rank)
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    if not dist.is_initialized():
        kwargs = {"device_id": torch.device("cuda", local_rank)} if backend == "nccl" else {}
        dist.init_process_group(backend, **kwargs)
    return dist.get_rank(), dist.get_world_size()
