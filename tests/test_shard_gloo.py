"""The N > 1 path on CPU: world_size-2 gloo.  Each rank runs ITS shard of a Monte-Carlo batch (the CPU oracle stands in for the
engine, which has no CPU path) with global instance offsets, the statistics blocks are summed with the same all-reduce
bench.py issues over NCCL, and the result equals the unsharded run: bit-exact states (noise keyed by global instance id),
statistics equal up to summation order."""
import os
import socket

import numpy as np
import pytest
import torch.multiprocessing as mp

from tests.common import abi


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, total, NF, steps, q):
    import torch.distributed as dist
    import __graft_entry__ as g
    from oracle import oracle
    from tests import tapes
    pkg = g.load_package()
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    cfg = abi.default_config()
    ev, idp, f, nb = tapes.consistent_tape(oracle, cfg, NF, steps=steps, seed=21, extra_models=False)
    for i in range(len(ev)):
        if ev[i]["kind"] == abi.EV_PROP:
            ev[i]["sigma"][:6] = [0.05, 0.05, 0.05, 0.002, 0.002, 0.002]
    first, count = pkg.shard.shard_range(total, rank, world)
    r = oracle.run_batch(cfg, NF, ev, idp, count=count, nthreads=2, n_buckets=nb, seed=0x5EED, bias_range=[0.01] * 6, inst_off=first,
                         perturb_seed=5)
    summed = pkg.shard.allreduce_stats(r["stats"])
    q.put((rank, first, count, r["x"], summed))
    dist.barrier()
    dist.destroy_process_group()


def test_shard_range_partitions():
    import __graft_entry__ as g
    shard = g.load_package().shard
    for total, world in [(10, 2), (7, 3), (8, 8), (5, 8), (262144, 8)]:
        blocks = [shard.shard_range(total, r, world) for r in range(world)]
        assert blocks[0][0] == 0 and sum(c for _, c in blocks) == total
        for (f0, c0), (f1, _) in zip(blocks, blocks[1:]):
            assert f0 + c0 == f1
        for inst in range(0, total, max(1, total // 50)):
            r = shard.owner_of(inst, total, world)
            assert blocks[r][0] <= inst < blocks[r][0] + blocks[r][1]
    with pytest.raises(ValueError):
        shard.shard_range(4, 2, 2)


@pytest.mark.timeout(300)
def test_two_rank_sharded_batch_equals_unsharded(oracle_mod):
    from tests import tapes
    world, total, NF, steps = 2, 7, 4, 30
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, total, NF, steps, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = sorted([q.get(timeout=240) for _ in range(world)], key=lambda t: t[0])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    cfg = abi.default_config()
    ev, idp, f, nb = tapes.consistent_tape(oracle_mod, cfg, NF, steps=steps, seed=21, extra_models=False)
    for i in range(len(ev)):
        if ev[i]["kind"] == abi.EV_PROP:
            ev[i]["sigma"][:6] = [0.05, 0.05, 0.05, 0.002, 0.002, 0.002]
    full = oracle_mod.run_batch(cfg, NF, ev, idp, count=total, nthreads=2, n_buckets=nb, seed=0x5EED, bias_range=[0.01] * 6, perturb_seed=5)
    assert [g[1] for g in got] == [0, 4] and [g[2] for g in got] == [4, 3]
    assert np.array_equal(np.concatenate([g[3] for g in got]), full["x"])
    for g in got:  # every rank holds the same reduced block
        np.testing.assert_allclose(g[4], full["stats"], rtol=1e-12, atol=1e-12)
    assert full["stats"][:, 0].max() == total
