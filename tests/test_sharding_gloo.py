"""CPU test of the N>1 host logic with torch.distributed (gloo, world_size 2): the hash shards partition the
conversations exactly as ssegpu_shard routes them, every rank's stream contains only its own conversations, and the
counter gather sums over ranks.  The per-rank 'counters' come from the CPU oracle here (there is no GPU on the build
box); on a GPU box the same gather runs on the library's counters (bench.py --gpus N)."""
import os
import sys

import numpy as np
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, n_conv, n_tok, q):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch.distributed as dist
    from __graft_entry__ import load_package
    import oracle_lib
    pkg = load_package()
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
    try:
        arena, off = pkg.sharding.rank_workload(1, n_conv, n_tok, rank, world)
        members = pkg.workload.shard_members(1, n_conv, rank, world)
        n = len(off) - 1
        assert n == len(members) * n_tok
        # every record of this rank's stream belongs to a conversation that hashes to this rank
        ids = [pkg.workload.conversation_id(1, int(i)) for i in members]
        assert all(pkg.shard(c, world) == rank for c in ids)
        for r in (0, n // 2, n - 1):
            rec = arena[off[r]:off[r + 1]].tobytes()
            assert ids[r % len(members)] in rec
        oracle = oracle_lib.load()
        out, out_off, status = oracle.adapter(arena, off)
        local = {k: 0 for k in pkg.sharding.COUNTER_KEYS}
        local.update(records=n, frames=int((status & 0xF == 0).sum()), bytes_in=int(off[n]), bytes_out=int(out_off[n]),
                     done=int((status & 0x100 != 0).sum()), batches=1)
        total = pkg.sharding.gather_counters(local, dist)
        q.put((rank, len(members), local, total))
    finally:
        dist.destroy_process_group()


def test_hash_shards_partition_and_counter_gather():
    world, n_conv, n_tok = 2, 3000, 7
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29400 + os.getpid() % 500
    procs = [ctx.Process(target=_worker, args=(r, world, port, n_conv, n_tok, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert sum(r[1] for r in res) == n_conv                       # shards cover every conversation exactly once
    assert abs(res[0][1] - res[1][1]) < 0.1 * n_conv              # and are balanced
    want = {k: sum(r[2][k] for r in res) for k in res[0][2]}
    for r in res:
        assert r[3] == want                                       # every rank sees the same global totals
    assert want["records"] == n_conv * n_tok == want["frames"] and want["done"] == n_conv


def test_route_matches_shard_members():
    sys.path.insert(0, ROOT)
    from __graft_entry__ import load_package
    pkg = load_package()
    ids = [pkg.workload.conversation_id(1, i) for i in range(500)]
    for world in (1, 2, 4, 8):
        owner = pkg.sharding.route(ids, world)
        for r in range(world):
            assert np.array_equal(np.nonzero(owner == r)[0].astype(np.uint32), pkg.workload.shard_members(1, 500, r, world))
