"""Multi-GPU host logic on CPU: sequence dealing, tuple ranges, and the world_size-2 gather of
final scores / handler shards over gloo (the compute leg is played by the oracle here; on GPUs
it is the CUDA path -- see bench.py)."""
import os

import numpy as np
import pytest


def test_shard_sequences_partition_and_balance():
    from spec_viterbi_b200 import sharding
    rng = np.random.default_rng(0)
    lens = rng.integers(1, 8000, 333)
    for world in (1, 2, 3, 8):
        parts = sharding.shard_sequences(lens, world)
        allidx = np.sort(np.concatenate(parts))
        assert np.array_equal(allidx, np.arange(lens.size))
        loads = np.array([lens[p].sum() for p in parts])
        assert loads.max() - loads.min() <= lens.max()
    covid = [38, 805, 275, 419, 61, 121, 43, 121, 97, 7096, 4405, 1273, 492, 75, 222, 73]
    parts = sharding.shard_sequences(covid, 2)
    assert 9 in parts[0] and 9 not in parts[1]


def test_shard_tuples():
    from spec_viterbi_b200 import sharding
    for n, w in ((400, 8), (400, 3), (20, 8), (5, 8)):
        ranges = [sharding.shard_tuples(n, w, r) for r in range(w)]
        assert ranges[0][0] == 0 and ranges[-1][1] == n
        assert all(ranges[i][1] == ranges[i + 1][0] for i in range(w - 1))
        sizes = [e - b for b, e in ranges]
        assert max(sizes) - min(sizes) <= 1


def _worker(rank, world, port, q):
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from oracle import pyoracle
        from spec_viterbi_b200 import generators, sharding
        hmm = generators.random_hmm(states_num=60, seed=3)
        rng = np.random.default_rng(5)
        lens = rng.integers(1, 90, 23)
        seqs = [rng.integers(0, 20, int(L)).astype(np.uint64) for L in lens]
        offs = np.zeros(len(seqs) + 1, np.uint64)
        offs[1:] = np.cumsum(lens)
        syms = np.concatenate(seqs)
        mine = sharding.shard_sequences(lens, world)[rank]
        s_loc, o_loc = sharding.take_sequences(syms, offs, mine)
        oracle = pyoracle.Oracle()
        out = oracle.plain_batch(hmm, s_loc, o_loc)
        best = torch.from_numpy(out.min(1))
        arg = torch.from_numpy(out.argmin(1).astype(np.int32))
        gbest, garg = sharding.gather_best(best, arg, mine, len(seqs))
        # the sync-free gather bench.py runs inside its timed region (strong scaling): same result, twice in a row
        bg = sharding.BestGather(sharding.shard_sequences(lens, world), len(seqs), torch.device("cpu"))
        for _ in range(2):
            b2, a2 = bg.run(best, arg)
        assert torch.equal(b2, gbest) and torch.equal(a2, garg)
        # handler shards: rank r "computes" its tuple range, all-gather completes the table
        ntup, n = 9, 4
        full = np.arange(ntup * n * n, dtype=np.float32).reshape(ntup, n, n)
        table = torch.zeros((ntup, n, n))
        b, e = sharding.shard_tuples(ntup, world, rank)
        table[b:e] = torch.from_numpy(full[b:e])
        sharding.all_gather_table(table, ntup)
        ntup2 = 8
        table2 = torch.zeros((ntup2, n, n))
        b, e = sharding.shard_tuples(ntup2, world, rank)
        table2[b:e] = torch.from_numpy(full[b:e])
        sharding.all_gather_table(table2, ntup2)
        if rank == 0:
            ref = oracle.plain_batch(hmm, syms, offs)
            ok = (np.array_equal(gbest.numpy(), ref.min(1)) and np.array_equal(garg.numpy(), ref.argmin(1))
                  and np.array_equal(table.numpy(), full) and np.array_equal(table2.numpy(), full[:ntup2]))
            q.put(bool(ok))
    finally:
        dist.destroy_process_group()


def test_world2_gloo_gather():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(180)
        assert p.exitcode == 0
    assert q.get(timeout=5) is True
