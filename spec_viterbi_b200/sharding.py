"""Multi-GPU plumbing: one process per GPU, torch.distributed for the (small) exchanges.

The path shards on two axes with no exchange inside the time loop:
  * independent observation sequences are dealt to ranks (length-balanced), the HMM is replicated;
  * the spec precompute is split by symbol-tuple index, every rank builds a contiguous range of
    handlers into its slice of the full table, then one all-gather completes the table everywhere.
Collectives are used only to gather final scores and handler shards (NCCL over NVLink on GPUs,
gloo in the CPU tests).
"""
from __future__ import annotations

from typing import List, Sequence, Tuple

import numpy as np


def shard_sequences(lengths: Sequence[int], world: int) -> List[np.ndarray]:
    """Longest-processing-time deal of sequences to ranks: returns, per rank, the (ascending)
    indices of its sequences.  Every index appears exactly once; total lengths are balanced."""
    lengths = np.asarray(lengths, dtype=np.int64)
    order = np.argsort(-lengths, kind="stable")
    loads = np.zeros(world, dtype=np.int64)
    counts = np.zeros(world, dtype=np.int64)
    owner = np.empty(lengths.size, dtype=np.int64)
    for i in order:
        # least loaded rank; ties -> fewest sequences, then lowest rank (deterministic)
        r = int(np.lexsort((np.arange(world), counts, loads))[0])
        owner[i] = r
        loads[r] += lengths[i]
        counts[r] += 1
    return [np.flatnonzero(owner == r) for r in range(world)]


def shard_tuples(ntuples: int, world: int, rank: int) -> Tuple[int, int]:
    """Contiguous, equal (+-1) range of symbol-tuple indices of `rank`."""
    base, rem = divmod(int(ntuples), world)
    begin = rank * base + min(rank, rem)
    return begin, begin + base + (1 if rank < rem else 0)


def take_sequences(symbols: np.ndarray, offsets: np.ndarray, idx: np.ndarray):
    """Packed sub-batch (symbols, offsets) of the sequences `idx`."""
    lens = (offsets[1:] - offsets[:-1])[idx].astype(np.uint64)
    offs = np.zeros(idx.size + 1, np.uint64)
    offs[1:] = np.cumsum(lens)
    out = np.empty(int(offs[-1]), symbols.dtype)
    for k, i in enumerate(idx):
        out[int(offs[k]):int(offs[k + 1])] = symbols[int(offsets[i]):int(offsets[i + 1])]
    return out, offs


def gather_best(local_best, local_arg, idx: np.ndarray, nseq_total: int, group=None):
    """All ranks receive (best[nseq_total], arg[nseq_total]) in the original sequence order.
    local_best / local_arg: torch tensors (CUDA with nccl, CPU with gloo) for the sequences `idx`."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size(group)
    dev = local_best.device
    n_local = torch.tensor([idx.size], dtype=torch.int64, device=dev)
    sizes = [torch.zeros(1, dtype=torch.int64, device=dev) for _ in range(world)]
    dist.all_gather(sizes, n_local, group=group)
    cap = int(max(int(s.item()) for s in sizes))
    pack = torch.zeros((cap, 3), dtype=torch.float64, device=dev)  # (index, best, arg): exact in f64
    pack[: idx.size, 0] = torch.as_tensor(idx, dtype=torch.float64, device=dev)
    pack[: idx.size, 1] = local_best.to(torch.float64)
    pack[: idx.size, 2] = local_arg.to(torch.float64)
    parts = [torch.empty_like(pack) for _ in range(world)]
    dist.all_gather(parts, pack, group=group)
    best = torch.empty(nseq_total, dtype=torch.float32, device=dev)
    arg = torch.empty(nseq_total, dtype=torch.int32, device=dev)
    for r in range(world):
        k = int(sizes[r].item())
        ids = parts[r][:k, 0].to(torch.int64)
        best[ids] = parts[r][:k, 1].to(torch.float32)
        arg[ids] = parts[r][:k, 2].to(torch.int32)
    return best, arg


class BestGather:
    """Device-side all-gather of (best score, best state) with NO host synchronisation: every rank knows the
    deal (shard_sequences is deterministic), so only the payload travels -- 8 bytes per sequence.  Buffers are
    allocated once; `run` may sit inside a timed region / be enqueued behind the step kernel."""

    def __init__(self, parts, nseq_total: int, device, group=None):
        import torch
        self.group = group
        self.world = len(parts)
        self.cap = max(int(p.size) for p in parts)
        self.recv_best = torch.empty((self.world, self.cap), dtype=torch.float32, device=device)
        self.recv_arg = torch.empty((self.world, self.cap), dtype=torch.int32, device=device)
        self.send_best = torch.full((self.cap,), float("inf"), dtype=torch.float32, device=device)
        self.send_arg = torch.full((self.cap,), -1, dtype=torch.int32, device=device)
        # scatter map: position (rank, k) of the receive buffer -> original sequence index
        src = np.concatenate([r * self.cap + np.arange(p.size) for r, p in enumerate(parts)])
        dst = np.concatenate(parts)
        self.src = torch.as_tensor(src, dtype=torch.int64, device=device)
        self.dst = torch.as_tensor(dst, dtype=torch.int64, device=device)
        self.best = torch.empty(nseq_total, dtype=torch.float32, device=device)
        self.arg = torch.empty(nseq_total, dtype=torch.int32, device=device)

    def run(self, local_best, local_arg):
        """local_*: this rank's results in the order of its shard.  Returns (best, arg) in original order."""
        import torch.distributed as dist
        k = local_best.numel()
        self.send_best[:k].copy_(local_best)
        self.send_arg[:k].copy_(local_arg)
        if self.world > 1:
            dist.all_gather_into_tensor(self.recv_best.view(-1), self.send_best, group=self.group)
            dist.all_gather_into_tensor(self.recv_arg.view(-1), self.send_arg, group=self.group)
        else:
            self.recv_best[0].copy_(self.send_best)
            self.recv_arg[0].copy_(self.send_arg)
        self.best.index_copy_(0, self.dst, self.recv_best.view(-1).index_select(0, self.src))
        self.arg.index_copy_(0, self.dst, self.recv_arg.view(-1).index_select(0, self.src))
        return self.best, self.arg


def all_gather_table(table, ntuples: int, group=None) -> None:
    """In-place all-gather of a [ntuples, n, n] handler table whose rank-r slice
    [shard_tuples(ntuples, world, r)] has been computed locally."""
    import torch.distributed as dist
    if not dist.is_available() or not dist.is_initialized():
        return  # single process: the local range is the whole table
    world = dist.get_world_size(group)
    if world == 1:
        return
    base, rem = divmod(int(ntuples), world)
    if rem == 0:
        rank = dist.get_rank(group)
        dist.all_gather_into_tensor(table.view(-1), table[rank * base:(rank + 1) * base].reshape(-1), group=group)
        return
    for r in range(world):  # ragged ranges: one broadcast per owner
        b, e = shard_tuples(ntuples, world, r)
        dist.broadcast(table[b:e], src=dist.get_global_rank(group, r) if group is not None else r, group=group)
