"""Row-sharding of the embedding tables over the GPUs of one box and the exchange schedule.

The reference is single-process (no collectives anywhere: grep for nccl/horovod/tf.distribute = 0 hits), so this is
new work with the 1-GPU result on the concatenated batch as its parity target:

    owner(id) = id mod N          local_row(id) = id div N          (bit-exact part of the parity check)
    key(id)   = owner << local_bits | local_row                      (what ctr_sort_ids emits with n_shards = N)

Per step, each rank r:
  1. sorts + de-duplicates the ids of ITS samples; the unique keys come out owner-major, so the slice for owner o
     is contiguous and `counts[o]` long                                  (ctr_sort_ids / ctr_unique_segments /
                                                                          ctr_shard_counts, on the GPU)
  2. all-to-all #1: counts, then the de-duplicated keys                  (exchange_counts / exchange)
  3. owners gather the requested rows; all-to-all #2 returns the vectors in the same owner-major order, i.e. the
     order of the requester's unique list                                (ctr_gather_rows + exchange)
  4. backward: all-to-all #3 carries one gradient row per unique key back to its owner, which sums the
     contributions of all peers in fixed peer order and applies Adam to its rows.

Everything here is plain torch.distributed on whatever device the tensors live on: NCCL on the GPUs, gloo in the
CPU tests (tests/test_shard_gloo.py) that cover this host-side logic with world_size 2.
"""
from __future__ import annotations

from typing import List, Optional, Sequence, Tuple

import numpy as np
import torch
import torch.distributed as dist

PAD_KEY = 0xFFFFFFFF


class ShardPlan:
    """The shard map of one table of `table_rows` rows over `world` ranks."""

    def __init__(self, table_rows: int, world: int):
        if world < 1 or table_rows < 1:
            raise ValueError("ShardPlan: need world >= 1 and table_rows >= 1")
        self.table_rows, self.world = int(table_rows), int(world)
        self.rows_local = (self.table_rows + self.world - 1) // self.world
        self.local_bits = max(1, int(self.rows_local - 1).bit_length())
        self.owner_bits = max(0, int(self.world - 1).bit_length())
        if self.local_bits + self.owner_bits > 31:
            raise ValueError("ShardPlan: owner|local key needs %d bits, at most 31 fit beside the padding key"
                             % (self.local_bits + self.owner_bits))
        self.key_bits = self.local_bits + self.owner_bits

    def owner(self, ids):
        return ids % self.world

    def local_row(self, ids):
        return ids // self.world

    def key(self, ids):
        return (self.owner(ids) << self.local_bits) | self.local_row(ids)

    def id_of_key(self, keys):
        mask = (1 << self.local_bits) - 1
        return (keys & mask) * self.world + (keys >> self.local_bits)


def exchange_counts(send_counts: torch.Tensor, group=None) -> torch.Tensor:
    """all-to-all of one integer per peer: recv[o] = what rank o will send me."""
    recv = torch.empty_like(send_counts)
    dist.all_to_all_single(recv, send_counts, group=group)
    return recv


def exchange(send: torch.Tensor, send_counts: Sequence[int], recv_counts: Sequence[int], out: Optional[torch.Tensor] = None,
             group=None) -> torch.Tensor:
    """Variable-size all-to-all of rows: `send` is [sum(send_counts), ...] laid out peer-major; returns
    [sum(recv_counts), ...] peer-major (rank 0's rows first: the fixed peer order the owner-side sum relies on)."""
    n_recv = int(sum(recv_counts))
    tail = tuple(send.shape[1:])
    if out is None:
        out = torch.empty((n_recv,) + tail, dtype=send.dtype, device=send.device)
    else:
        out = out[:n_recv]
    n_send = int(sum(send_counts))
    dist.all_to_all_single(out, send[:n_send].contiguous(), output_split_sizes=[int(c) for c in recv_counts],
                           input_split_sizes=[int(c) for c in send_counts], group=group)
    return out


def split_sizes(counts_dev: torch.Tensor, group=None) -> Tuple[List[int], List[int]]:
    """(send_counts, recv_counts) as host lists from the device-side per-owner unique counts: ONE all-to-all and ONE
    device->host read per step (NCCL needs the split sizes on the host)."""
    send = counts_dev.to(torch.int64)
    recv = exchange_counts(send, group)
    both = torch.stack([send, recv]).cpu().tolist()
    return [int(x) for x in both[0]], [int(x) for x in both[1]]


# ---- peer-memory exchange (NVLink stores instead of all-to-all #2 / #3) ----------------------------------------------
def gather_counts(send_counts: torch.Tensor, group=None) -> torch.Tensor:
    """all-gather of every rank's per-owner counts: mat[p, o] = unique rows rank p requests from owner o.  With the
    whole matrix on every rank, each rank can place its rows inside a PEER's buffer without asking (peer_routes)."""
    world = dist.get_world_size(group)
    mat = torch.empty(world * send_counts.numel(), dtype=send_counts.dtype, device=send_counts.device)
    dist.all_gather_into_tensor(mat, send_counts.contiguous(), group=group)
    return mat.view(world, send_counts.numel())


def peer_routes(mat: Sequence[Sequence[int]], rank: int):
    """From the count matrix, the two placements of rank `rank` (each a pair (seg_end[p], dst_row[p])):

    vectors   -- source = the rows peers requested of me, peer-major (recv_counts = column `rank`); the rows of peer p
                 land at p's unique-list offset of owner `rank`:  sum_{o < rank} mat[p][o]
    gradients -- source = my unique list, owner-major (send_counts = row `rank`); my rows land in owner o's receive
                 buffer behind those of the ranks before me:       sum_{p < rank} mat[p][o]
    Both reproduce exactly the layout `exchange` (all_to_all_single, peer-major) produces."""
    world = len(mat)
    send = [int(mat[rank][o]) for o in range(world)]
    recv = [int(mat[p][rank]) for p in range(world)]

    def ends(c):
        out, acc = [], 0
        for x in c:
            acc += x
            out.append(acc)
        return out
    vec = (ends(recv), [sum(int(mat[p][o]) for o in range(rank)) for p in range(world)])
    grad = (ends(send), [sum(int(mat[p][o]) for p in range(rank)) for o in range(world)])
    return send, recv, vec, grad


def reshard_rows(n_old_local: int, w_old: int, r_old: int, w_new: int, r_new: int, table_rows: int):
    """Rows of an old shard (rank r_old of w_old: local row l holds id l * w_old + r_old) that rank r_new of w_new owns
    now: (boolean mask over the old shard's rows, their new local rows id // w_new).  Used when a per-shard checkpoint
    is restored at another world size (train on 8, serve on 1)."""
    ids = np.arange(n_old_local, dtype=np.int64) * w_old + r_old
    keep = (ids % w_new == r_new) & (ids < table_rows)
    return keep, ids[keep] // w_new
