"""CPU, world_size 2, gloo: the host-side half of the multi-GPU path (deep_learning_b200/shard.py) -- shard map,
split sizes, the three all-to-alls and the fixed-peer-order owner sum -- with numpy standing in for the CUDA
kernels (np.unique for ctr_sort_ids+ctr_unique_segments, fancy indexing for ctr_gather_rows, np.add.at in peer
order for ctr_segment_sum_rows).  Parity target: the unsharded gather / scatter-add on the concatenated batch."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from deep_learning_b200.shard import PAD_KEY, ShardPlan, exchange, split_sizes


def test_shard_plan_bit_exact_map():
    for R, N in [(1, 1), (10, 2), (1000, 8), (100_000_000, 8), (7, 4)]:
        pl = ShardPlan(R, N)
        ids = np.unique(np.concatenate([np.arange(min(R, 50)), np.random.RandomState(0).randint(0, R, 200), [R - 1]])).astype(np.int64)
        assert np.array_equal(pl.owner(ids), ids % N) and np.array_equal(pl.local_row(ids), ids // N)
        keys = pl.key(ids)
        assert keys.max() < PAD_KEY and pl.local_row(ids).max() < pl.rows_local
        assert np.array_equal(pl.id_of_key(keys), ids)                      # the key is invertible
        # sorting by key is owner-major, local row minor: each owner's slice is contiguous
        ks = np.sort(keys)
        assert np.all(np.diff(ks >> pl.local_bits) >= 0)
    with pytest.raises(ValueError):
        ShardPlan(1 << 32, 8)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _batch_ids(rank, B, F, R, hot):
    g = np.random.RandomState(100 + rank)
    ids = g.randint(0, R, size=(B, F)).astype(np.int64)
    ids[:, 0] = g.randint(0, hot, size=B)          # a hot field: heavy duplication inside and across ranks
    return ids


def _worker(rank, world, port, R, B, F, w, q):
    try:
        os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
        dist.init_process_group("gloo", rank=rank, world_size=world)
        pl = ShardPlan(R, world)
        table = np.random.RandomState(7).standard_normal((R, w)).astype(np.float32)
        shard = np.zeros((pl.rows_local, w), dtype=np.float32)
        mine = table[rank::world]
        shard[:len(mine)] = mine
        all_ids = [_batch_ids(r, B, F, R, hot=5) for r in range(world)]
        ids = all_ids[rank]
        # --- GPU-kernel roles in numpy: de-dup, owner-major unique keys, per-owner counts, inverse map
        keys = pl.key(ids).astype(np.uint32)
        uniq, inv = np.unique(keys.reshape(-1), return_inverse=True)
        counts = np.bincount((uniq >> pl.local_bits).astype(np.int64), minlength=world)
        # --- all-to-all #1: counts then keys
        send_counts, recv_counts = split_sizes(torch.from_numpy(counts.astype(np.int32)))
        assert send_counts == counts.tolist()
        req = exchange(torch.from_numpy(uniq.astype(np.int64)), send_counts, recv_counts).numpy()
        assert np.all((req >> pl.local_bits) == rank)                    # I only receive keys I own
        # --- owner gather + all-to-all #2
        rows_send = shard[req & ((1 << pl.local_bits) - 1)]
        rows_recv = exchange(torch.from_numpy(rows_send), recv_counts, send_counts).numpy()
        assert np.array_equal(rows_recv, table[pl.id_of_key(uniq.astype(np.int64))])   # bit-exact, unique-list order
        assert np.array_equal(rows_recv[inv].reshape(B, F, w), table[ids])             # == unsharded gather
        # --- backward: one gradient row per unique key -> owner, summed in peer order
        def lookup_grads(r):
            return np.random.RandomState(900 + r).standard_normal((B * F, w)).astype(np.float32)
        g_u = np.zeros((len(uniq), w), dtype=np.float32)
        np.add.at(g_u, inv, lookup_grads(rank))
        g_recv = exchange(torch.from_numpy(g_u), send_counts, recv_counts).numpy()
        order = np.argsort(req & ((1 << pl.local_bits) - 1), kind="stable")            # stable: keeps peer order
        own_rows, start = np.unique((req & ((1 << pl.local_bits) - 1))[order], return_index=True)
        g_own = np.add.reduceat(g_recv[order], start, axis=0) if len(order) else np.zeros((0, w), np.float32)
        # reference: the same per-unique partial sums, added rank 0 first then rank 1, on the full table
        ref = np.zeros((R, w), dtype=np.float32)
        for r in range(world):
            k_r = pl.key(all_ids[r]).astype(np.uint32).reshape(-1)
            u_r, inv_r = np.unique(k_r, return_inverse=True)
            gu_r = np.zeros((len(u_r), w), dtype=np.float32)
            np.add.at(gu_r, inv_r, lookup_grads(r))
            ref[pl.id_of_key(u_r.astype(np.int64))] += gu_r
        got = np.zeros_like(ref)
        got[own_rows * world + rank] = g_own
        mine_mask = np.zeros(R, dtype=bool)
        mine_mask[rank::world] = True
        assert np.array_equal(got[mine_mask], ref[mine_mask])            # bit-exact: same addends, same order
        # dense side: a SUM allreduce of per-rank shares of the global mean
        share = torch.tensor([float(rank + 1) / world])
        dist.all_reduce(share)
        assert abs(share.item() - sum(range(1, world + 1)) / world) < 1e-6
        dist.destroy_process_group()
        q.put((rank, "ok"))
    except Exception as e:   # surface the failure in the parent
        import traceback
        q.put((rank, traceback.format_exc()))


@pytest.mark.parametrize("R,B,F", [(1000, 64, 5), (17, 33, 3)])
def test_exchange_schedule_world2_gloo(R, B, F):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, R, B, F, 4, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    for rank, msg in res:
        assert msg == "ok", "rank %d:\n%s" % (rank, msg)
