"""The peer-exchange kernels on ONE GPU: the "peers" of a route are separate local buffers, which exercises exactly the
addressing the NVLink path uses (tests/mgpu_parity.py runs the same kernels across real peers).  Bit-exact."""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def hd():
    from deep_learning_b200 import _lib
    return _lib.Handle(0)


def _st():
    return torch.cuda.current_stream().cuda_stream


def _route(bufs, seg_counts, dst_row, rank=0):
    from deep_learning_b200._lib import PeerRoute
    rt = PeerRoute()
    rt.world, rt.rank, rt.dst_off = len(bufs), rank, 0
    acc = 0
    for p, b in enumerate(bufs):
        acc += int(seg_counts[p])
        rt.slab[p], rt.seg_end[p], rt.dst_row[p] = b.data_ptr(), acc, int(dst_row[p])
    return rt


@pytest.mark.parametrize("world,counts", [(1, [700]), (3, [300, 0, 211]), (8, [17, 40, 0, 1, 99, 256, 3, 64])])
@pytest.mark.parametrize("k", [16, 8])
def test_gather_rows_push_routes_rows(hd, world, counts, k):
    from deep_learning_b200._lib import ptr
    g = np.random.RandomState(world + k)
    width, stride, wire, R, lb = k + 1, 24 if k == 16 else 12, 20 if k == 16 else 12, 3000, 12
    table = g.standard_normal((R, stride)).astype(np.float32)
    n = sum(counts)
    rows = g.randint(0, R, size=n)
    keys = (rows | (5 << lb)).astype(np.uint32)            # owner bits above local_bits are masked away
    dst_row = g.randint(0, 50, size=world)
    bufs = [torch.full((dst_row[p] + counts[p] + 3, wire), 7.0, device="cuda") for p in range(world)]
    rt = _route(bufs, counts, dst_row)
    t_table, t_keys = torch.from_numpy(table).cuda(), torch.from_numpy(keys.view(np.int32)).cuda()
    hd.call("ctr_gather_rows_push", _st(), ptr(t_table), stride, ptr(t_keys), n, lb, width, wire, C.byref(rt))
    torch.cuda.synchronize()
    begin = 0
    for p in range(world):
        got = bufs[p].cpu().numpy()
        want = np.full_like(got, 7.0)
        seg = rows[begin:begin + counts[p]]
        want[dst_row[p]:dst_row[p] + counts[p], :width] = table[seg, :width]
        want[dst_row[p]:dst_row[p] + counts[p], width:] = 0.0
        assert np.array_equal(got, want), "peer %d" % p
        begin += counts[p]


@pytest.mark.parametrize("world", [1, 2, 8])
def test_merge_sorted_segments_is_the_stable_sort(hd, world):
    from deep_learning_b200._lib import ptr
    g = np.random.RandomState(90 + world)
    segs = [np.unique(g.randint(0, 4000, size=g.randint(0, 3000))).astype(np.uint32) for _ in range(world)]
    if world > 1:
        segs[1] = segs[1][:0]                                 # an empty segment
    keys = np.concatenate(segs) if segs else np.zeros(0, np.uint32)
    n = len(keys)
    ends = (C.c_uint32 * world)(*np.cumsum([len(s) for s in segs]).tolist())
    t_keys = torch.from_numpy(keys.view(np.int32)).cuda()
    ok, oi = torch.zeros(n, dtype=torch.int32, device="cuda"), torch.zeros(n, dtype=torch.int32, device="cuda")
    hd.call("ctr_merge_sorted_segments", _st(), ptr(t_keys), n, ends, world, ptr(ok), ptr(oi), None)
    order = np.argsort(keys, kind="stable")
    assert np.array_equal(ok.cpu().numpy().view(np.uint32), keys[order])
    assert np.array_equal(oi.cpu().numpy().view(np.uint32), order.astype(np.uint32))


def test_finalize_push_equals_finalize_then_exchange(hd):
    """runs of a sorted lookup list -> gradient rows; the pushed rows must equal ctr_embed_bwd_finalize's, placed per route,
    and the run of padding keys past the routed list must be dropped"""
    from deep_learning_b200._lib import ptr
    g = np.random.RandomState(5)
    CH, gs, width = 32, 20, 18
    run_len = np.r_[g.randint(1, 6, size=400), [700, 90], g.randint(1, 4, size=100)]      # two hot rows
    U = len(run_len) - 1                                     # the last run plays the padding key
    n = int(run_len.sum())
    seg = np.r_[0, np.cumsum(run_len)].astype(np.uint32)
    part = g.standard_normal((n + n // CH + 2, gs)).astype(np.float32)
    t_seg = torch.from_numpy(np.concatenate([seg, np.zeros(1, np.uint32)]).view(np.int32)).cuda()
    t_nu = torch.tensor([U + 1], dtype=torch.int32, device="cuda")
    t_part = torch.from_numpy(part).cuda()
    ref = torch.zeros(n, gs, device="cuda")
    hd.call("ctr_embed_bwd_finalize", _st(), ptr(t_seg), ptr(t_nu), n, width, gs, ptr(t_part), ptr(ref))
    counts, dst_row = [150, 0, U - 150], [4, 9, 0]
    bufs = [torch.full((dst_row[p] + counts[p] + 2, gs), 7.0, device="cuda") for p in range(3)]
    rt = _route(bufs, counts, dst_row, rank=1)
    hd.call("ctr_embed_bwd_finalize_push", _st(), ptr(t_seg), ptr(t_nu), n, width, gs, ptr(t_part), C.byref(rt))
    torch.cuda.synchronize()
    ref = ref.cpu().numpy()
    begin = 0
    for p in range(3):
        got = bufs[p].cpu().numpy()
        rows = slice(dst_row[p], dst_row[p] + counts[p])
        assert np.array_equal(got[rows, :width], ref[begin:begin + counts[p], :width]), "peer %d" % p
        assert np.all(got[:dst_row[p]] == 7.0) and np.all(got[dst_row[p] + counts[p]:] == 7.0)
        begin += counts[p]


def test_barrier_single_rank_and_timeout_flag(hd):
    """world 1 passes at once; a peer that never arrives sets the status word instead of hanging"""
    from deep_learning_b200._lib import ptr
    flags = [torch.zeros(64, dtype=torch.int32, device="cuda") for _ in range(2)]
    status = torch.zeros(1, dtype=torch.int32, device="cuda")
    rt = _route(flags[:1], [0], [0])
    hd.call("ctr_peer_barrier", _st(), C.byref(rt), 0, 1, ptr(status), 100)
    assert int(status.item()) == 0 and int(flags[0][0].item()) == 1
    rt = _route(flags, [0, 0], [0, 0], rank=0)               # "rank 1" never writes flags[0][1]
    hd.call("ctr_peer_barrier", _st(), C.byref(rt), 0, 2, ptr(status), 50)
    assert int(status.item()) == 2                            # 1 + the missing peer's rank
    assert int(flags[1][0].item()) == 2                       # the own arrival was still published


# ---- device-side routing (the sharded step without a host read of the split sizes) ---------------------------------------
@pytest.mark.parametrize("world", [1, 2, 3, 8])
def test_plan_routes_equals_the_host_formula(hd, world):
    """ctr_peer_plan_routes == deep_learning_b200.shard.peer_routes for every rank of a random count matrix"""
    from deep_learning_b200 import _lib
    from deep_learning_b200._lib import ptr
    from deep_learning_b200.shard import peer_routes
    g = np.random.RandomState(40 + world)
    mat = g.randint(1, 500, size=(world, world)).astype(np.int32)
    if world > 1:
        mat[0, -1] = 0
    t_mat = torch.from_numpy(mat.reshape(-1).copy()).cuda()
    for rank in range(world):
        plan = torch.full((_lib.PLAN_WORDS,), -1, dtype=torch.int32, device="cuda")
        status = torch.zeros(_lib.STATUS_WORDS, dtype=torch.int32, device="cuda")
        hd.call("ctr_peer_plan_routes", _st(), ptr(t_mat), world, rank, 10 ** 9, 10 ** 9, ptr(plan), ptr(status))
        p = plan.cpu().numpy()
        send, recv, vec, grad = peer_routes(mat.tolist(), rank)
        assert p[_lib.PLAN_VEC:_lib.PLAN_VEC + world].tolist() == vec[0] and p[_lib.PLAN_VEC + 16:_lib.PLAN_VEC + 16 + world].tolist() == vec[1]
        assert p[_lib.PLAN_GRAD:_lib.PLAN_GRAD + world].tolist() == grad[0] and p[_lib.PLAN_GRAD + 16:_lib.PLAN_GRAD + 16 + world].tolist() == grad[1]
        assert p[_lib.PLAN_VEC + 32] == p[_lib.PLAN_NREQ] == sum(recv) and p[_lib.PLAN_GRAD + 32] == p[_lib.PLAN_U] == sum(send)
        assert status.cpu().numpy().tolist() == [0, 0, 0, 0]
    # capacity: the first owner that would receive more than the buffers hold is reported (1 + owner), same on every rank
    worst = mat.sum(axis=0)
    cap = int(np.sort(worst)[-1]) - 1
    first = int(np.argmax(worst > cap))
    for rank in range(world):
        status = torch.zeros(_lib.STATUS_WORDS, dtype=torch.int32, device="cuda")
        hd.call("ctr_peer_plan_routes", _st(), ptr(t_mat), world, rank, 10 ** 9, cap, ptr(plan), ptr(status))
        assert status.cpu().numpy().tolist() == [0, 0, 0, 1 + first]


def test_kernels_follow_a_device_route_block(hd):
    """gather+push, key push, merge and run structure with placement and counts read from a plan block on the device
    == the same calls with host-side routes"""
    from deep_learning_b200 import _lib
    from deep_learning_b200._lib import ptr
    g = np.random.RandomState(77)
    world, rank = 3, 1
    mat = np.array([[40, 25, 0], [31, 7, 64], [0, 90, 12]], dtype=np.int32)
    t_mat = torch.from_numpy(mat.reshape(-1).copy()).cuda()
    plan = torch.zeros(_lib.PLAN_WORDS, dtype=torch.int32, device="cuda")
    status = torch.zeros(_lib.STATUS_WORDS, dtype=torch.int32, device="cuda")
    hd.call("ctr_peer_plan_routes", _st(), ptr(t_mat), world, rank, 10 ** 6, 10 ** 6, ptr(plan), ptr(status))
    from deep_learning_b200.shard import peer_routes
    send, recv, vec, grad = peer_routes(mat.tolist(), rank)
    n_req, U, cap = sum(recv), sum(send), 400
    # gather + push along the vector route
    k, stride, wire, R, lb = 16, 24, 20, 500, 10
    table = g.standard_normal((R, stride)).astype(np.float32)
    req = g.randint(0, R, size=cap).astype(np.uint32)
    t_table, t_req = torch.from_numpy(table).cuda(), torch.from_numpy(req.view(np.int32)).cuda()
    rows_of = [int(max(vec[1])) + int(max(recv)) + 4] * world

    def bufs():
        return [torch.full((rows_of[p], wire), 7.0, device="cuda") for p in range(world)]
    a, b = bufs(), bufs()
    rt_host = _route(a, recv, vec[1], rank=rank)
    hd.call("ctr_gather_rows_push", _st(), ptr(t_table), stride, ptr(t_req), n_req, lb, k + 1, wire, C.byref(rt_host))
    rt_dev = _route(b, [0] * world, [0] * world, rank=rank)
    rt_dev.dyn = plan.data_ptr() + 4 * _lib.PLAN_VEC
    hd.call("ctr_gather_rows_push", _st(), ptr(t_table), stride, ptr(t_req), cap, lb, k + 1, wire, C.byref(rt_dev))
    for x, y in zip(a, b):
        assert torch.equal(x, y)
    # key push along the gradient route (width 1)
    uniq = torch.arange(1000, 1000 + cap, dtype=torch.int32, device="cuda")
    ka = [torch.full((300,), -1, dtype=torch.int32, device="cuda") for _ in range(world)]
    kb = [t.clone() for t in ka]
    rt_host = _route(ka, send, grad[1], rank=rank)
    hd.call("ctr_peer_push_rows", _st(), ptr(uniq), U, 1, C.byref(rt_host))
    rt_dev = _route(kb, [0] * world, [0] * world, rank=rank)
    rt_dev.dyn = plan.data_ptr() + 4 * _lib.PLAN_GRAD
    hd.call("ctr_peer_push_rows", _st(), ptr(uniq), cap, 1, C.byref(rt_dev))
    for x, y in zip(ka, kb):
        assert torch.equal(x, y)
    # merge + run structure of a request list whose length only the device knows
    segs = [np.sort(g.choice(5000, size=c, replace=False)).astype(np.uint32) for c in recv]
    keys = np.concatenate(segs + [np.full(cap - n_req, 0xDEADBEEF, np.uint32)])
    t_keys = torch.from_numpy(keys.view(np.int32)).cuda()
    ok, oi = (torch.full((cap,), -1, dtype=torch.int32, device="cuda") for _ in range(2))
    hd.call("ctr_merge_sorted_segments", _st(), ptr(t_keys), cap, None, world, ptr(ok), ptr(oi), plan.data_ptr() + 4 * _lib.PLAN_VEC)
    order = np.argsort(keys[:n_req], kind="stable")
    assert np.array_equal(ok.cpu().numpy().view(np.uint32)[:n_req], keys[:n_req][order])
    assert np.array_equal(oi.cpu().numpy().view(np.uint32)[:n_req], order.astype(np.uint32))
    assert np.all(ok.cpu().numpy()[n_req:] == -1)
    uq, seg, uidx = (torch.full((cap + 2,), -1, dtype=torch.int32, device="cuda") for _ in range(3))
    nu = torch.zeros(1, dtype=torch.int32, device="cuda")
    ws = torch.zeros(4096, dtype=torch.uint8, device="cuda")
    hd.call("ctr_unique_segments", _st(), ptr(ok), cap, ptr(uq), ptr(seg), ptr(uidx), ptr(nu), ptr(ws), ws.numel(),
            plan.data_ptr() + 4 * _lib.PLAN_NREQ)
    want, first = np.unique(keys[:n_req], return_index=False), None
    Uo = int(nu.item())
    assert Uo == len(want) and np.array_equal(uq.cpu().numpy().view(np.uint32)[:Uo], want)
    assert int(seg[Uo].item()) == n_req
    # the all-zero staging row behind the unique rows
    buf = torch.full((U + 20, wire), 3.0, device="cuda")
    hd.call("ctr_zero_row_at", _st(), ptr(buf), wire, wire, plan.data_ptr() + 4 * _lib.PLAN_U)
    got = buf.cpu().numpy()
    assert np.all(got[U] == 0.0) and np.all(np.delete(got, U, axis=0) == 3.0)
