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
    hd.call("ctr_merge_sorted_segments", _st(), ptr(t_keys), n, ends, world, ptr(ok), ptr(oi))
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
