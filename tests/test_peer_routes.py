"""Host logic of the peer-memory exchange (shard.peer_routes): storing row i of a source list at
`dst_row[p] + (i - seg_begin[p])` of peer p's buffer reproduces, for every rank, exactly the layout the variable-size
all-to-all (`shard.exchange`, peer-major) delivers -- for the vectors (owner -> requester) and the gradients
(requester -> owner).  Pure numpy, no GPU, no process group."""
import numpy as np
import pytest

from deep_learning_b200 import shard


def _segments(ends):
    begin = 0
    for p, e in enumerate(ends):
        yield p, begin, e
        begin = e


@pytest.mark.parametrize("world", [1, 2, 3, 8])
def test_routes_reproduce_the_all_to_all_layout(world):
    g = np.random.RandomState(10 + world)
    mat = g.randint(0, 50, size=(world, world))
    mat[g.rand(world, world) < 0.2] = 0                      # empty segments
    # unique lists (requester side, owner-major) and request lists (owner side, peer-major) of tagged rows
    uniq = [[("row", p, o, i) for o in range(world) for i in range(mat[p][o])] for p in range(world)]
    req = [[("row", p, o, i) for p in range(world) for i in range(mat[p][o])] for o in range(world)]
    rows_recv = [[None] * len(uniq[p]) for p in range(world)]
    g_recv = [[None] * len(req[o]) for o in range(world)]
    for r in range(world):
        send, recv, vec, grad = shard.peer_routes(mat.tolist(), r)
        assert send == mat[r].tolist() and recv == mat[:, r].tolist()
        assert vec[0][-1] == len(req[r]) and grad[0][-1] == len(uniq[r])
        for p, b, e in _segments(vec[0]):                   # owner r pushes the rows peer p asked for
            for i in range(b, e):
                rows_recv[p][vec[1][p] + i - b] = req[r][i]
        for o, b, e in _segments(grad[0]):                  # requester r pushes its gradient rows to owner o
            for i in range(b, e):
                g_recv[o][grad[1][o] + i - b] = uniq[r][i]
    for r in range(world):
        assert rows_recv[r] == uniq[r]                      # vectors arrive in the requester's unique-list order
        assert g_recv[r] == req[r]                          # gradients arrive peer-major, rank 0's rows first


def test_reshard_rows_partition_every_id_exactly_once():
    """per-shard checkpoints restored at another world size (CtrEngine.load_shard): every row id of the table ends up on
    exactly one new rank, at local row id // w_new"""
    import numpy as np
    from deep_learning_b200 import shard as S
    R = 1003
    for w_old, w_new in ((8, 1), (2, 1), (1, 4), (8, 3), (3, 8)):
        seen = np.zeros(R, dtype=np.int64)
        for r_new in range(w_new):
            rows_new = (R + w_new - 1) // w_new
            got = np.full(rows_new, -1, dtype=np.int64)
            for r_old in range(w_old):
                n_old = (R + w_old - 1) // w_old            # shards are allocated to the rounded-up size
                keep, loc = S.reshard_rows(n_old, w_old, r_old, w_new, r_new, R)
                ids = np.arange(n_old, dtype=np.int64)[keep] * w_old + r_old
                assert (ids % w_new == r_new).all() and (ids < R).all()
                assert (got[loc] == -1).all()
                got[loc] = ids
                seen[ids] += 1
            valid = got[got >= 0]
            assert np.array_equal(valid, np.arange(len(got))[got >= 0] * w_new + r_new)
        assert (seen == 1).all()
