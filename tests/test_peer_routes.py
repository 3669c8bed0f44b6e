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
