"""CPU: the thread rendezvous of deep_learning_b200/comm.py (LoopbackWorld) that the 1-GPU world-N step test rides on."""
import pytest
import torch

from deep_learning_b200.comm import LoopbackWorld


@pytest.fixture(autouse=True)
def _no_cuda(monkeypatch):
    monkeypatch.setattr(torch.cuda, "set_device", lambda d: None)


def test_objects_travel_between_ranks():
    def body(c):
        return c.all_gather_object(c.rank * 10), c.broadcast_object("from%d" % c.rank, src=2), c.world
    out = LoopbackWorld(4).run(body)
    assert out == [([0, 10, 20, 30], "from2", 4)] * 4


def test_a_failing_rank_does_not_hang_the_others():
    def body(c):
        if c.rank == 1:
            raise ValueError("boom")
        c.barrier()
        return 1
    with pytest.raises(ValueError, match="boom"):
        LoopbackWorld(3).run(body)
