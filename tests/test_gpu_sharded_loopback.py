"""-m gpu, ONE GPU: the row-sharded world-N train step with the N ranks as N threads of this process
(deep_learning_b200/comm.py LoopbackComm).  Each rank is a complete ShardedCtrEngine -- own shard of the table, own peer
slab, id plans, fused gather+push / finalize+push kernels, owner-side merge + Adam, dense-gradient sum out of the slabs;
only the spinning inter-rank barrier kernel is replaced by a host rendezvous plus stream event waits (kernels that wait
on one another must not share a GPU).  The body is tests/mgpu_parity.py's: N ranks on N shards of the batch == the fp64
oracle on the concatenated batch (unique key lists and gathered rows bit-exact, loss / parameters 1e-5), free-running
steps with the id plan prefetched, per-shard checkpoint restore and N -> 1 re-shard."""
import pytest

import mgpu_parity

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("world,alg,adam_mode", [
    (2, "deepfm", "lazy"), (2, "deepfm", "tf_dense"), (2, "wdl", "lazy"), (2, "deepfm_multi", "tf_dense"),
    (4, "deepfm", "lazy"), (8, "deepfm", "lazy"), (8, "deepfm", "tf_dense"), (3, "dnn", "lazy")])
def test_sharded_step_world_n_on_one_gpu(world, alg, adam_mode, capsys):
    from deep_learning_b200.comm import LoopbackWorld
    out = LoopbackWorld(world).run(lambda comm: mgpu_parity.run(comm, alg, adam_mode))
    assert out == [True] * world
    assert "MGPU PARITY OK" in capsys.readouterr().out


@pytest.mark.parametrize("world,alg,adam_mode", [(2, "deepfm", "lazy"), (4, "deepfm", "tf_dense")])
def test_sharded_step_split_backward_order(world, alg, adam_mode, capsys, monkeypatch):
    """CTR_BWD_ORDER=split (opt-in): the gradient rows are pushed ahead of the weight-gradient GEMMs, barrier and owner-side
    update run behind them (CtrEngine._backward_split) -- same results as the plain order"""
    from deep_learning_b200.comm import LoopbackWorld
    monkeypatch.setenv("CTR_BWD_ORDER", "split")
    out = LoopbackWorld(world).run(lambda comm: mgpu_parity.run(comm, alg, adam_mode))
    assert out == [True] * world
    assert "MGPU PARITY OK" in capsys.readouterr().out


@pytest.mark.parametrize("world,alg,adam_mode", [(2, "deepfm", "lazy"), (8, "deepfm", "lazy"), (3, "dnn", "lazy")])
def test_sharded_step_early_global_loss(world, alg, adam_mode, capsys, monkeypatch):
    """peer-memory mode hands the GLOBAL loss to the host right behind the forward pass (each rank's share summed out of the
    peers' slabs in rank order on the loss stream).  The batches of the parity body are small, so the batch-size threshold
    of the loss stream is lowered; the body checks last_loss() == the returned loss == the oracle's, step after step."""
    from deep_learning_b200.comm import LoopbackWorld
    monkeypatch.setenv("CTR_LOSS_STREAM_MIN_BATCH", "1")
    out = LoopbackWorld(world).run(lambda comm: mgpu_parity.run(comm, alg, adam_mode))
    assert out == [True] * world
    assert "MGPU PARITY OK" in capsys.readouterr().out
