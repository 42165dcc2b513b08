"""API semantics (SURVEY §4.3 'Unit (CPU)'): init/rank/size, LR scaling, step budgets."""
import os

import pytest
import torch

import b200dp as hvd
from b200dp.parallel.collectives import adasum_pair, adasum_tree


def test_init_single_process_is_idempotent(monkeypatch):
    for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK", "LOCAL_WORLD_SIZE"):
        monkeypatch.delenv(k, raising=False)
    w1 = hvd.init(device="cpu")
    w2 = hvd.init(device="cpu")
    assert w1 is w2
    assert (hvd.rank(), hvd.size(), hvd.local_rank(), hvd.local_size()) == (0, 1, 0, 1)
    assert hvd.nccl_built() in (True, False)
    hvd.shutdown()
    assert not hvd.is_initialized()


def test_uninitialised_access_raises():
    with pytest.raises(RuntimeError):
        hvd.rank()


def test_rank_env_from_mpirun_names(monkeypatch):
    for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK", "LOCAL_WORLD_SIZE"):
        monkeypatch.delenv(k, raising=False)
    monkeypatch.setenv("OMPI_COMM_WORLD_RANK", "0")
    monkeypatch.setenv("OMPI_COMM_WORLD_SIZE", "1")
    monkeypatch.setenv("OMPI_COMM_WORLD_LOCAL_RANK", "0")
    monkeypatch.setenv("OMPI_COMM_WORLD_LOCAL_SIZE", "1")
    hvd.init(device="cpu")
    assert hvd.size() == 1 and hvd.local_size() == 1


def test_lr_scaler_rules(monkeypatch):
    hvd.init(device="cpu")
    from b200dp.runtime import world as W
    w = W.world()
    w.size, w.local_size = 8, 4
    assert hvd.lr_scaler(False) == 8
    monkeypatch.setattr(W, "nccl_built", lambda: True)
    assert hvd.lr_scaler(True) == 4
    monkeypatch.setattr(W, "nccl_built", lambda: False)
    assert hvd.lr_scaler(True) == 1
    w.size, w.local_size = 1, 1


def test_single_rank_collectives_are_identity():
    hvd.init(device="cpu")
    t = torch.arange(5.0)
    assert torch.equal(hvd.allreduce(t), t)
    assert torch.equal(hvd.allreduce(t, op=hvd.Adasum), t)
    assert torch.equal(hvd.broadcast(t, 0), t)
    assert hvd.broadcast_object({"a": 1}) == {"a": 1}


def test_adasum_pair_properties():
    a = torch.tensor([1.0, 0.0])
    b = torch.tensor([0.0, 2.0])
    assert torch.allclose(adasum_pair(a, b), a + b)           # orthogonal -> sum
    assert torch.allclose(adasum_pair(a, a), a)               # identical -> average (= itself)
    c = torch.randn(100)
    assert torch.allclose(adasum_pair(c, 3 * c), 2 * c, atol=1e-5)  # parallel -> mean
    with pytest.raises(ValueError):
        adasum_tree([a, b, a])
    r = adasum_tree([a, b, a, b])
    assert r.shape == a.shape


def test_distributed_optimizer_factory():
    opt = hvd.DistributedOptimizer(hvd.Adam(lr=0.01), op=hvd.Adasum)
    assert opt.op is hvd.Adasum and opt.lr == 0.01 and opt.base.epsilon == 1e-8
    assert hvd.Adam.keras().epsilon == 1e-7
    assert hvd.Adam(lr=0.001).scaled(8).lr == pytest.approx(0.008)
    lin = torch.nn.Linear(3, 2)
    topt = hvd.DistributedOptimizer(torch.optim.SGD(lin.parameters(), lr=0.1))
    hvd.init(device="cpu")
    lin(torch.ones(1, 3)).sum().backward()
    before = lin.weight.detach().clone()
    topt.step()
    assert not torch.equal(before, lin.weight)
    with pytest.raises(TypeError):
        hvd.DistributedOptimizer(object())


def test_alias_package_matches():
    import k8s_distributed_deeplearning_b200 as k
    assert k.init is hvd.init
    from k8s_distributed_deeplearning_b200.models import spec
    assert spec.get_spec().num_params == 3274634


def test_gpu_data_plane_refuses_ranks_on_different_hosts(monkeypatch):
    """The CUDA data plane is one NVSwitch domain (peer memory mapped with CUDA IPC).  Ranks on different hosts must get a
    sentence from init(), not an opaque cudaIpcOpenMemHandle failure."""
    import torch.distributed as dist
    from b200dp.runtime import world as W

    def fake_gather(out, obj):
        out[0], out[1] = "pod-a", "pod-b"

    monkeypatch.setattr(dist, "all_gather_object", fake_gather)
    w = W.World(rank=0, size=2, local_rank=0, local_size=1, host="pod-a")
    with pytest.raises(RuntimeError, match="single-node"):
        W._require_single_node(w)

    def same_host(out, obj):
        for i in range(len(out)):
            out[i] = "pod-a"

    monkeypatch.setattr(dist, "all_gather_object", same_host)
    W._require_single_node(w)
    with pytest.raises(RuntimeError, match="exceeds"):
        W._require_single_node(W.World(rank=0, size=9, local_rank=0, local_size=9, host="pod-a"))
