"""End-to-end: the sm_100a engine against the fp32/bf16-rounded PyTorch oracle (SURVEY §4.3 'End-to-end')."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

import b200dp as hvd
from b200dp.data import mnist
from b200dp.models.b200_engine import B200Engine
from b200dp.train.engine import TorchEngine


@pytest.fixture(scope="module")
def data():
    (x, y), (xt, yt) = mnist.synthetic_mnist(1024, 256, seed=9)
    return mnist.preprocess(x), y.astype(np.int32), mnist.preprocess(xt), yt.astype(np.int32)


def _pair(B, dropout, use_tc=True, use_graph=False, width=1.0, lr=1e-3):
    spec = hvd.get_spec("mnist_cnn", width)
    opt = hvd.DistributedOptimizer(hvd.Adam(lr))
    g = B200Engine(spec, opt, seed=3, batch_size=B, dropout=dropout, use_tc=use_tc, use_graph=use_graph)
    c = TorchEngine(spec, opt, seed=3, device="cuda", dropout=dropout, compute_dtype=torch.bfloat16)
    return g, c


@pytest.mark.parametrize("use_tc", [False, True])
@pytest.mark.parametrize("B,dropout", [(128, False), (100, True), (32, True)])
def test_steps_match_oracle(data, B, dropout, use_tc):
    x, y, _, _ = data
    g, c = _pair(B, dropout, use_tc)
    assert torch.allclose(g.master, c.params.data)
    for s in range(3):
        xb, yb = x[s * B:(s + 1) * B], y[s * B:(s + 1) * B]
        rg, rc = g.train_step(xb, yb), c.train_step(xb, yb)
        assert abs(rg.loss_value() - rc.loss_value()) < 3e-2 * max(1.0, abs(rc.loss_value())), (s, rg.loss_value(), rc.loss_value())
        assert abs(int(rg.correct) - int(rc.correct)) <= max(2, B // 20)
    torch.cuda.synchronize()
    # after 3 Adam steps every weight moved by <= ~3*lr; compare the update direction tensor by tensor
    from b200dp.models.spec import init_flat
    init = init_flat(g.spec, 3, "cuda")
    for p in g.spec.params:
        dg = (g.master - init)[p.offset:p.offset + p.numel]
        dc = (c.params.data - init)[p.offset:p.offset + p.numel]
        cos = torch.nn.functional.cosine_similarity(dg, dc, dim=0).item()
        assert cos > 0.9, f"{p.name}: update direction cosine {cos}"


def test_graph_replay_equals_eager(data):
    x, y, _, _ = data
    B = 64
    a, _ = _pair(B, True, use_graph=False)
    b, _ = _pair(B, True, use_graph=True)
    for s in range(6):
        xb, yb = x[s * B:(s + 1) * B], y[s * B:(s + 1) * B]
        la, lb = a.train_step(xb, yb).loss_value(), b.train_step(xb, yb).loss_value()
        assert abs(la - lb) < 2e-3 * max(1.0, abs(la)), (s, la, lb)
    assert len(b._graphs) == 2                                        # one graph per gradient-buffer parity
    assert b.launches_per_step >= 10


def test_loss_decreases_eval_and_checkpoint_roundtrip(data, tmp_path):
    x, y, xt, yt = data
    B = 128
    g, _ = _pair(B, True, use_graph=True, lr=2e-3)
    first = g.train_step(x[:B], y[:B]).loss_value()
    for it in range(60):
        i = (it % 8) * B
        last = g.train_step(x[i:i + B], y[i:i + B])
    assert last.loss_value() < 0.5 * first
    ev = g.eval_batch(xt, yt)
    assert ev.accuracy() > 0.8 and np.isfinite(ev.loss_value())
    assert g.predict(xt[:10]).shape == (10, 10)
    st = g.state()
    nxt = g.train_step(x[:B], y[:B]).loss_value()
    h, _ = _pair(B, True, use_graph=True, lr=2e-3)
    h.load_state(st["tensors"], st["meta"])
    assert h.global_step == 61
    assert torch.equal(h.master, st["tensors"]["params"].cuda())
    got = h.train_step(x[:B], y[:B]).loss_value()
    assert abs(got - nxt) < 1e-3 * max(1.0, abs(nxt))                   # same next-step loss after resume
    # and the checkpoint loads into the CPU engine (shared layout)
    c = TorchEngine(g.spec, hvd.Adam(1e-3), seed=0)
    c.load_state(st["tensors"], st["meta"])
    assert c.global_step == 61


def test_keras_front_end_on_gpu(data):
    x, y, xt, yt = data
    hvd.init(device="cuda")
    model = hvd.MNISTModel(engine="b200", batch_size=64)
    model.compile(optimizer=hvd.DistributedOptimizer(hvd.Adam.keras(2e-3)), loss=hvd.SparseCategoricalCrossentropy(), metrics=["accuracy"])
    h = model.fit(x, y, batch_size=64, epochs=2, steps_per_epoch=16, validation_data=(xt, yt), validation_steps=2,
                  callbacks=[hvd.callbacks.BroadcastGlobalVariablesCallback(0), hvd.callbacks.MetricAverageCallback()], verbose=0)
    assert h.history["loss"][1] < h.history["loss"][0]
    assert 0 <= h.history["val_accuracy"][-1] <= 1


# ---- keras128 (conv3x3x32 -> conv3x3x64 -> pool -> dropout -> dense128 -> dropout -> dense10) on the same engine ----
def _kpair(B, dropout, use_tc=True, use_graph=False, width=1.0, lr=1e-3):
    spec = hvd.get_spec("keras128", width)
    opt = hvd.DistributedOptimizer(hvd.Adam.keras(lr))
    g = B200Engine(spec, opt, seed=5, batch_size=B, dropout=dropout, use_tc=use_tc, use_graph=use_graph)
    c = TorchEngine(spec, opt, seed=5, device="cuda", dropout=dropout, compute_dtype=torch.bfloat16)
    return g, c


@pytest.mark.parametrize("use_tc", [False, True])
@pytest.mark.parametrize("B,dropout", [(128, False), (100, True), (32, True)])
def test_keras128_steps_match_oracle(data, B, dropout, use_tc):
    x, y, _, _ = data
    g, c = _kpair(B, dropout, use_tc)
    assert g.keras and g.tc_conv == use_tc and g.F == 9216 and g.D == 128
    for s in range(3):
        xb, yb = x[s * B:(s + 1) * B], y[s * B:(s + 1) * B]
        rg, rc = g.train_step(xb, yb), c.train_step(xb, yb)
        assert abs(rg.loss_value() - rc.loss_value()) < 3e-2 * max(1.0, abs(rc.loss_value())), (s, rg.loss_value(), rc.loss_value())
        assert abs(int(rg.correct) - int(rc.correct)) <= max(2, B // 20)
    torch.cuda.synchronize()
    from b200dp.models.spec import init_flat
    init = init_flat(g.spec, 5, "cuda")
    for p in g.spec.params:
        dg = (g.master - init)[p.offset:p.offset + p.numel]
        dc = (c.params.data - init)[p.offset:p.offset + p.numel]
        cos = torch.nn.functional.cosine_similarity(dg, dc, dim=0).item()
        assert cos > 0.9, f"{p.name}: update direction cosine {cos}"
    # inference path (no dropout kernels) against the oracle
    lg, lc = g.predict(x[:B]), c.predict(x[:B])
    assert (lg - lc).abs().max().item() < 5e-2 * max(1.0, lc.abs().max().item())


def test_keras128_trains_from_graph_and_roundtrips(data, tmp_path):
    x, y, xt, yt = data
    B = 128
    g, _ = _kpair(B, True, use_graph=True, lr=2e-3)
    first = g.train_step(x[:B], y[:B]).loss_value()
    for it in range(60):
        i = (it % 8) * B
        last = g.train_step(x[i:i + B], y[i:i + B])
    assert last.loss_value() < 0.5 * first
    assert len(g._graphs) == 2 and g.launches_per_step >= 10
    ev = g.eval_batch(xt, yt)
    assert ev.accuracy() > 0.8 and np.isfinite(ev.loss_value())
    st = g.state()
    nxt = g.train_step(x[:B], y[:B]).loss_value()
    h, _ = _kpair(B, True, use_graph=True, lr=2e-3)
    h.load_state(st["tensors"], st["meta"])
    assert abs(h.train_step(x[:B], y[:B]).loss_value() - nxt) < 1e-3 * max(1.0, abs(nxt))


def test_keras128_through_model_api(data):
    x, y, xt, yt = data
    m = hvd.Model(hvd.get_spec("keras128"), seed=2, batch_size=64)
    m.compile(optimizer=hvd.DistributedOptimizer(hvd.Adam.keras(2e-3)), loss=hvd.SparseCategoricalCrossentropy(from_logits=True),
              metrics=["accuracy"])
    hist = m.fit(x, y, batch_size=64, epochs=2, verbose=0)
    assert type(m.engine).__name__ == "B200Engine" and m.engine.keras
    assert hist.history["loss"][-1] < hist.history["loss"][0]


def test_set_lr_takes_effect_without_graph_recapture(data):
    """LR schedules: the kernels read the rate from device memory, so set_lr() needs no new CUDA graph."""
    x, y, _, _ = data
    B = 64
    g, _ = _pair(B, True, use_graph=True)
    for s in range(4):
        g.train_step(x[s * B:(s + 1) * B], y[s * B:(s + 1) * B])
    graphs = dict(g._graphs)
    assert len(graphs) == 2
    g.set_lr(0.0)
    torch.cuda.synchronize()
    before = g.master.clone()
    for s in range(2):
        g.train_step(x[s * B:(s + 1) * B], y[s * B:(s + 1) * B])
    torch.cuda.synchronize()
    assert torch.equal(g.master, before)                       # zero rate: Adam's update is exactly zero
    g.set_lr(1e-3)
    g.train_step(x[:B], y[:B])
    torch.cuda.synchronize()
    assert not torch.equal(g.master, before) and g._graphs == graphs and g.state()["meta"]["optimizer"]["lr"] == 1e-3


def test_poison_mode_changes_nothing_when_the_protocol_holds(data, monkeypatch):
    """B200DP_POISON=1 overwrites consumed gradient buffers and gathered activations with NaN (SURVEY §5.2 debug build).
    A correct step never reads them again, so the run must track the normal one (up to the order of the fp32 atomics in
    the split-K and weight-gradient reductions, which differs from run to run)."""
    x, y, _, _ = data
    B = 64
    a, _ = _pair(B, True, use_graph=True)
    monkeypatch.setenv("B200DP_POISON", "1")
    b, _ = _pair(B, True, use_graph=True)
    assert b.poison and not a.poison
    for s in range(8):
        xb, yb = x[s * B:(s + 1) * B], y[s * B:(s + 1) * B]
        la, lb = a.train_step(xb, yb).loss_value(), b.train_step(xb, yb).loss_value()
        assert np.isfinite(lb) and abs(la - lb) < 2e-3 * max(1.0, abs(la)), (s, la, lb)
    torch.cuda.synchronize()
    assert bool(torch.isfinite(b.master).all())
    assert (a.w16.float() - b.w16.float()).abs().max().item() < 1e-2


def test_state_async_snapshot_equals_state_and_is_not_disturbed_by_later_steps(data):
    x, y, _, _ = data
    B = 64
    g, _ = _pair(B, True, use_graph=True)
    for s in range(4):
        g.train_step(x[s * B:(s + 1) * B], y[s * B:(s + 1) * B])
    want = g.state()
    pend = g.state_async()
    for s in range(4, 8):                       # training goes on while the snapshot drains
        g.train_step(x[s * B:(s + 1) * B], y[s * B:(s + 1) * B])
    got = pend.result()
    for k in ("params", "adam_m", "adam_v"):
        assert torch.equal(got["tensors"][k], want["tensors"][k]), k
    assert got["meta"]["global_step"] == 4
    assert not torch.equal(g.state()["tensors"]["params"], want["tensors"]["params"])
