"""Single-process behaviour of the TF1-style session and the Keras-style fit."""
import json
import os

import numpy as np
import pytest
import torch

import b200dp as hvd
from b200dp.data import mnist
from b200dp.utils.eventlog import read_events


@pytest.fixture(scope="module")
def data():
    (x, y), (xt, yt) = mnist.synthetic_mnist(640, 128, seed=5)
    return mnist.preprocess(x), y, mnist.preprocess(xt), yt


def _engine(lr=2e-3, **kw):
    return hvd.make_engine(hvd.get_spec("mnist_cnn", 0.25), hvd.DistributedOptimizer(hvd.Adam(lr)), engine="torch", **kw)


def test_session_stops_at_step_logs_and_checkpoints(tmp_path, data, capsys):
    hvd.init(device="cpu")
    x, y, _, _ = data
    eng = _engine()
    hooks = [hvd.BroadcastGlobalVariablesHook(0), hvd.StopAtStepHook(last_step=25 // hvd.size()),
             hvd.LoggingTensorHook(every_n_iter=10)]
    gen = mnist.train_input_generator(x, y, batch_size=32, seed=0)
    losses = []
    with hvd.MonitoredTrainingSession(eng, checkpoint_dir=str(tmp_path / "ck"), hooks=hooks,
                                      save_checkpoint_secs=None, save_checkpoint_steps=10) as sess:
        while not sess.should_stop():
            losses.append(sess.run(*next(gen)).loss_value())
    assert eng.global_step == 25 and len(losses) == 25
    assert np.mean(losses[-5:]) < np.mean(losses[:5])                   # it learns
    out = capsys.readouterr()
    steps = [json.loads(l)["step"] for l in out.out.splitlines() if '"train_step"' in l]
    assert steps == [1, 11, 21]                                         # every 10 iterations
    recs = [json.loads(l) for l in out.out.splitlines() if '"train_step"' in l]
    # the record carries what the reference's log line carries, in fields: the interval as time per step, plus rate and lr
    assert recs[0]["lr"] == pytest.approx(2e-3) and recs[0]["batch"] == 32 and "step_ms" not in recs[0]
    for r in recs[1:]:
        assert r["step_ms"] > 0 and r["images_per_sec"] == pytest.approx(32 * hvd.size() / (r["step_ms"] / 1e3), rel=0.05)
    files = sorted(os.listdir(tmp_path / "ck"))
    assert "latest" in files and "ckpt-25.safetensors" in files and "ckpt-10.safetensors" in files
    with pytest.raises(RuntimeError):
        sess.run(*next(gen))


def test_session_auto_resume_reproduces_next_step(tmp_path, data):
    hvd.init(device="cpu")
    x, y, _, _ = data
    ck = str(tmp_path / "ck")
    a = _engine(seed=1)
    with hvd.MonitoredTrainingSession(a, ck, [hvd.StopAtStepHook(last_step=6)], save_checkpoint_secs=None) as s:
        i = 0
        while not s.should_stop():
            s.run(x[i * 32:(i + 1) * 32], y[i * 32:(i + 1) * 32]); i += 1
    ref_next = a.train_step(x[192:224], y[192:224]).loss_value()         # step 7 on the live engine
    b = _engine(seed=99)                                                # different init: must be overwritten
    with hvd.MonitoredTrainingSession(b, ck, [hvd.StopAtStepHook(last_step=7)], save_checkpoint_secs=None) as s:
        assert b.global_step == 6
        got = s.run(x[192:224], y[192:224]).loss_value()
        assert s.should_stop()
    assert got == pytest.approx(ref_next, rel=0, abs=0)                 # bitwise-equal continuation
    # restored past the budget -> stops immediately
    c = _engine()
    with hvd.MonitoredTrainingSession(c, ck, [hvd.StopAtStepHook(last_step=5)], save_checkpoint_secs=None) as s:
        assert s.should_stop()


def test_session_error_does_not_write_final_checkpoint(tmp_path, data):
    hvd.init(device="cpu")
    eng = _engine()
    with pytest.raises(ZeroDivisionError):
        with hvd.MonitoredTrainingSession(eng, str(tmp_path / "ck"), [], save_checkpoint_secs=None):
            1 / 0
    assert not [f for f in os.listdir(tmp_path / "ck") if f.startswith("ckpt-")]


def test_keras_fit_callbacks_eval_export(tmp_path, data):
    hvd.init(device="cpu")
    x, y, xt, yt = data
    model = hvd.MNISTModel(width=0.25, engine="torch", seed=2)
    opt = hvd.DistributedOptimizer(hvd.Adam.keras(lr=2e-3 * hvd.lr_scaler()), op=hvd.Average)
    model.compile(optimizer=opt, loss=hvd.SparseCategoricalCrossentropy(from_logits=True), metrics=["accuracy"])
    ck = hvd.callbacks.ModelCheckpoint(str(tmp_path / "ck" / "mnist-{epoch}.safetensors"), save_best_only=True)
    cbs = [hvd.callbacks.BroadcastGlobalVariablesCallback(0), hvd.callbacks.MetricAverageCallback(),
           hvd.callbacks.TensorBoard(str(tmp_path / "logs")), ck]
    batch = 32
    steps_per_epoch = max(1, len(x) // (batch * hvd.size()))
    epochs = max(1, 60 // steps_per_epoch)
    assert (steps_per_epoch, epochs) == (20, 3)
    h = model.fit(x, y, batch_size=batch, epochs=epochs, steps_per_epoch=steps_per_epoch,
                  validation_data=(xt, yt), validation_steps=max(1, len(xt) // batch), callbacks=cbs, verbose=0)
    assert model.engine.global_step == 60
    assert set(h.history) == {"loss", "accuracy", "val_loss", "val_accuracy"}
    assert h.history["loss"][-1] < h.history["loss"][0]
    assert len(ck.saved) >= 1 and all(os.path.exists(p) for p in ck.saved)
    best = min(h.history["val_loss"])
    assert ck.best == pytest.approx(best)
    ev = [f for f in os.listdir(tmp_path / "logs") if f.startswith("events.out.tfevents")]
    assert len(ev) == 1
    recs = read_events(str(tmp_path / "logs" / ev[0]))
    assert [r[0] for r in recs[1:]] == [1, 2, 3] and "epoch_val_loss" in recs[1][1]
    score = model.evaluate(xt, yt, batch_size=batch)
    assert len(score) == 2 and 0 <= score[1] <= 1 and score[1] > 0.5
    model.save(str(tmp_path / "final_model"))
    m2 = hvd.MNISTModel(width=0.25, engine="torch", seed=77)
    m2.load_weights(str(tmp_path / "final_model"))
    assert m2.evaluate(xt, yt, batch_size=batch)[0] == pytest.approx(score[0], abs=1e-6)
    assert model.predict(xt[:7]).shape == (7, 10)


def test_best_only_skips_worse_epochs(tmp_path):
    cb = hvd.callbacks.ModelCheckpoint(str(tmp_path / "m-{epoch}.safetensors"), save_best_only=True)

    class Fake:
        class engine:
            @staticmethod
            def state():
                return {"tensors": {"params": torch.zeros(2)}, "meta": {}}
    cb.set_model(Fake)
    for e, v in enumerate([1.0, 0.5, 0.7, 0.4]):
        cb.on_epoch_end(e, {"val_loss": v})
    assert [os.path.basename(p) for p in cb.saved] == ["m-1.safetensors", "m-2.safetensors", "m-4.safetensors"]


def test_timeline_env_writes_host_spans_for_the_cpu_engine(tmp_path, monkeypatch):
    """HOROVOD_TIMELINE on the CPU/gloo engine: a Chrome trace with forward / backward / adam spans per step."""
    import json
    import numpy as np
    import b200dp as hvd
    out = tmp_path / "tl.json"
    monkeypatch.setenv("HOROVOD_TIMELINE", str(out))
    eng = hvd.make_engine(hvd.get_spec("mnist_cnn", 0.25), hvd.Adam(1e-3), engine="torch")
    x, y = np.random.rand(8, 784).astype(np.float32), np.arange(8) % 10
    with hvd.MonitoredTrainingSession(eng, hooks=[hvd.StopAtStepHook(last_step=3)]) as sess:
        while not sess.should_stop():
            sess.run(x, y)
    assert eng.host_timeline is None                                   # detached at exit
    tr = json.load(open(out))["traceEvents"]
    names = [e["name"] for e in tr]
    assert names.count("forward") == 3 and names.count("backward") == 3 and names.count("adam") == 3
    assert all(e["ph"] == "X" and e["dur"] > 0 for e in tr)
    fw = [e for e in tr if e["name"] == "forward"]
    assert fw[0]["ts"] < fw[1]["ts"] < fw[2]["ts"]


def test_lr_schedule_and_warmup_callbacks(monkeypatch):
    import numpy as np
    import b200dp as hvd
    from b200dp.data import mnist
    from b200dp.train import trainer as T
    (x, y), _ = mnist.synthetic_mnist(64, 8, seed=2)
    x = mnist.preprocess(x)

    def model():
        m = hvd.Model(hvd.get_spec("mnist_cnn", 0.25), engine="torch", seed=1)
        m.compile(optimizer=hvd.Adam(1e-3), loss=hvd.SparseCategoricalCrossentropy(from_logits=True), metrics=["accuracy"])
        return m
    # staircase schedule: x1 for epoch 0, x0 afterwards -> the weights stop moving after the first epoch
    m = model()
    sched = hvd.callbacks.LearningRateScheduleCallback(1e-3, lambda e: 1.0 if e < 1 else 0.0)
    snaps = []

    class Snap(hvd.callbacks.Callback):
        def on_epoch_end(self, epoch, logs=None):
            snaps.append(self.model.engine.params.data.clone())
    h = m.fit(x, y, batch_size=16, epochs=3, verbose=0, callbacks=[sched, Snap()])
    assert sched.history == [1e-3, 0.0, 0.0] and h.history["lr"] == [1e-3, 0.0, 0.0]
    assert torch_equal(snaps[1], snaps[2]) and not torch_equal(snaps[0], m.engine.params.data * 0)
    # warm-up in a world of 4: starts at lr/4, climbs every batch, lands on lr exactly
    monkeypatch.setattr(T._w, "is_initialized", lambda: True)
    monkeypatch.setattr(T._w, "size", lambda: 4)
    warm = hvd.callbacks.LearningRateWarmupCallback(4e-3, warmup_epochs=2)
    monkeypatch.undo()
    m2 = model()
    m2.fit(x, y, batch_size=16, epochs=3, verbose=0, callbacks=[warm])
    hs = warm.history
    assert abs(hs[0] - 1e-3) < 1e-12 and all(b >= a for a, b in zip(hs, hs[1:])) and max(hs) <= 4e-3 + 1e-12
    assert m2.engine.adam.lr == 4e-3 and m2.engine.state()["meta"]["optimizer"]["lr"] == 4e-3


def torch_equal(a, b):
    import torch
    return bool(torch.equal(a, b))
