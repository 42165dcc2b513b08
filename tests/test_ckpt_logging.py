import io
import json
import os

import numpy as np
import pytest
import torch

import b200dp as hvd
from b200dp.ckpt.manager import CheckpointManager, export_model, load_tensors
from b200dp.utils.eventlog import EventWriter, crc32c, read_events
from b200dp.utils.logging import JsonLogger
from b200dp.data import mnist


def _state(step, val=1.0):
    return {"tensors": {"params": torch.full((8,), val), "adam_m": torch.zeros(8), "adam_v": torch.ones(8)},
            "meta": {"arch": "mnist_cnn", "global_step": step}}


def test_roundtrip_bitwise_and_latest_pointer(tmp_path):
    m = CheckpointManager(str(tmp_path), keep=2, save_secs=None)
    assert m.latest() is None and m.restore() is None
    t = torch.randn(1000)
    p = m.save({"tensors": {"params": t}, "meta": {"x": 1}}, step=7)
    assert os.path.basename(p) == "ckpt-7.safetensors"
    r = m.restore()
    assert torch.equal(r["tensors"]["params"], t) and r["meta"]["step"] == 7 and r["meta"]["x"] == 1
    assert open(tmp_path / "latest").read() == "ckpt-7.safetensors"
    for s in (8, 9, 10):
        m.save(_state(s), s)
    assert [s for s, _ in m.list()] == [9, 10]             # keep-last-2
    assert m.latest().endswith("ckpt-10.safetensors")
    os.remove(tmp_path / "latest")                         # lost pointer -> directory scan
    assert m.latest().endswith("ckpt-10.safetensors")
    assert not [f for f in os.listdir(tmp_path) if ".tmp." in f]


def test_disabled_manager_is_noop(tmp_path):
    m = CheckpointManager(None)
    assert not m.enabled and m.save(_state(1), 1) is None and m.restore() is None and not m.due(5)


def test_due_by_steps_and_secs(tmp_path):
    m = CheckpointManager(str(tmp_path), save_secs=None, save_steps=5)
    assert not m.due(3) and m.due(5)
    m.save(_state(5), 5)
    assert not m.due(5)
    m2 = CheckpointManager(str(tmp_path / "b"), save_secs=0.0)
    assert m2.due(1)


def test_export_model(tmp_path):
    p = export_model(str(tmp_path / "final_model"), {"params": torch.arange(4.0)}, {"arch": "mnist_cnn"})
    t, meta = load_tensors(p)
    assert torch.equal(t["params"], torch.arange(4.0)) and meta["arch"] == "mnist_cnn"
    assert json.load(open(tmp_path / "final_model" / "model.json"))["arch"] == "mnist_cnn"


def test_json_logger_schema(capsys):
    buf = io.StringIO()
    lg = JsonLogger(rank=3, size=8, local_rank=3, stream=buf, tf_style=True)
    lg.step(10, 0.25, lr=0.008)
    lg.step(20, torch.tensor(0.125))
    lines = [json.loads(l) for l in buf.getvalue().splitlines()]
    assert len(lines) == 2
    for k in ("ts", "level", "event", "rank", "local_rank", "size", "host", "step", "loss"):
        assert k in lines[0]
    assert lines[0]["rank"] == 3 and lines[0]["size"] == 8 and lines[0]["event"] == "train_step"
    assert lines[1]["loss"] == 0.125
    err = capsys.readouterr().err
    assert "INFO:tensorflow:loss = 0.25, step = 10" in err and "sec)" in err.splitlines()[1]


def test_logger_level_filter():
    buf = io.StringIO()
    lg = JsonLogger(stream=buf, level="WARNING", tf_style=False)
    lg.info("x"); lg.warning("y")
    assert [json.loads(l)["event"] for l in buf.getvalue().splitlines()] == ["y"]


def test_event_file_roundtrip(tmp_path):
    assert crc32c(b"123456789") == 0xE3069283
    w = EventWriter(str(tmp_path))
    w.add_scalars({"epoch_loss": 0.5, "epoch_accuracy": 0.75}, step=1)
    w.add_scalars({"epoch_loss": 0.25}, step=2)
    w.close()
    ev = read_events(w.path)
    assert ev[0] == (0, {})
    assert ev[1][0] == 1 and ev[1][1]["epoch_loss"] == 0.5 and ev[1][1]["epoch_accuracy"] == 0.75
    assert ev[2] == (2, {"epoch_loss": 0.25})
    assert len(open(tmp_path / "scalars.jsonl").read().splitlines()) == 2


def test_data_pipeline(tmp_path):
    d = mnist.ensure_cache_dir(str(tmp_path / "a" / "datasets"))
    assert mnist.ensure_cache_dir(d) == d                              # EEXIST tolerated
    (xtr, ytr), (xte, yte) = mnist.load_data("MNIST-data-0", cache_dir=d, n_train=300, n_test=50)
    assert xtr.shape == (300, 28, 28) and xtr.dtype == np.uint8 and yte.shape == (50,)
    assert os.path.exists(os.path.join(d, "MNIST-data-0.npz"))           # per-rank cache file
    (xtr2, _), _ = mnist.load_data("MNIST-data-0", cache_dir=d)
    assert np.array_equal(xtr, xtr2)
    x = mnist.preprocess(xtr)
    assert x.shape == (300, 784) and x.dtype == np.float32 and 0 <= x.min() and x.max() <= 1
    gen = mnist.train_input_generator(x, ytr, batch_size=64, seed=0)
    seen = [next(gen) for _ in range(9)]
    assert all(b[0].shape == (64, 784) and b[1].shape == (64,) for b in seen)
    first_epoch = np.concatenate([b[1] for b in seen[:4]])               # 300//64 = 4 full batches, tail dropped
    assert len(first_epoch) == 256


def test_stall_watchdog_warns_and_recovers(capsys):
    import time
    from b200dp.utils.watchdog import StallWatchdog
    wd = StallWatchdog(warn_secs=0.2, poll_secs=0.05).start()
    wd.beat(7)
    time.sleep(0.55)
    assert wd.warnings >= 2
    before = wd.warnings
    for _ in range(6):
        wd.beat(8)
        time.sleep(0.05)
    assert wd.warnings == before                     # beating keeps it quiet
    wd.stop()
    out = capsys.readouterr().out
    assert '"event":"stall"' in out and '"step":7' in out


def test_topology_discovery_on_cpu():
    from b200dp.runtime import topology
    t = topology.discover()
    assert "numa" in t and t["host_cpus"] >= 1 and isinstance(t["gpus"], list)


def test_keras128_falls_back_to_torch_engine():
    import b200dp as hvd
    eng = hvd.make_engine(hvd.get_spec("keras128", 0.25), hvd.Adam(1e-3), engine="torch")
    r = eng.train_step(np.random.rand(4, 784).astype(np.float32), np.array([1, 2, 3, 4]))
    assert np.isfinite(r.loss_value())


def test_prometheus_exporter_serves_step_metrics(tmp_path):
    """the reference defines no Prometheus metrics; ours are optional and served per rank."""
    import urllib.request
    import b200dp as hvd
    from b200dp.utils.metrics import MetricsExporter
    exp = MetricsExporter(rank=0, size=2)
    port = exp.serve(0, addr="127.0.0.1")
    try:
        eng = hvd.make_engine(hvd.get_spec("mnist_cnn", 0.25), hvd.Adam(1e-3), engine="torch")
        x, y = np.random.rand(8, 784).astype(np.float32), np.arange(8) % 10
        hooks = [hvd.StopAtStepHook(last_step=6), hvd.MetricsHook(exp, every_n_iter=2)]
        with hvd.MonitoredTrainingSession(eng, checkpoint_dir=str(tmp_path), hooks=hooks, save_checkpoint_secs=None) as sess:
            sess.metrics = exp
            while not sess.should_stop():
                sess.run(x, y)
        body = urllib.request.urlopen(f"http://127.0.0.1:{port}/metrics", timeout=5).read().decode()
    finally:
        exp.close()
    vals = {l.split("{")[0]: float(l.rsplit(" ", 1)[1]) for l in body.splitlines() if l.startswith("b200dp_") and "{" in l}
    assert vals["b200dp_global_step"] == 5.0                       # sampled at iterations 0, 2, 4 -> steps 1, 3, 5
    assert vals["b200dp_steps_total"] == 4.0 and vals["b200dp_world_size"] == 2.0
    assert vals["b200dp_images_per_second"] > 0 and vals["b200dp_step_seconds"] > 0
    assert np.isfinite(vals["b200dp_loss"]) and vals["b200dp_checkpoints_total"] == 1.0      # the final save
    assert 'rank="0"' in body and vals["b200dp_last_step_timestamp_seconds"] > 1.6e9


def test_restore_skips_a_damaged_newest_checkpoint(tmp_path, capsys):
    m = CheckpointManager(str(tmp_path), keep=5, save_secs=None)
    for step in (10, 20, 30):
        m.save({"tensors": {"params": torch.full((4,), float(step))}, "meta": {"arch": "x"}}, step)
    newest = m.latest()
    with open(newest, "r+b") as f:                      # truncate the newest file in place (bit rot / short write on NFS)
        f.truncate(16)
    got = CheckpointManager(str(tmp_path)).restore()
    assert got["meta"]["step"] == 20 and torch.equal(got["tensors"]["params"], torch.full((4,), 20.0))
    assert "checkpoint_unreadable" in capsys.readouterr().out
    for _, p in m.list():
        with open(p, "r+b") as f:
            f.truncate(8)
    with pytest.raises(RuntimeError, match="no readable checkpoint"):
        CheckpointManager(str(tmp_path)).restore()


def test_async_save_runs_on_a_writer_thread_and_surfaces_errors(tmp_path):
    """CheckpointManager.save_async: the caller hands over a pending-state handle and returns at once; the file appears
    after wait(); an error in the writer is re-raised by wait(), not lost."""
    import threading
    import time
    import torch
    from b200dp.ckpt.manager import CheckpointManager

    gate = threading.Event()

    class Pending:
        def result(self):
            gate.wait(10)
            return {"tensors": {"params": torch.arange(8, dtype=torch.float32)}, "meta": {"arch": "x"}}

    m = CheckpointManager(str(tmp_path), save_steps=5, save_secs=None)
    t0 = time.monotonic()
    assert m.maybe_save_async(Pending, 5) is True
    assert time.monotonic() - t0 < 1.0 and m.latest() is None          # returned without waiting for the state
    assert m.maybe_save_async(Pending, 5) is False                    # same step is not saved twice
    gate.set()
    path = m.wait()
    assert path and m.latest() == path
    r = m.restore()
    assert torch.equal(r["tensors"]["params"], torch.arange(8, dtype=torch.float32)) and r["meta"]["step"] == 5

    class Broken:
        def result(self):
            raise RuntimeError("device lost")

    m.save_async(Broken(), 10)
    import pytest
    with pytest.raises(RuntimeError, match="device lost"):
        m.wait()
