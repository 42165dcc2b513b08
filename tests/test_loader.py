"""The native minibatch loader (csrc/loader.cpp) against its pure-Python twin and the reference generator's semantics
(horovod/tensorflow_mnist.py:76-85: fresh permutation per pass, consecutive full batches, tail dropped)."""
import numpy as np
import pytest
import torch

from b200dp.data.loader import BatchLoader, epoch_permutation


def _data(n=103, row=12):
    x = np.arange(n * row, dtype=np.float32).reshape(n, row)
    y = (np.arange(n) % 10).astype(np.int64)
    return x, y


def test_native_loader_is_used_and_matches_python_twin_batch_for_batch():
    x, y = _data()
    nat = BatchLoader(x, y, 16, seed=7, native=True, pin=False)
    py = BatchLoader(x, y, 16, seed=7, native=False, pin=False)
    assert nat.native and not py.native and nat.batches_per_epoch == 6
    for _ in range(20):                                   # more than three passes
        (xa, ya), (xb, yb) = nat.next(), py.next()
        assert torch.equal(xa, xb) and torch.equal(ya, yb)
    nat.close()


def test_epochs_are_permutations_with_the_tail_dropped_and_rows_stay_paired():
    x, y = _data(50, 3)
    ld = BatchLoader(x, y, 8, seed=1, pin=False)
    for epoch in range(3):
        seen = []
        for _ in range(ld.batches_per_epoch):
            xb, yb = ld.next()
            rows = (xb[:, 0] / 3).long()
            assert torch.equal(yb.long(), rows % 10)       # labels travel with their rows
            seen += rows.tolist()
        assert len(seen) == 48 and len(set(seen)) == 48      # 6 full batches, 2 rows dropped, no repeats inside a pass
        assert seen == epoch_permutation(50, 1, epoch)[:48].tolist()
    ld.close()
    assert epoch_permutation(50, 1, 0).tolist() != epoch_permutation(50, 1, 1).tolist()
    assert epoch_permutation(50, 1, 0).tolist() != epoch_permutation(50, 2, 0).tolist()      # per-rank seeds differ
    assert sorted(epoch_permutation(50, 1, 0).tolist()) == list(range(50))


def test_unshuffled_order_and_held_slots_are_not_overwritten():
    x, y = _data(40, 2)
    ld = BatchLoader(x, y, 4, shuffle=False, nslots=4, hold=3, pin=False)
    a = ld.next()[0]
    b = ld.next()[0]
    c = ld.next()[0]
    snap = (a.clone(), b.clone(), c.clone())
    import time
    time.sleep(0.05)                                      # give the producer time to run ahead as far as it may
    assert torch.equal(a, snap[0]) and torch.equal(b, snap[1]) and torch.equal(c, snap[2])
    assert a[0, 0].item() == 0 and b[0, 0].item() == 8 and c[0, 0].item() == 16
    d = ld.next()[0]                                      # releases the oldest slot only now
    assert d[0, 0].item() == 24 and torch.equal(b, snap[1]) and torch.equal(c, snap[2])
    ld.close()


def test_bad_arguments():
    x, y = _data(10, 2)
    for kw in (dict(batch_size=0), dict(batch_size=11), dict(batch_size=2, nslots=2, hold=2)):
        with pytest.raises(ValueError):
            BatchLoader(x, y, **kw)
    with pytest.raises(ValueError):
        BatchLoader(x, y[:5], 2)


def test_loader_feeds_a_training_loop():
    import b200dp as hvd
    from b200dp.data import mnist
    (x, y), _ = mnist.synthetic_mnist(256, 16, seed=3)
    ld = BatchLoader(mnist.preprocess(x), y, 32, seed=0, pin=False)
    eng = hvd.make_engine(hvd.get_spec("mnist_cnn", 0.25), hvd.Adam(2e-3), engine="torch")
    losses = [eng.train_step(*ld.next()).loss_value() for _ in range(24)]
    assert np.mean(losses[-4:]) < np.mean(losses[:4])
    ld.close()


@pytest.mark.parametrize("flag", ["1", "0"])
def test_model_fit_draws_from_the_native_loader(monkeypatch, flag):
    import b200dp as hvd
    from b200dp.data import loader as L, mnist
    made = []
    orig = L.BatchLoader.__init__

    def spy(self, *a, **k):
        orig(self, *a, **k)
        made.append(self.native)
    monkeypatch.setattr(L.BatchLoader, "__init__", spy)
    monkeypatch.setenv("B200DP_NATIVE_LOADER", flag)
    (x, y), _ = mnist.synthetic_mnist(128, 16, seed=5)
    m = hvd.Model(hvd.get_spec("mnist_cnn", 0.25), engine="torch", seed=1)
    m.compile(optimizer=hvd.Adam(2e-3), loss=hvd.SparseCategoricalCrossentropy(from_logits=True), metrics=["accuracy"])
    h = m.fit(mnist.preprocess(x), y, batch_size=32, epochs=3, verbose=0)
    assert made == ([True] if flag == "1" else [])
    assert h.history["loss"][-1] < h.history["loss"][0]
