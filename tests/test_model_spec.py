import math

import pytest
import torch

from b200dp.models import spec as S
from b200dp.models import torch_ref
from b200dp.utils.philox import dropout_keep_mask
from b200dp.parallel.buckets import plan_buckets, plan_arena_buckets


def test_param_counts_match_reference_model():
    sp = S.get_spec("mnist_cnn")
    assert sp.num_params == 3_274_634
    assert sp.fwd_flops_per_image() == 2 * (784 * 25 * 32 + 196 * 800 * 64 + 3136 * 1024 + 1024 * 10)
    shapes = {p.name: p.shape for p in sp.params}
    assert shapes["conv1.kernel"] == (5, 5, 1, 32) and shapes["conv2.kernel"] == (5, 5, 32, 64)
    assert shapes["dense.kernel"] == (3136, 1024) and shapes["out.kernel"] == (1024, 10)
    assert all(p.offset % S.ALIGN_ELEMS == 0 for p in sp.params)
    assert S.get_spec("keras128").num_params == 1_199_882
    assert S.get_spec("mnist_cnn", 6.0).num_params == 117_490_954 or S.get_spec("mnist_cnn", 6.0).num_params > 1e8
    with pytest.raises(ValueError):
        S.get_spec("resnet")


def test_width_16_param_count():
    assert abs(S.get_spec("mnist_cnn", 16.0).num_params - 835e6) / 835e6 < 0.01


@pytest.mark.parametrize("arch", ["mnist_cnn", "keras128"])
def test_forward_shapes_and_eval_determinism(arch):
    sp = S.get_spec(arch, 0.25)
    flat = S.init_flat(sp, seed=3)
    x = torch.rand(5, 784)
    a = torch_ref.forward(sp, sp.views(flat), x, training=False)
    b = torch_ref.forward(sp, sp.views(flat), x, training=False)
    assert a.shape == (5, 10) and torch.equal(a, b)
    t1 = torch_ref.forward(sp, sp.views(flat), x, training=True, seed=1, step=0)
    t2 = torch_ref.forward(sp, sp.views(flat), x, training=True, seed=1, step=1)
    assert not torch.equal(t1, t2)                          # dropout mask changes per step


def test_philox_mask_statistics_and_determinism():
    m1 = dropout_keep_mask(100_000, 0.5, seed=7, step=3)
    m2 = dropout_keep_mask(100_000, 0.5, seed=7, step=3)
    assert torch.equal(m1, m2)
    assert abs(m1.float().mean().item() - 0.5) < 0.01
    assert abs(dropout_keep_mask(100_000, 0.25, 7, 3).float().mean().item() - 0.75) < 0.01
    assert not torch.equal(m1, dropout_keep_mask(100_000, 0.5, seed=7, step=4))
    # Philox4x32-10 known-answer test (Random123 kat: counter=0,key=0)
    from b200dp.utils.philox import philox4x32
    import numpy as np
    z = np.zeros(1, np.uint32)
    out = [int(v[0]) for v in philox4x32(z, z, z, z, 0, 0)]
    assert out == [0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8]


def test_adam_tf_matches_manual():
    p = torch.tensor([1.0, -2.0]); g = torch.tensor([0.5, 0.25])
    m = torch.zeros(2); v = torch.zeros(2)
    torch_ref.adam_step_(p, g, m, v, 1, lr=0.1)
    lr_t = 0.1 * math.sqrt(1 - 0.999) / (1 - 0.9)
    exp = torch.tensor([1.0, -2.0]) - lr_t * (0.1 * g) / ((0.001 * g * g).sqrt() + 1e-8)
    assert torch.allclose(p, exp, atol=1e-6)
    # torch placement agrees with torch.optim.Adam
    w = torch.nn.Parameter(torch.tensor([1.0, -2.0]))
    o = torch.optim.Adam([w], lr=0.1)
    w.grad = g.clone(); o.step()
    p2 = torch.tensor([1.0, -2.0]); m2 = torch.zeros(2); v2 = torch.zeros(2)
    torch_ref.adam_step_(p2, g, m2, v2, 1, lr=0.1, eps_mode="torch")
    assert torch.allclose(p2, w.detach(), atol=1e-6)


def test_bucket_planner():
    assert plan_buckets([10, 10, 10], 20) == [[0, 1], [2]]
    assert plan_buckets([50, 1, 1], 20) == [[0], [1, 2]]
    assert plan_buckets([], 20) == []
    sp = S.get_spec()
    bs = plan_arena_buckets(sp, [["dense.bias", "out."], ["dense.kernel"], ["conv1.", "conv2."]])
    assert [b.name for b in bs] == ["dense.bias+out.", "dense.kernel", "conv1.+conv2."]
    assert sum(b.numel for b in bs) == sp.total
    dk = sp.param("dense.kernel")
    assert bs[1].start == dk.offset and bs[1].numel >= dk.numel
    with pytest.raises(ValueError):
        plan_arena_buckets(sp, [["conv1.", "dense."], ["conv2.", "out."]])


def test_custom_spec_trains_through_the_torch_engine_and_checks_its_layers():
    import numpy as np
    import b200dp as hvd
    L = hvd.LayerSpec
    spec = hvd.custom_spec("tiny3", [L("conv", dict(k=3, cout=4, padding="VALID")), L("pool"), L("conv", dict(k=3, cout=8)), L("pool"),
                                     L("flatten"), L("dense", dict(units=16, relu=True, name="fc")), L("dropout", dict(rate=0.1)),
                                     L("dense", dict(units=3, name="out"))])
    assert spec.num_classes == 3 and spec.param("fc.kernel").shape == (7 * 7 * 8, 16)      # 28 -> 26 -> 13 -> 13 -> 7
    rng = np.random.RandomState(0)
    x, y = rng.rand(48, 784).astype(np.float32), rng.randint(0, 3, 48)
    x[np.arange(48), y * 200] = 5.0                          # an easy signal
    m = hvd.Model(spec, seed=1)
    m.compile(optimizer=hvd.Adam(5e-3), loss=hvd.SparseCategoricalCrossentropy(from_logits=True), metrics=["accuracy"])
    h = m.fit(x, y, batch_size=16, epochs=6, verbose=0)
    assert type(m.engine).__name__ == "TorchEngine" and h.history["loss"][-1] < h.history["loss"][0]
    assert m.predict(x[:5]).shape == (5, 3)
    for bad in ([L("dense", dict(units=3))], [L("flatten"), L("conv", dict(k=3, cout=2)), L("dense", dict(units=3))],
                [L("conv", dict(k=3, cout=2)), L("flatten")], [L("attention"), L("flatten"), L("dense", dict(units=3))]):
        with pytest.raises(ValueError):
            hvd.custom_spec("bad", bad)
    with pytest.raises(ValueError):
        hvd.custom_spec("mnist_cnn", [L("flatten"), L("dense", dict(units=10))])


from hypothesis import given, settings, strategies as st


@settings(max_examples=200, deadline=None)
@given(st.lists(st.integers(min_value=0, max_value=1 << 22), max_size=40), st.integers(min_value=1, max_value=1 << 22))
def test_plan_buckets_invariants(sizes, cap):
    """Every tensor lands in exactly one bucket, ready order is kept, a bucket only exceeds the cap when it is one tensor."""
    bs = plan_buckets(sizes, cap)
    assert [i for b in bs for i in b] == list(range(len(sizes)))
    for b in bs:
        assert b and (sum(sizes[i] for i in b) <= cap or len(b) == 1)
    for a, b in zip(bs, bs[1:]):                      # greedy: the next tensor really did not fit
        assert sum(sizes[i] for i in a) + sizes[b[0]] > cap


@settings(max_examples=60, deadline=None)
@given(st.sampled_from(["mnist_cnn", "keras128"]), st.sampled_from([0.25, 0.5, 1.0, 1.5, 2.0, 6.0]))
def test_arena_buckets_tile_every_width(arch, width):
    sp = S.get_spec(arch, width)
    bs = plan_arena_buckets(sp, [["dense.bias", "out."], ["dense.kernel"], ["conv1.", "conv2."]])
    assert sorted((b.start, b.end) for b in bs)[0][0] == 0 and sum(b.numel for b in bs) == sp.total
    for b in bs:
        assert b.start % S.ALIGN_ELEMS == 0                   # 256-byte aligned ranges: vectorised kernels need no peeling
        assert all(b.start <= o and o + n <= b.end for o, n in b.segments)
    assert sum(n for b in bs for _, n in b.segments) == sp.num_params
