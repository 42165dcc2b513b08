"""world_size 2 / 4 over the launcher with the gloo backend (SURVEY §4.3 'Distributed (CPU)';
BASELINE.json config 1: world_size=2 on CPU/gloo, plumbing, no GPU)."""
import glob
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
RUN = [sys.executable, os.path.join(ROOT, "bin", "b200run")]
W = os.path.join(ROOT, "tests", "workers")


def _launch(np_, script_args, timeout=300, extra=(), env=None):
    e = dict(os.environ, B200DP_LOG_TF_STYLE="0", OMP_NUM_THREADS="2", **(env or {}))
    return subprocess.run(RUN + ["-np", str(np_), "--cpu", *extra, sys.executable, *script_args],
                          capture_output=True, text=True, timeout=timeout, env=e, cwd=ROOT)


@pytest.mark.parametrize("np_", [2, 4])
def test_collectives(tmp_path, np_):
    r = _launch(np_, [os.path.join(W, "collectives_worker.py"), str(tmp_path)])
    assert r.returncode == 0, r.stderr[-3000:]
    res = [json.load(open(p)) for p in sorted(glob.glob(str(tmp_path / "r*.json")))]
    assert len(res) == np_
    for i, d in enumerate(res):
        assert d["rank"] == i and d["size"] == np_ and d["local_size"] == np_
        assert d["overlapped_launches"] >= 3                              # buckets were reduced from the hooks
        for k in ("avg_ok", "sum_ok", "input_untouched", "adasum_ok", "bcast_ok", "replicas_identical", "accumulation_ok"):
            assert d[k], (i, k)
        assert d["obj"] == 0 and d["gather"] == list(range(np_))
        assert d["bf16_avg"] == pytest.approx((np_ - 1) / 2)
        for k in ("async_avg_ok", "async_bcast_ok", "grouped_ok", "grouped_adasum_ok", "alltoall_ok", "alltoall_even_ok",
                  "reducescatter_ok", "minmax_ok", "compression_ok", "ragged_gather_ok"):
            assert d[k], (i, k)
        assert 0 <= d["join"] < np_ and d["cross"] == [0, 1]


def _example(name, tmp_path, np_, args, env=None, extra=()):
    return _launch(np_, [os.path.join(ROOT, "examples", name), "--synthetic", "--engine", "torch", "--width", "0.25",
                         "--train-samples", "512", *args], env=env, extra=extra)


def test_session_example_rank0_only_files_and_resume(tmp_path):
    ck = str(tmp_path / "ck")
    r = _example("tensorflow_mnist.py", tmp_path, 2, ["--num-steps", "24", "--batch-size", "32", "--checkpoint-dir", ck])
    assert r.returncode == 0, r.stderr[-3000:]
    recs = [json.loads(l) for l in r.stdout.splitlines() if l.startswith("{")]
    steps = {(d["rank"], d["step"]) for d in recs if d["event"] == "train_step"}
    assert {(0, 1), (1, 1), (0, 11), (1, 11)} <= steps                  # every rank logs every 10 iterations
    assert max(s for _, s in steps) <= 12                               # 24 // 2 steps
    saved = [d for d in recs if d["event"] == "checkpoint_saved"]
    assert saved and all(d["rank"] == 0 for d in saved) and saved[-1]["step"] == 12
    assert sorted(os.listdir(ck)) == ["ckpt-12.safetensors", "latest"]
    # restart with a bigger budget: rank 0 restores, broadcast syncs rank 1, both continue from 12
    r2 = _example("tensorflow_mnist.py", tmp_path, 2, ["--num-steps", "40", "--batch-size", "32", "--checkpoint-dir", ck])
    assert r2.returncode == 0, r2.stderr[-3000:]
    recs2 = [json.loads(l) for l in r2.stdout.splitlines() if l.startswith("{")]
    assert [d for d in recs2 if d["event"] == "checkpoint_restored"][0]["step"] == 12
    first = {d["rank"]: d["step"] for d in reversed(recs2) if d["event"] == "train_step"}
    assert first == {0: 13, 1: 13}
    assert "ckpt-20.safetensors" in os.listdir(ck)


def test_keras_example_metric_average_and_export(tmp_path):
    args = ["--num-steps", "16", "--batch-size", "32", "--test-samples", "64", "--log-dir", str(tmp_path / "logs"),
            "--checkpoint-dir", str(tmp_path / "ck"), "--export-dir", str(tmp_path / "final_model")]
    r = _example("tensorflow_mnist_gpu.py", tmp_path, 2, args)
    assert r.returncode == 0, r.stderr[-3000:]
    assert "Test loss:" in r.stdout and "Test accuracy:" in r.stdout
    recs = [json.loads(l) for l in r.stdout.splitlines() if l.startswith("{")]
    ep = [d for d in recs if d["event"] == "epoch_end"]
    assert ep and all(d["rank"] == 0 for d in ep)                       # verbose only on rank 0
    assert ep[0]["epochs"] == 2                                          # 16 // (512 // (32*2))
    assert os.path.exists(tmp_path / "final_model" / "model.safetensors")
    assert any(f.startswith("events.out.tfevents") for f in os.listdir(tmp_path / "logs"))
    assert any(f.startswith("mnist-") for f in os.listdir(tmp_path / "ck"))


def test_adasum_example_runs(tmp_path):
    r = _example("tensorflow_mnist.py", tmp_path, 2, ["--num-steps", "8", "--batch-size", "16", "--use-adasum",
                                                      "--checkpoint-dir", str(tmp_path / "ck")])
    assert r.returncode == 0, r.stderr[-3000:]


def test_killed_rank_fails_job_then_restart_resumes(tmp_path):
    ck = str(tmp_path / "ck")
    base = ["--num-steps", "40", "--batch-size", "16", "--checkpoint-dir", ck, "--checkpoint-secs", "0"]
    r = _example("tensorflow_mnist.py", tmp_path, 2, base, env={"B200DP_FAULT": "1:7"}, extra=["--grace-secs", "2"])
    assert r.returncode == 17                                            # the dead rank's code propagates
    assert "terminating" in r.stderr
    assert any(f.startswith("ckpt-") for f in os.listdir(ck))            # rank 0 had been saving
    r2 = _example("tensorflow_mnist.py", tmp_path, 2, base, env={"B200DP_FAULT": "1:7"},
                  extra=["--max-restarts", "1", "--grace-secs", "2"])
    # first attempt restores (>= step 6), dies again only if it passes step 7 at restart 0; either way ends clean
    assert r2.returncode == 0, r2.stderr[-3000:]
    assert "ckpt-20.safetensors" in os.listdir(ck)


def test_elastic_state_resumes_from_last_commit_after_a_rank_dies(tmp_path):
    """hvd.elastic on top of the restart supervisor: rank 1 dies at batch 6 (last commit: 4); the relaunched job syncs the
    persisted commit from rank 0, runs the reset callbacks and finishes; replicas end bitwise identical."""
    r = _launch(2, [os.path.join(W, "elastic_worker.py"), str(tmp_path), "1", "6"], extra=["--max-restarts", "1", "--grace-secs", "2"])
    assert r.returncode == 0, r.stderr[-3000:]
    assert "restart 1/1" in r.stderr
    res = [json.load(open(p)) for p in sorted(glob.glob(str(tmp_path / "r*.json")))]
    assert len(res) == 2
    for d in res:
        assert d["restart"] == 1 and d["first_batch"] == 4 and d["final_batch"] == 12
        assert d["seen"] == list(range(12))                  # batches 4 and 5 were replayed, nothing skipped
        assert d["global_step"] == 12 and d["resets"] == [1]
        assert d["commits"] == 3                             # 4 (persisted before the failure), 8, 12
    assert res[0]["digest"] == res[1]["digest"]
    recs = [json.loads(l) for l in r.stdout.splitlines() if l.startswith("{")]
    assert any(d.get("event") == "elastic_resume" and d.get("commits") == 1 for d in recs)


def test_elastic_state_single_process_commit_restore(tmp_path):
    import torch
    import b200dp as hvd
    lin = torch.nn.Linear(3, 2)
    opt = torch.optim.SGD(lin.parameters(), lr=0.1, momentum=0.9)
    st = hvd.elastic.TorchState(lin, opt, directory=str(tmp_path), epoch=0, best=1e9)
    lin(torch.ones(1, 3)).sum().backward(); opt.step()
    st.epoch, st.best = 2, 0.5
    st.commit()
    w_commit = lin.weight.detach().clone()
    lin(torch.ones(1, 3)).sum().backward(); opt.step()
    st.epoch = 3
    assert not torch.equal(lin.weight, w_commit)
    st.restore()
    assert st.epoch == 2 and st.best == 0.5 and torch.equal(lin.weight, w_commit)
    # a fresh process adopts the persisted commit
    lin2 = torch.nn.Linear(3, 2)
    st2 = hvd.elastic.TorchState(lin2, torch.optim.SGD(lin2.parameters(), lr=0.1, momentum=0.9), directory=str(tmp_path), epoch=0, best=1e9)
    assert st2.load_persisted() and st2.epoch == 2 and st2.commits == 1 and torch.equal(lin2.weight, w_commit)
    # resize request at a commit point
    flag = tmp_path / "resize"
    flag.write_text("4")
    os.environ["B200DP_RESIZE_FILE"] = str(flag)
    try:
        with pytest.raises(hvd.elastic.HostsUpdatedInterrupt):
            st.commit()
        with pytest.raises(SystemExit) as ei:
            hvd.elastic.run(lambda s: s.commit())(st)
        assert ei.value.code == 75
    finally:
        del os.environ["B200DP_RESIZE_FILE"]


@pytest.mark.parametrize("extra_args", [[], ["--fp16-allreduce", "--batches-per-allreduce", "2"], ["--use-adasum"]])
def test_pytorch_example_keeps_replicas_identical(tmp_path, extra_args):
    """Horovod's PyTorch recipe (any nn.Module + torch.optim) over the generic DistributedOptimizer, 2 ranks on gloo."""
    r = _launch(2, [os.path.join(ROOT, "examples", "pytorch_mnist.py"), "--synthetic", "--train-samples", "512", "--test-samples", "128",
                    "--batch-size", "32", "--epochs", "1", *extra_args])
    assert r.returncode == 0, r.stderr[-3000:]
    assert "replicas identical: True" in r.stdout and "Test set: Average loss:" in r.stdout


def test_elastic_resize_relaunches_with_the_new_world_size(tmp_path):
    """A resize request (file naming the new -np) is honoured at the next commit: every rank exits 75, b200run relaunches
    with 4 ranks, the new incarnation syncs the last commit from rank 0 and finishes."""
    flag = tmp_path / "resize"
    flag.write_text("4\n")
    r = _launch(2, [os.path.join(W, "elastic_worker.py"), str(tmp_path), "-1", "0"], extra=["--resize-file", str(flag), "--grace-secs", "2"])
    assert r.returncode == 0, r.stderr[-3000:]
    assert "resizing 2 -> 4 ranks" in r.stderr and not flag.exists()
    res = [json.load(open(p)) for p in sorted(glob.glob(str(tmp_path / "r*.json")))]
    assert len(res) == 4
    for d in res:
        assert d["restart"] == 1 and d["first_batch"] == 4 and d["final_batch"] == 12 and d["global_step"] == 12
        assert d["resets"] == [1] and d["commits"] == 3
    assert len({d["digest"] for d in res}) == 1


def test_keras_example_resumes_from_its_epoch_checkpoint_after_a_restart(tmp_path):
    """`b200run --max-restarts` relaunches the Keras script after a rank died; the relaunch must continue from the newest
    epoch checkpoint instead of quietly starting at epoch 0 (the reference script has no resume path of its own)."""
    args = ["--num-steps", "32", "--batch-size", "32", "--test-samples", "64", "--log-dir", str(tmp_path / "logs"),
            "--checkpoint-dir", str(tmp_path / "ck"), "--export-dir", str(tmp_path / "final_model")]
    # 512 // (32*2) = 8 steps per epoch, 4 epochs; rank 1 dies at step 20 (inside epoch 3), epochs 1-2 were checkpointed
    # whenever val_loss improved (the first epoch always is)
    r = _example("tensorflow_mnist_gpu.py", tmp_path, 2, args, env={"B200DP_FAULT": "1:20"},
                 extra=["--max-restarts", "1", "--grace-secs", "2"])
    assert r.returncode == 0, r.stderr[-3000:]
    assert "restart 1/1" in r.stderr and "resuming from mnist-" in r.stderr
    recs = [json.loads(l) for l in r.stdout.splitlines() if l.startswith("{")]
    ep = [d["epoch"] for d in recs if d["event"] == "epoch_end"]
    assert ep[-1] == 4 and ep.count(1) == 1, ep    # epoch 1 ran once: the relaunch did not start over
    assert os.path.exists(tmp_path / "final_model" / "model.safetensors")
