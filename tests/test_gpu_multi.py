"""Real multi-GPU runs through the launcher: CUDA-IPC peer mapping, NVLink collectives, fused allreduce+Adam
(SURVEY §4.3 'Collective (GPU)' / 'End-to-end').  Skipped when the box has fewer GPUs than needed."""
import glob
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
RUN = [sys.executable, os.path.join(ROOT, "bin", "b200run")]
W = os.path.join(ROOT, "tests", "workers", "gpu_multi_worker.py")


def _run(np_, algo, tmp_path):
    env = dict(os.environ, B200DP_LOG_LEVEL="WARNING")
    r = subprocess.run(RUN + ["-np", str(np_), sys.executable, W, str(tmp_path), algo], capture_output=True, text=True,
                       timeout=600, env=env, cwd=ROOT)
    assert r.returncode == 0, (r.stdout[-2000:], r.stderr[-4000:])
    res = [json.load(open(p)) for p in sorted(glob.glob(str(tmp_path / "r*.json")))]
    assert len(res) == np_
    art = os.path.join(ROOT, "gpurun_out")
    if os.path.isdir(art):
        json.dump(res, open(os.path.join(art, f"multi_{np_}_{algo}.json"), "w"), indent=1)
    return res


def _check(res, np_, algo):
    for d in res:
        assert d["nccl_built"] is True
        for k, v in d.items():
            if k.startswith(("avg_", "sum_", "bcast_")) or k in ("adasum", "losses_finite", "replicas_bitwise_identical", "still_identical",
                                                                               "identical_after_skewed_stress", "stress_loss_finite",
                                                                               "identical_after_compute_delay_stress", "delay_loss_finite"):
                assert v is True, (d["rank"], k)
        assert d["bf16_avg"] == pytest.approx((np_ - 1) / 2)
        assert d["int_sum"] == np_ * (np_ + 1) // 2
    r0 = res[0]
    assert r0["ckpt_step"] == 6
    algo = algo.split("@")[0]
    if algo.endswith("+poison"):
        algo = algo[:-len("+poison")]
        assert r0["poison"]
    if algo != "adasum":
        assert r0["oracle_cos"] > 0.9 and r0["adam_m_cos"] > 0.95, r0
    if algo in ("twoshot", "gemm_rs"):
        assert r0["sharded_ranges"] >= 1
    if algo == "gemm_rs":
        assert r0["gemm_rs"] and 3 in r0["algos"], r0           # dense wgrad GEMM fused with its reduce-scatter
    if algo == "auto":
        assert r0["dense_fused"] and r0["sharded_ranges"] == 1, r0     # gather -> wgrad GEMM + Adam epilogue -> weight push
    assert r0["late_state_finite"] and r0["late_state_matches_weights"], r0    # final save after the peers left for shutdown
    if algo == "nvls" and r0.get("nvls_supported"):
        assert r0["engine_nvls"] and 2 in r0["algos"], r0            # the dense bucket really went through the switch
        assert r0["nvls_bcast_used"], r0                                   # and so did the large public broadcasts


@pytest.mark.multigpu(2)
@pytest.mark.parametrize("algo", ["auto", "auto+poison", "twoshot+poison", "oneshot", "twoshot", "adasum", "nvls", "gemm_rs"])
def test_two_gpus(tmp_path, algo):
    _check(_run(2, algo, tmp_path), 2, algo)


@pytest.mark.multigpu(2)
@pytest.mark.parametrize("algo", ["auto@keras128", "oneshot@keras128", "gemm_rs@keras128"])
def test_two_gpus_keras128(tmp_path, algo):
    """The keras128 family through the same fused allreduce+Adam / GEMM+reduce-scatter kernels (F=9216, D=128)."""
    res = _run(2, algo, tmp_path)
    assert res[0]["arch"] == "keras128"
    _check(res, 2, algo)


@pytest.mark.multigpu(4)
@pytest.mark.parametrize("algo", ["auto", "adasum"])
def test_four_gpus(tmp_path, algo):
    _check(_run(4, algo, tmp_path), 4, algo)


@pytest.mark.multigpu(8)
@pytest.mark.parametrize("algo", ["auto", "auto+poison", "twoshot", "adasum", "nvls", "gemm_rs"])
def test_eight_gpus(tmp_path, algo):
    _check(_run(8, algo, tmp_path), 8, algo)
