"""Run under ``b200run -np N`` on N GPUs.  Real peer-memory path (CUDA IPC over NVLink)."""
import hashlib
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import b200dp as hvd
from b200dp.parallel.collectives import adasum_tree
from b200dp.train.engine import TorchEngine

out_dir, algo = sys.argv[1], sys.argv[2]
arch = "mnist_cnn"
if "@" in algo:                             # "<algo>@<arch>", e.g. gemm_rs@keras128
    algo, arch = algo.split("@")
if algo == "nvls":                         # also route large public broadcasts through the switch (multimem.st)
    os.environ["B200DP_BCAST"] = "nvls"
if algo.endswith("+poison"):               # consumed buffers are overwritten with NaN: a late reader poisons the loss
    os.environ["B200DP_POISON"] = "1"
    algo = algo[:-len("+poison")]
if algo == "gemm_rs":                      # two-shot with the dense wgrad GEMM fused with its reduce-scatter
    os.environ["B200DP_GEMM_RS"] = "1"
    algo = "twoshot"
hvd.init(device="cuda")
r, n = hvd.rank(), hvd.size()
dev = hvd.device()
res = {"rank": r, "size": n, "device": str(dev), "nccl_built": hvd.nccl_built(), "arch": arch}
hvd.world().arena.set_timeout_ms(20000)


def gen(seed, numel):
    return torch.randn(numel, generator=torch.Generator().manual_seed(seed))


# ---- public collectives on CUDA tensors -------------------------------------------------------------
for numel in (5, 1000, 300_001, 5_000_000):
    all_t = [gen(100 + i, numel) for i in range(n)]
    t = all_t[r].to(dev)
    avg = hvd.allreduce(t, op=hvd.Average)
    res[f"avg_{numel}"] = bool(torch.allclose(avg.cpu(), torch.stack(all_t).mean(0), atol=1e-5))
    sm = hvd.allreduce(t, op=hvd.Sum)
    res[f"sum_{numel}"] = bool(torch.allclose(sm.cpu(), torch.stack(all_t).sum(0), atol=1e-4))
    b = torch.full((numel,), float(r), device=dev)
    hvd.broadcast_(b, root_rank=n - 1)
    res[f"bcast_{numel}"] = bool(torch.equal(b.cpu(), torch.full((numel,), float(n - 1))))
ad = [gen(500 + i, 3000) * (i + 1) for i in range(n)]
got = hvd.allreduce(ad[r].to(dev), op=hvd.Adasum)
res["adasum"] = bool(torch.allclose(got.cpu(), adasum_tree(ad), rtol=1e-4, atol=1e-5))
h16 = hvd.allreduce(torch.full((64,), float(r), device=dev, dtype=torch.bfloat16))
res["bf16_avg"] = float(h16[0])
ints = hvd.allreduce(torch.tensor([r + 1], device=dev), op=hvd.Sum)
res["int_sum"] = int(ints[0])

# ---- data-parallel training through the fused allreduce+Adam kernels --------------------------------
B = 32
spec = hvd.get_spec(arch)
use_adasum = algo == "adasum"
if algo == "nvls":
    res["nvls_supported"] = bool(hvd.world().arena.nvls_supported())
    if not res["nvls_supported"]:
        algo = "twoshot"
opt = hvd.DistributedOptimizer(hvd.Adam(1e-3 * (1 if use_adasum else n)), op=hvd.Adasum if use_adasum else hvd.Average,
                               algo="auto" if use_adasum else algo)
eng = hvd.make_engine(spec, opt, engine="b200", batch_size=B, seed=11 + r, dropout=False)   # different init per rank on purpose
eng.broadcast_state(0)
rng = np.random.RandomState(1000)
xs = rng.rand(6, n, B, 784).astype(np.float32)
ys = rng.randint(0, 10, (6, n, B)).astype(np.int32)
losses = []
for s in range(6):
    losses.append(eng.train_step(xs[s, r], ys[s, r]).loss_value())
torch.cuda.synchronize()
res["losses_finite"] = bool(np.isfinite(losses).all())
digest = hashlib.sha256(eng.w16.cpu().view(torch.int16).numpy().tobytes()).hexdigest()
digests = hvd.allgather_object(digest)
res["replicas_bitwise_identical"] = len(set(digests)) == 1
res["sharded_ranges"] = len(eng._sharded_ranges)
res["engine_nvls"] = bool(eng.nvls)
res["nvls_bcast_used"] = bool(getattr(hvd.world().arena, "_staging_mc", False))
res["gemm_rs"] = bool(eng.gemm_rs)
res["dense_fused"] = bool(eng.dense_fused)
res["poison"] = bool(eng.poison)
res["algos"] = [r[2] for r in eng.comm_dense + eng.comm_rest]

# rank 0 alone takes a checkpoint (one-sided gather of sharded Adam state) while the others wait at a barrier
if r == 0:
    st = eng.state()
    if not use_adasum:
        oracle = TorchEngine(spec, hvd.Adam(1e-3 * n), seed=11, device="cuda", dropout=False, compute_dtype=torch.bfloat16)
        for s in range(6):
            oracle.train_step(xs[s].reshape(n * B, 784), ys[s].reshape(n * B))     # one big batch == averaged gradients
        d = (st["tensors"]["params"].cuda() - oracle.params.data)
        init = hvd.models.spec.init_flat(spec, 11, "cuda")
        moved = (oracle.params.data - init)
        cos = torch.nn.functional.cosine_similarity(st["tensors"]["params"].cuda() - init, moved, dim=0).item()
        res["oracle_cos"] = cos
        half = spec.total // 2
        got_mv = st["tensors"]["params"].cuda() - init
        res["oracle_cos_halves"] = [torch.nn.functional.cosine_similarity(got_mv[:half], moved[:half], dim=0).item(),
                                    torch.nn.functional.cosine_similarity(got_mv[half:], moved[half:], dim=0).item()]
        res["oracle_max_diff"] = float(d.abs().max())
        res["adam_m_cos"] = torch.nn.functional.cosine_similarity(st["tensors"]["adam_m"].cuda(), oracle.m, dim=0).item()
    res["ckpt_step"] = st["meta"]["global_step"]
hvd.barrier()
# keep training after the checkpoint: the protocol must still be in step
for s in range(3):
    eng.train_step(xs[s, r], ys[s, r])
torch.cuda.synchronize()
digests = hvd.allgather_object(hashlib.sha256(eng.w16.cpu().view(torch.int16).numpy().tobytes()).hexdigest())
res["still_identical"] = len(set(digests)) == 1
# race stress on the real fabric: hundreds of graph-replayed steps with every barrier arrival randomly delayed, so
# fast ranks run a step ahead of slow ones; the replicas must still agree bit for bit (double-buffered gradients)
hvd.world().arena.set_debug_skew_ns(30000)
for s in range(300):
    eng.train_step(xs[s % 6, r], ys[s % 6, r])
torch.cuda.synchronize()
digests = hvd.allgather_object(hashlib.sha256(eng.w16.cpu().view(torch.int16).numpy().tobytes()).hexdigest())
res["identical_after_skewed_stress"] = len(set(digests)) == 1
res["stress_loss_finite"] = bool(np.isfinite(eng.train_step(xs[0, r], ys[0, r]).loss_value()))
hvd.world().arena.set_debug_skew_ns(0)
# compute delay instead of barrier skew: one rank at a time starts its step 100-300 us late (a spin kernel in front of the
# graph launch), so the others sit in the gather / closing barriers with their buffers of the previous step still live
for s in range(400):
    if r == (s // 20) % n:
        torch.cuda._sleep(int(2e5) * (1 + s % 3))
    eng.train_step(xs[s % 6, r], ys[s % 6, r])
torch.cuda.synchronize()
digests = hvd.allgather_object(hashlib.sha256(eng.w16.cpu().view(torch.int16).numpy().tobytes()).hexdigest())
res["identical_after_compute_delay_stress"] = len(set(digests)) == 1
res["delay_loss_finite"] = bool(np.isfinite(eng.train_step(xs[0, r], ys[0, r]).loss_value()))
# End of run: the other ranks go straight to shutdown while rank 0 dawdles and only then reads the (sharded) state for
# its final save.  shutdown() is collective, so the peers' memory is still mapped and valid when rank 0 pulls from it.
if r == 0:
    import time
    time.sleep(1.5)
    late = eng.state()
    res["late_state_finite"] = bool(torch.isfinite(late["tensors"]["params"]).all() and torch.isfinite(late["tensors"]["adam_v"]).all())
    w_from_master = late["tensors"]["params"].to(torch.bfloat16)
    res["late_state_matches_weights"] = bool(torch.equal(w_from_master, eng.w16.cpu()))
json.dump(res, open(os.path.join(out_dir, f"r{r}.json"), "w"))
hvd.shutdown()
