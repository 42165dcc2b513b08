import os, sys, json
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import b200dp as hvd
from b200dp.models.spec import init_flat
algo = sys.argv[1]
hvd.init(device="cuda"); r, n = hvd.rank(), hvd.size()
spec = hvd.get_spec("mnist_cnn")
opt = hvd.DistributedOptimizer(hvd.Adam(1e-3 * n), op=hvd.Average, algo=algo)
eng = hvd.make_engine(spec, opt, engine="b200", batch_size=32, seed=11 + r, dropout=False, use_graph=False)
eng.broadcast_state(0)
rng = np.random.RandomState(1000)
xs = rng.rand(6, n, 32, 784).astype(np.float32); ys = rng.randint(0, 10, (6, n, 32)).astype(np.int32)
losses = [eng.train_step(xs[s, r], ys[s, r]).loss_value() for s in range(6)]
torch.cuda.synchronize()
hvd.barrier()
out = {"rank": r, "algo": algo, "losses": losses, "ranges": [eng.comm_dense, eng.comm_rest], "sharded": eng._sharded_ranges}
pre = eng.master.clone()
if r == 0:
    st = eng.state()
    post = st["tensors"]["params"].cuda()
    init = init_flat(spec, 11, "cuda")
    w16 = eng.w16.float()
    per = {}
    for p in spec.params:
        sl = slice(p.offset, p.offset + p.numel)
        per[p.name] = {"moved_pre": float((pre[sl] - init[sl]).abs().mean()), "moved_post": float((post[sl] - init[sl]).abs().mean()),
                       "w16_vs_post": float((w16[sl] - post[sl]).abs().max()), "w16_vs_pre": float((w16[sl] - pre[sl]).abs().max()),
                       "frac_changed_by_gather": float(((post[sl] - pre[sl]) != 0).float().mean())}
    out["per_param"] = per
hvd.barrier()
json.dump(out, open(f"gpurun_out/dbg_{algo}_{r}.json", "w"), indent=1)
hvd.shutdown()
