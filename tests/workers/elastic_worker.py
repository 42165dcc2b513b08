"""Run under `b200run -np N --cpu --max-restarts 1`: an elastic-style loop that dies once and resumes from its last commit."""
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import b200dp as hvd

out_dir, die_rank, die_step = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
hvd.init(device="cpu")
r, n = hvd.rank(), hvd.size()
restart = int(os.environ.get("B200DP_RESTART_COUNT", "0"))
eng = hvd.make_engine(hvd.get_spec("mnist_cnn", 0.25), hvd.DistributedOptimizer(hvd.Adam(1e-3 * n)), engine="torch", seed=3 + r)
resets = []
state = hvd.elastic.EngineState(eng, directory=os.path.join(out_dir, "elastic"), batch=0, seen=[])
state.register_reset_callbacks([lambda: resets.append(restart)])
rng = np.random.RandomState(5)
xs = rng.rand(12, n, 8, 784).astype(np.float32)
ys = rng.randint(0, 10, (12, n, 8)).astype(np.int32)


@hvd.elastic.run
def train(state):
    first = state.batch
    while state.batch < 12:
        eng.train_step(xs[state.batch, r], ys[state.batch, r])
        state.seen.append(state.batch)
        state.batch += 1
        if state.batch % 4 == 0:
            state.commit()
        if restart == 0 and r == die_rank and state.batch == die_step:
            os._exit(23)                                   # a rank dies between commits
    return first


first = train(state)
import hashlib
digest = hashlib.sha256(eng.params.data.numpy().tobytes()).hexdigest()
json.dump({"rank": r, "restart": restart, "first_batch": first, "final_batch": state.batch, "seen": state.seen, "commits": state.commits,
           "global_step": eng.global_step, "resets": resets, "digest": digest}, open(os.path.join(out_dir, f"r{r}.json"), "w"))
hvd.shutdown()
