"""Host-side logic of the row-sharded mode on CPU (gloo, no GPU): handle exchange, row ranges, alpha all-gather.
The solver is a stand-in that owns rows [n*r/P, n*(r+1)/P) and returns alpha = row index + 1 on its own rows."""
import json
import os
import sys

import numpy as np
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as entry  # noqa: E402

n, out_path = int(sys.argv[1]), sys.argv[2]
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
entry.load_package()
from svmb200 import parallel  # noqa: E402


class StandIn:
    def __init__(self):
        self.n = n
        self.connected = None
        self.model = None

    def shard_setup(self, r, w):
        assert (r, w) == (rank, world)
        return bytes([r]) * 64

    def shard_connect(self, w, handles):
        self.connected = [handles[64 * p] for p in range(w)]

    def smo_train(self, **kw):
        a = np.zeros(n)
        r0, r1 = n * rank // world, n * (rank + 1) // world
        a[r0:r1] = np.arange(r0, r1) + 1.0
        return {"alpha": a, "b": 0.5, "num_iter": 7, "smo_ms": 1.0}

    def set_model(self, alpha, b):
        self.model = (alpha.copy(), b)


s = StandIn()
sh = parallel.ShardedSolver(s)
out = sh.smo_train(cache_rows=0)
ok = (s.connected == list(range(world)) and np.array_equal(out["alpha"], np.arange(n) + 1.0)
      and np.array_equal(s.model[0], out["alpha"]) and s.model[1] == 0.5 and sh.row_range(n) == (n * rank // world, n * (rank + 1) // world))
res = [None] * world
dist.all_gather_object(res, bool(ok))
if rank == 0:
    json.dump(res, open(out_path, "w"))
dist.barrier()
dist.destroy_process_group()
