"""One full OnPolicyRunner iteration on a bench workload (bench.train_iteration_record): collection and learning time as
the runner's own log reports them, with the env step issued eagerly and as a CUDA graph.
usage: python tools/train_iter_probe.py [cfg=cfg3] [iterations=5]"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402

cfg_id = sys.argv[1] if len(sys.argv) > 1 else "cfg3"
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 5
dev = torch.device("cuda", 0)
cfg, spec, slabs, init = bench.build_workload(cfg_id, seed=0, device=dev)
print(json.dumps(bench.train_iteration_record(cfg_id, cfg, spec, slabs, init, dev, iters=iters)))
