import json, sys
sys.path.insert(0, ".")
import torch
import bench
precs = [int(x) for x in sys.argv[1].split(",")] if len(sys.argv) > 1 else [0, 2, 1]
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
for p in precs:
    print(json.dumps(bench.time_ppo_update(torch.device("cuda", 0), precision=p, reps=reps)), flush=True)
