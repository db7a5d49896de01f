"""PPO update time of one BASELINE config (default cfg3: Rough-Anymal-C, 4096 envs x 24 steps, [512, 256, 128]) at the given
GEMM precisions:  python tools/ppo_time.py [precisions=2,1] [reps=3] [cfg=cfg3].   B200LAB_GEMM_TMA=0 selects the cp.async
operand path of the tcgen05 GEMM for an A/B against the TMA path."""
import json
import sys

sys.path.insert(0, ".")
import torch

import bench

precs = [int(x) for x in sys.argv[1].split(",")] if len(sys.argv) > 1 else [2, 1]
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
cfg_id = sys.argv[3] if len(sys.argv) > 3 else "cfg3"
from isaaclab_b200.envs.compile import compile_env_cfg
from isaaclab_b200.tasks import TASKS
cc = bench.CONFIGS[cfg_id]
spec = compile_env_cfg(TASKS[cc["task"]](), num_envs=cc["envs"], seed=0)
c = bench.ppo_cfg(cfg_id, spec)
for p in precs:
    r = bench.time_ppo_update(torch.device("cuda", 0), c, precision=p, reps=reps)
    print(json.dumps({k: r[k] for k in ("s_per_iter", "update_ms", "tflops", "precision") if k in r}), flush=True)
