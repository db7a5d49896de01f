"""CPU oracle: a restatement of the reference's algorithms for the hot path (godk1122/IsaacLab, Isaac Lab 2.1.0 +
rsl_rl 3.0.0), in plain torch-on-CPU / numpy / C.

TEST INFRASTRUCTURE ONLY.  Nothing under `isaaclab_b200/` imports this package; only `tests/`,
`__graft_entry__.smoke()` and `bench.py`'s `cpu_baseline` / `--impl reference` legs may use it, and only as the
checker or as the timed CPU baseline -- never as the product path.

Pinning (see DESIGN.md "Oracle"):
  * rsl_rl (GAE, minibatch generator, PPO update), MDP terms / managers, actuators, contact sensor, math utils:
    pinned against the UNMODIFIED reference Python executed in the authoring container
    (tests/golden/gen_*.py -> tests/golden/*.pt; the reference ships no golden vectors of its own for this path).
  * raycast_mesh: the reference delegates to warp-lang's `wp.mesh_query_ray` (third-party, unpinned version, absent
    here and not installable) -> PARITY UNPINNED for the ray query; the oracle is a brute-force fp64
    Moeller-Trumbore closest-hit following the wrapper's contract (LAB/utils/warp/kernels.py:65-69, ops.py:70).
"""
