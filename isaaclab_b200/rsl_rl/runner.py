"""`OnPolicyRunner` with the rsl_rl 3.0.0 surface (RSL/runners/on_policy_runner.py:25-460): same constructor
(`env, train_cfg, log_dir, device`), `learn / save / load / get_inference_policy / train_mode / eval_mode`, the same
checkpoint dictionary and the same `WORLD_SIZE / RANK / LOCAL_RANK` multi-GPU convention.

What differs is where the loop's bookkeeping runs.  The reference syncs the host every env step
(`(dones > 0).nonzero()`, `.cpu().numpy().tolist()`, per-reset `.item()` in `extras["log"]`); here the finished-
episode deques are device rings filled by one kernel per step (`b200lab_episode_stats`), the episode log is
accumulated on the device by the step kernels, and the host reads both once per iteration.  Rollout timing uses
CUDA events, so an iteration has exactly two host syncs: the loss statistics of `PPO.update` and the log read.
"""
from __future__ import annotations

import json
import os
import statistics
import time
import warnings

import torch

from .. import _lib
from .modules import ActorCritic, ActorCriticRecurrent
from .ppo import PPO

_POLICIES = {"ActorCritic": ActorCritic, "ActorCriticRecurrent": ActorCriticRecurrent}
_ALGORITHMS = {"PPO": PPO}
_RING = 100  # deque(maxlen=100), on_policy_runner.py:83-84


def resolve_obs_groups(obs, obs_groups: dict, default_sets: list) -> dict:
    """Observation-set resolution of RSL/utils/utils.py:202-309: 'policy' is mandatory (or defaulted from a group of
    that name), every listed group must exist in `obs`, missing default sets fall back to the same-named group or
    to the policy set."""
    obs_groups = dict(obs_groups or {})
    if "policy" not in obs_groups:
        if "policy" not in obs:
            raise ValueError("The observation configuration dictionary 'obs_groups' must contain the 'policy' key."
                             f" Found keys: {list(obs_groups.keys())}")
        warnings.warn("'obs_groups' has no 'policy' key; using the observation group named 'policy'.")
        obs_groups["policy"] = ["policy"]
    for set_name, groups in obs_groups.items():
        if len(groups) == 0:
            raise ValueError(f"The '{set_name}' key in the 'obs_groups' dictionary can not be an empty list.")
        for group in groups:
            if group not in obs:
                raise ValueError(f"Observation '{group}' in observation set '{set_name}' not found in the observations"
                                 f" from the environment. Available observations from the environment: {list(obs.keys())}")
    for name in default_sets:
        if name not in obs_groups:
            obs_groups[name] = [name] if name in obs else list(obs_groups["policy"])
    return obs_groups


class EpisodeStats:
    """cur_reward_sum / cur_episode_length / rewbuffer / lenbuffer of `learn` (on_policy_runner.py:83-86,113-132) on
    the device."""

    def __init__(self, num_envs: int, device, ring: int = _RING):
        self.N, self.ring = num_envs, ring
        self.cur_reward_sum = torch.zeros(num_envs, dtype=torch.float32, device=device)
        self.cur_episode_length = torch.zeros(num_envs, dtype=torch.float32, device=device)
        # one allocation => one device->host copy per read: [ring rewards | ring lengths]
        self._rings = torch.zeros(2, ring, dtype=torch.float32, device=device)
        self.pushed = torch.zeros(1, dtype=torch.int64, device=device)

    def update(self, rewards: torch.Tensor, dones: torch.Tensor):
        if dones.dtype == torch.bool:
            dones = dones.view(torch.uint8)
        if dones.element_size() not in (1, 8):
            dones = dones.to(torch.long)
        rewards, dones = rewards.contiguous(), dones.contiguous()
        p = _lib.ptr
        _lib.check(_lib.lib().b200lab_episode_stats(p(rewards), p(dones), dones.element_size(), self.N,
                                                    p(self.cur_reward_sum), p(self.cur_episode_length),
                                                    p(self._rings[0]), p(self._rings[1]), self.ring, p(self.pushed),
                                                    _lib.current_stream()))

    def read(self):
        """(rewbuffer, lenbuffer) as lists, oldest first -- the deque contents."""
        n = int(self.pushed.item())
        rings = self._rings.cpu()
        k = min(n, self.ring)
        order = [(n - k + i) % self.ring for i in range(k)]
        return [float(rings[0, j]) for j in order], [float(rings[1, j]) for j in order]


class _JsonlWriter:
    """Minimal scalar writer (`add_scalar(tag, value, step)`), one JSON object per line in `log_dir/scalars.jsonl`.
    Used when tensorboard is not importable."""

    def __init__(self, log_dir):
        os.makedirs(log_dir, exist_ok=True)
        self._f = open(os.path.join(log_dir, "scalars.jsonl"), "a")

    def add_scalar(self, tag, value, step):
        self._f.write(json.dumps({"tag": tag, "value": float(value), "step": int(step)}) + "\n")

    def flush(self):
        self._f.flush()

    def close(self):
        self._f.close()


class OnPolicyRunner:
    def __init__(self, env, train_cfg: dict, log_dir: str | None = None, device="cpu"):
        self.cfg = train_cfg
        self.alg_cfg = dict(train_cfg["algorithm"])
        self.policy_cfg = dict(train_cfg["policy"])
        self.device = device
        self.env = env
        self._configure_multi_gpu()
        self.num_steps_per_env = self.cfg["num_steps_per_env"]
        self.save_interval = self.cfg["save_interval"]
        obs = self.env.get_observations()
        self.cfg["obs_groups"] = resolve_obs_groups(obs, self.cfg.get("obs_groups"), ["critic"])
        self.alg = self._construct_algorithm(obs)
        self.disable_logs = self.is_distributed and self.gpu_global_rank != 0
        self.log_dir = log_dir
        self.writer = None
        self.logger_type = None
        self.tot_timesteps = 0
        self.tot_time = 0.0
        self.current_learning_iteration = 0
        self.git_status_repos = []
        self.stats = EpisodeStats(self.env.num_envs, self.device)
        self.last_log = {}

    # ------------------------------------------------------------------------------------------------
    def learn(self, num_learning_iterations: int, init_at_random_ep_len: bool = False):
        self._prepare_logging_writer()
        if init_at_random_ep_len:  # on_policy_runner.py:66-69
            self.env.episode_length_buf = torch.randint_like(self.env.episode_length_buf,
                                                             high=int(self.env.max_episode_length))
        obs = self.env.get_observations().to(self.device)
        self.train_mode()
        if self.is_distributed:
            print(f"Synchronizing parameters for rank {self.gpu_global_rank}...")
            self.alg.broadcast_parameters()
        on_gpu = torch.device(self.device).type == "cuda"
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)] if on_gpu else None
        start_iter = self.current_learning_iteration
        tot_iter = start_iter + num_learning_iterations
        for it in range(start_iter, tot_iter):
            t0 = time.perf_counter()
            if ev:
                ev[0].record()
            ep_infos = []
            for _ in range(self.num_steps_per_env):
                actions = self.alg.act(obs)
                obs, rewards, dones, extras = self.env.step(actions.to(self.env.device))
                obs, rewards, dones = obs.to(self.device), rewards.to(self.device), dones.to(self.device)
                self.alg.process_env_step(obs, rewards, dones, extras)
                if self.log_dir is not None:
                    # envs without a device-side log hand one over per step, as the reference's do
                    if "episode" in extras:
                        ep_infos.append(extras["episode"])
                    elif extras.get("log"):
                        ep_infos.append(extras["log"])
                self.stats.update(rewards, dones)
            self.alg.compute_returns(obs)
            if ev:
                ev[1].record()
            t1 = time.perf_counter()
            loss_dict = self.alg.update()  # ends with the iteration's first host sync (loss statistics)
            if ev:
                ev[2].record()
                ev[2].synchronize()
                collection_time = ev[0].elapsed_time(ev[1]) * 1e-3
                learn_time = ev[1].elapsed_time(ev[2]) * 1e-3
            else:
                t2 = time.perf_counter()
                collection_time, learn_time = t1 - t0, t2 - t1
            self.current_learning_iteration = it
            if self.log_dir is not None and not self.disable_logs:
                if hasattr(self.env, "episode_log"):
                    dev_log = self.env.episode_log()
                    if dev_log:
                        ep_infos.append(dev_log)
                self.log(dict(it=it, tot_iter=tot_iter, start_iter=start_iter, collection_time=collection_time,
                              learn_time=learn_time, loss_dict=loss_dict, ep_infos=ep_infos,
                              num_learning_iterations=num_learning_iterations))
                if it % self.save_interval == 0:
                    self.save(os.path.join(self.log_dir, f"model_{it}.pt"))
        if self.log_dir is not None and not self.disable_logs:
            self.save(os.path.join(self.log_dir, f"model_{self.current_learning_iteration}.pt"))

    # ------------------------------------------------------------------------------------------------
    def log(self, locs: dict, width: int = 80, pad: int = 35):
        collection_size = self.num_steps_per_env * self.env.num_envs * self.gpu_world_size
        iteration_time = locs["collection_time"] + locs["learn_time"]
        self.tot_timesteps += collection_size
        self.tot_time += iteration_time
        it = locs["it"]
        scalars = {}
        ep_string = ""
        if locs["ep_infos"]:
            for key in locs["ep_infos"][0]:
                vals = []
                for info in locs["ep_infos"]:
                    if key not in info:
                        continue
                    v = info[key]
                    vals.extend(v.detach().reshape(-1).cpu().tolist() if isinstance(v, torch.Tensor) else [float(v)])
                value = statistics.fmean(vals) if vals else float("nan")
                if "/" in key:
                    scalars[key] = value
                    ep_string += f"{f'{key}:':>{pad}} {value:.4f}\n"
                else:
                    scalars["Episode/" + key] = value
                    ep_string += f"{f'Mean episode {key}:':>{pad}} {value:.4f}\n"
        mean_std = float(self.alg.policy.action_std.mean())
        fps = int(collection_size / max(iteration_time, 1e-9))
        for key, value in locs["loss_dict"].items():
            scalars[f"Loss/{key}"] = value
        scalars["Loss/learning_rate"] = self.alg.learning_rate
        scalars["Policy/mean_noise_std"] = mean_std
        scalars["Perf/total_fps"] = fps
        scalars["Perf/collection time"] = locs["collection_time"]
        scalars["Perf/learning_time"] = locs["learn_time"]
        rewbuffer, lenbuffer = self.stats.read()
        if rewbuffer:
            scalars["Train/mean_reward"] = statistics.mean(rewbuffer)
            scalars["Train/mean_episode_length"] = statistics.mean(lenbuffer)
            scalars["Train/mean_reward/time"] = (scalars["Train/mean_reward"], self.tot_time)
            scalars["Train/mean_episode_length/time"] = (scalars["Train/mean_episode_length"], self.tot_time)
        for tag, value in scalars.items():
            if isinstance(value, tuple):
                self.writer.add_scalar(tag, value[0], int(value[1]))
            else:
                self.writer.add_scalar(tag, value, it)
        if hasattr(self.writer, "flush"):
            self.writer.flush()
        self.last_log = {k: (v[0] if isinstance(v, tuple) else v) for k, v in scalars.items()}

        head = f" \033[1m Learning iteration {it}/{locs['tot_iter']} \033[0m "
        lines = ["#" * width, head.center(width, " "), "",
                 f"{'Computation:':>{pad}} {fps:.0f} steps/s (collection: {locs['collection_time']:.3f}s, "
                 f"learning {locs['learn_time']:.3f}s)",
                 f"{'Mean action noise std:':>{pad}} {mean_std:.2f}"]
        for key, value in locs["loss_dict"].items():
            lines.append(f"{f'Mean {key} loss:':>{pad}} {value:.4f}")
        if rewbuffer:
            lines.append(f"{'Mean reward:':>{pad}} {scalars['Train/mean_reward']:.2f}")
            lines.append(f"{'Mean episode length:':>{pad}} {scalars['Train/mean_episode_length']:.2f}")
        done_iters = it + 1 - locs["start_iter"]
        eta = self.tot_time / done_iters * (locs["start_iter"] + locs["num_learning_iterations"] - it)
        tail = ["-" * width, f"{'Total timesteps:':>{pad}} {self.tot_timesteps}",
                f"{'Iteration time:':>{pad}} {iteration_time:.2f}s",
                f"{'Time elapsed:':>{pad}} {time.strftime('%H:%M:%S', time.gmtime(self.tot_time))}",
                f"{'ETA:':>{pad}} {time.strftime('%H:%M:%S', time.gmtime(eta))}"]
        print("\n".join(lines) + "\n" + ep_string + "\n".join(tail) + "\n")

    # ------------------------------------------------------------------------------------------------
    def save(self, path: str, infos=None):
        """Checkpoint dictionary of on_policy_runner.py:289-305 (keys and nested state-dict formats unchanged, so
        files interchange with the reference runner)."""
        torch.save({"model_state_dict": self.alg.policy.state_dict(),
                    "optimizer_state_dict": self.alg.optimizer.state_dict(),
                    "iter": self.current_learning_iteration,
                    "infos": infos}, path)

    def load(self, path: str, load_optimizer: bool = True, map_location: str | None = None):
        loaded = torch.load(path, weights_only=False, map_location=map_location)
        resumed = self.alg.policy.load_state_dict(loaded["model_state_dict"])
        if load_optimizer and resumed:
            self.alg.optimizer.load_state_dict(loaded["optimizer_state_dict"])
        if resumed:
            self.current_learning_iteration = loaded["iter"]
        return loaded["infos"]

    def get_inference_policy(self, device=None):
        self.eval_mode()
        if device is not None:
            self.alg.policy.to(device)
        return self.alg.policy.act_inference

    def train_mode(self):
        self.alg.policy.train()

    def eval_mode(self):
        self.alg.policy.eval()

    def add_git_repo_to_log(self, repo_file_path):
        self.git_status_repos.append(repo_file_path)

    # ------------------------------------------------------------------------------------------------
    def _configure_multi_gpu(self):
        """on_policy_runner.py:353-395: one process per GPU, rank/world from the torchrun environment."""
        self.gpu_world_size = int(os.getenv("WORLD_SIZE", "1"))
        self.is_distributed = self.gpu_world_size > 1
        if not self.is_distributed:
            self.gpu_local_rank = 0
            self.gpu_global_rank = 0
            self.multi_gpu_cfg = None
            return
        self.gpu_local_rank = int(os.getenv("LOCAL_RANK", "0"))
        self.gpu_global_rank = int(os.getenv("RANK", "0"))
        self.multi_gpu_cfg = {"global_rank": self.gpu_global_rank, "local_rank": self.gpu_local_rank,
                              "world_size": self.gpu_world_size}
        if self.device != f"cuda:{self.gpu_local_rank}":
            raise ValueError(f"Device '{self.device}' does not match expected device for local rank "
                             f"'{self.gpu_local_rank}'.")
        if self.gpu_local_rank >= self.gpu_world_size:
            raise ValueError(f"Local rank '{self.gpu_local_rank}' is greater than or equal to world size "
                             f"'{self.gpu_world_size}'.")
        if self.gpu_global_rank >= self.gpu_world_size:
            raise ValueError(f"Global rank '{self.gpu_global_rank}' is greater than or equal to world size "
                             f"'{self.gpu_world_size}'.")
        if not torch.distributed.is_initialized():
            torch.distributed.init_process_group(backend="nccl", rank=self.gpu_global_rank,
                                                 world_size=self.gpu_world_size)
        torch.cuda.set_device(self.gpu_local_rank)

    def _construct_algorithm(self, obs) -> PPO:
        """on_policy_runner.py:397-441.  `class_name` selects from the classes on the accelerated path; an unknown
        name is an error, not a silent fallback."""
        if self.alg_cfg.get("rnd_cfg") is not None or self.alg_cfg.get("symmetry_cfg") is not None:
            raise NotImplementedError("RND / symmetry augmentation are outside the accelerated path")
        if self.cfg.get("empirical_normalization") is not None:
            warnings.warn("`empirical_normalization` is deprecated; set `actor_obs_normalization` / "
                          "`critic_obs_normalization` in the policy configuration.", DeprecationWarning)
            self.policy_cfg.setdefault("actor_obs_normalization", self.cfg["empirical_normalization"])
            self.policy_cfg.setdefault("critic_obs_normalization", self.cfg["empirical_normalization"])
        policy_name = self.policy_cfg.pop("class_name", "ActorCritic")
        if policy_name not in _POLICIES:
            raise NotImplementedError(f"policy class '{policy_name}' is not on the accelerated path "
                                      f"(available: {sorted(_POLICIES)})")
        policy = _POLICIES[policy_name](obs, self.cfg["obs_groups"], self.env.num_actions, device=self.device,
                                        **self.policy_cfg)
        alg_name = self.alg_cfg.pop("class_name", "PPO")
        if alg_name not in _ALGORITHMS:
            raise NotImplementedError(f"algorithm class '{alg_name}' is not on the accelerated path "
                                      f"(available: {sorted(_ALGORITHMS)})")
        alg = _ALGORITHMS[alg_name](policy, device=self.device, **self.alg_cfg, multi_gpu_cfg=self.multi_gpu_cfg)
        alg.init_storage("rl", self.env.num_envs, self.num_steps_per_env, obs, [self.env.num_actions])
        return alg

    def _prepare_logging_writer(self):
        if self.log_dir is None or self.writer is not None or self.disable_logs:
            return
        self.logger_type = str(self.cfg.get("logger", "tensorboard")).lower()
        if self.logger_type in ("neptune", "wandb"):
            raise NotImplementedError("external logging services are out of scope (no network); use 'tensorboard'")
        try:
            from torch.utils.tensorboard import SummaryWriter
            self.writer = SummaryWriter(log_dir=self.log_dir, flush_secs=10)
        except Exception:  # tensorboard not installed in this image
            self.writer = _JsonlWriter(self.log_dir)
