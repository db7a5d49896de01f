from .modules import ActorCritic, ActorCriticRecurrent, EmpiricalNormalization, Memory, unpad_trajectories
from .ppo import PPO
from .runner import EpisodeStats, OnPolicyRunner, resolve_obs_groups
from .storage import RolloutStorage, split_and_pad_trajectories
from .tensordict import TensorDict
from .vecenv_wrapper import RslRlVecEnvWrapper

__all__ = ["ActorCritic", "ActorCriticRecurrent", "EmpiricalNormalization", "Memory", "split_and_pad_trajectories",
           "unpad_trajectories", "EpisodeStats", "OnPolicyRunner", "PPO", "RolloutStorage", "RslRlVecEnvWrapper",
           "TensorDict", "resolve_obs_groups"]
