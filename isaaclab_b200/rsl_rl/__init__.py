from .modules import ActorCritic, EmpiricalNormalization
from .ppo import PPO
from .runner import EpisodeStats, OnPolicyRunner, resolve_obs_groups
from .storage import RolloutStorage
from .tensordict import TensorDict
from .vecenv_wrapper import RslRlVecEnvWrapper

__all__ = ["ActorCritic", "EmpiricalNormalization", "EpisodeStats", "OnPolicyRunner", "PPO", "RolloutStorage", "RslRlVecEnvWrapper",
           "TensorDict", "resolve_obs_groups"]
