from .modules import ActorCritic
from .ppo import PPO
from .storage import RolloutStorage
from .tensordict import TensorDict

__all__ = ["ActorCritic", "PPO", "RolloutStorage", "TensorDict"]
