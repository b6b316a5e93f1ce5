"""Return estimators, the policies around them and the fused PPO loss (mirror of
``tianshou.policy`` for the hot path: BasePolicy <- PGPolicy <- A2CPolicy <- PPOPolicy, BasePolicy <- DQNPolicy)."""
from .a2c import A2CPolicy
from .base import BasePolicy
from .dqn import DQNPolicy
from .pg import PGPolicy, RunningMeanStd
from .ppo import PPOPolicy, ppo_loss

__all__ = ["BasePolicy", "PGPolicy", "A2CPolicy", "PPOPolicy", "DQNPolicy", "RunningMeanStd", "ppo_loss"]
