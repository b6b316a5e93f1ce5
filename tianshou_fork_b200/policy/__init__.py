"""Return estimators and the fused PPO loss (mirror of ``tianshou.policy`` for the hot path)."""
from .base import BasePolicy
from .ppo import PPOPolicy, RunningMeanStd, ppo_loss

__all__ = ["BasePolicy", "PPOPolicy", "RunningMeanStd", "ppo_loss"]
