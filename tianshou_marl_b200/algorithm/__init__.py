from .actions import BoxActionMapper
from .ctde import CentralizedCritic, CTDEPolicy, DecentralizedActor, GlobalStateConstructor
from .mappo import MAPPO
from .multiagent import FlexibleMultiAgentPolicyManager
from .ppo import A2C, PPO, AdamOptimizerFactory, ContinuousActorPolicy, DiscreteActorPolicy, Reinforce
from .training_coordinator import SimultaneousTrainer
from .returns import compute_episodic_return, episode_mc_return_to_go, gae_flat, gae_time_major

__all__ = ["compute_episodic_return", "episode_mc_return_to_go", "gae_flat", "gae_time_major", "PPO", "A2C", "Reinforce", "AdamOptimizerFactory",
           "DiscreteActorPolicy", "ContinuousActorPolicy", "FlexibleMultiAgentPolicyManager", "CTDEPolicy", "GlobalStateConstructor",
           "DecentralizedActor", "CentralizedCritic", "SimultaneousTrainer", "MAPPO", "BoxActionMapper"]
