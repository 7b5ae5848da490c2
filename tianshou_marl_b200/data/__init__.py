from .batch import Batch, split_indices
from .buffer import ReplayBuffer, VectorReplayBuffer
from .collector import Collector, EpisodeRolloutHookMCReturn

__all__ = ["Batch", "split_indices", "ReplayBuffer", "VectorReplayBuffer", "Collector", "EpisodeRolloutHookMCReturn"]
