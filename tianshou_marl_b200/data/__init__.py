from .batch import Batch, split_indices
from .buffer import ReplayBuffer, VectorReplayBuffer

__all__ = ["Batch", "split_indices", "ReplayBuffer", "VectorReplayBuffer"]
