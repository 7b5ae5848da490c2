from .net import ActorCriticNet
from .stats import A2CTrainingStats, CollectStats, SequenceSummaryStats

__all__ = ["ActorCriticNet", "SequenceSummaryStats", "A2CTrainingStats", "CollectStats"]
