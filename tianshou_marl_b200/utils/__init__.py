from .net import ActorCriticNet
from .stats import A2CTrainingStats, CollectStats, RunningMeanStd, SequenceSummaryStats

__all__ = ["ActorCriticNet", "SequenceSummaryStats", "A2CTrainingStats", "CollectStats", "RunningMeanStd"]
