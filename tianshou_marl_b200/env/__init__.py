from .mpe import SPREAD, TAG, Box, Discrete, EmptyInfo, MPEVectorEnv

__all__ = ["MPEVectorEnv", "Discrete", "Box", "EmptyInfo", "SPREAD", "TAG"]
