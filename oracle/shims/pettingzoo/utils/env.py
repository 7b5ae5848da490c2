class AECEnv: pass
class ParallelEnv: pass
