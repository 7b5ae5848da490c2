class BaseWrapper:
    def __init__(self, env): self.env = env
