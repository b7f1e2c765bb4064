class BaseCallback:
    def __init__(self, verbose=0):
        self.verbose = verbose
        self.locals = {}
        self.num_timesteps = 0
