"""stable_baselines3.common.logger.Logger: the reference's callback reads `model.logger.name_to_value`
(callbacks.py:54-55)."""


class Logger:
    def __init__(self, verbose=0):
        self.name_to_value = {}
        self.verbose = verbose

    def record(self, key, value, exclude=None):
        self.name_to_value[key] = value

    def record_mean(self, key, value, exclude=None):
        self.name_to_value[key] = value

    def dump(self, step=0):
        if self.verbose >= 1 and self.name_to_value:
            width = max(len(k) for k in self.name_to_value)
            print("-" * (width + 16))
            for k in sorted(self.name_to_value):
                v = self.name_to_value[k]
                print(f"| {k:<{width}} | {v:<10.4g} |" if isinstance(v, float) else f"| {k:<{width}} | {v!s:<10} |")
            print("-" * (width + 16), flush=True)
        self.name_to_value = {}
