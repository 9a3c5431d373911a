"""stable_baselines3.common.callbacks: BaseCallback (what callbacks.py:11 subclasses) and CallbackList."""


class BaseCallback:
    def __init__(self, verbose=0):
        self.model = None
        self.training_env = None
        self.n_calls = 0
        self.num_timesteps = 0
        self.verbose = verbose
        self.locals = {}
        self.globals = {}
        self.logger = None
        self.parent = None

    def init_callback(self, model):
        self.model = model
        self.training_env = model.get_env()
        self.logger = model.logger
        self._init_callback()

    def _init_callback(self):
        pass

    def on_training_start(self, locals_, globals_):
        self.locals, self.globals = locals_, globals_
        self._on_training_start()

    def _on_training_start(self):
        pass

    def on_rollout_start(self):
        self._on_rollout_start()

    def _on_rollout_start(self):
        pass

    def _on_step(self):
        return True

    def on_step(self):
        self.n_calls += 1
        self.num_timesteps = self.model.num_timesteps
        return self._on_step()

    def on_training_end(self):
        self._on_training_end()

    def _on_training_end(self):
        pass

    def on_rollout_end(self):
        self._on_rollout_end()

    def _on_rollout_end(self):
        pass

    def update_locals(self, locals_):
        self.locals.update(locals_)
        self.update_child_locals(locals_)

    def update_child_locals(self, locals_):
        pass


class CallbackList(BaseCallback):
    def __init__(self, callbacks):
        super().__init__()
        self.callbacks = list(callbacks)

    def _init_callback(self):
        for c in self.callbacks:
            c.init_callback(self.model)

    def _on_training_start(self):
        for c in self.callbacks:
            c.on_training_start(self.locals, self.globals)

    def _on_rollout_start(self):
        for c in self.callbacks:
            c.on_rollout_start()

    def _on_step(self):
        ok = True
        for c in self.callbacks:
            ok = (c.on_step() is not False) and ok     # SB3: `callback.on_step() and continue_training`; None counts as True here
        return ok

    def _on_rollout_end(self):
        for c in self.callbacks:
            c.on_rollout_end()

    def _on_training_end(self):
        for c in self.callbacks:
            c.on_training_end()

    def update_child_locals(self, locals_):
        for c in self.callbacks:
            c.update_locals(locals_)
