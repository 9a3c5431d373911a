"""Host-side mirror of the `torchcule.atari` surface that the reference touches
(policies/cule_bfs.py:3-4,29-31,55-62,71-96,110-118,143,154-167,188-196; environments/cule_env.py:7-8,18-22,
60-71,76-93; environments/cule_env_multiple.py:21-23,48-58), backed by libstm_b200.so.

`shims/torchcule/atari.py` re-exports `Env` and `Rom` from here, so the reference's own `policies/cule_bfs.py`
and `environments/*.py` run unmodified on top of this module.

Everything lives on the GPU: there is no CPU back-end.  An `Env(device="cpu")` — what the reference asks for below
its crossover depth (cule_bfs.py:39, cule_env.py:20) — is given the GPU as well and reports `device = cuda:<n>`,
so the reference's CPU->GPU hand-off (`copy_envs`, cule_bfs.py:71-79) degenerates into a device-to-device copy.

Two ways in:
  * the fused calls (`Env.step`, `CuleBFS.bfs` of this package): one K2 launch per agent step, K3 for the stack;
  * the raw-pointer, frame-by-frame calls the reference's `CuleBFS.bfs` makes — base-class `step(...)`, `get_data(...)`,
    `generate_frames(...)`, `expand(n)` — which map to stm_step_frame / stm_frame_obs / stm_expand one to one and give
    bit-identical results (tests/test_facade.py).
"""
import torch

from . import games
from .engine import Engine

_ENGINES = {}
_ENGINE_FACTORY = Engine          # tests substitute a CPU double here (tests/oracle_engine.py); the product never does


def engine_for(env_name, device=0):
    """One Engine (one stm_ctx) per (cartridge, device), shared by every Env / CuleBFS on it."""
    key = (games.resolve(env_name), int(device), _ENGINE_FACTORY)
    if key not in _ENGINES:
        _ENGINES[key] = _ENGINE_FACTORY(key[0], device=key[1])
    return _ENGINES[key]


class Rom:
    """torchcule.atari.Rom: `Rom(env_name).minimal_actions()` (cule_bfs.py:29-30)."""

    def __init__(self, env_name):
        self.desc, self.rom, self.info = games.load(env_name)

    def minimal_actions(self):
        return torch.tensor(self.info["actions"], dtype=torch.uint8)


class _Space:
    def __init__(self, n=None, shape=None):
        self.n, self.shape = n, shape


class _EnvBase:
    """What `super(AtariEnv, env)` reaches in the reference: the extension's own base class, whose `reset` and `step`
    work on raw pointers, one emulated frame at a time (cule_bfs.py:57,154-155; cule_env.py:21)."""

    def reset(self, _seed=0):
        """Power-on state in every env (cule_bfs.py:57 `super(AtariEnv, env).reset(0)`)."""
        n = self.num_envs
        self.blobs[:n], self.rewards[:n], self.done[:n], self.lives[:n] = self.eng.reset_states(n)
        self._frame_idx = 0

    def step(self, fire_reset, _unused_flag, actions_ptr, _unused_a, done_ptr, _unused_b):
        """One emulated frame for the first size() envs; `actions_ptr` = address of size() ALE action codes
        (cule_bfs.py:154-155).  The emulation is fused with the reward/done/lives update, so the launch happens
        in the `get_data` call that always follows (cule_bfs.py:157-158); here the arguments are kept.  The done
        flags of the first frame of an agent step decide which envs sit the whole step out."""
        if self._frame_idx % self.frameskip == 0:
            self._frozen = self.done[:self._n].clone()
        self._frame_idx += 1
        self._pending = (int(actions_ptr), int(done_ptr))


class Env(_EnvBase):
    """torchcule.atari.Env for the attributes and methods the reference uses.

    Tensors (all on the GPU): `blobs` (N, 256) u8 is the storage; `ram` = blobs[:, :128] and
    `states` = blobs[:, 128:] are views of it (`frame_states` is zero-width: the TIA needs no separate
    render state), so the reference's root injection `pool.states[0] = env.states[0]` etc.
    (cule_bfs.py:110-112) keeps working as plain tensor assignment.
    """

    def __init__(self, num_envs=1, device="cuda:0", action_set=None, env_name="BreakoutNoFrameskip-v4",
                 color_mode="gray", repeat_prob=0.0, rescale=True, episodic_life=True, frameskip=4, **_):
        if color_mode != "gray" or not rescale or repeat_prob != 0.0:
            raise NotImplementedError("only color_mode='gray', rescale=True, repeat_prob=0.0 (what main.py:37-38 uses)")
        dev = torch.device(device)
        self.eng = engine_for(env_name, (dev.index or 0) if dev.type == "cuda" else 0)     # "cpu" -> the GPU (module docstring)
        self.device, self.is_cuda = self.eng.device, self.eng.device.type == "cuda"
        self.env_name, self.num_envs = env_name, num_envs
        self.episodic_life, self.frameskip, self.rescale, self.num_channels = episodic_life, frameskip, rescale, 1
        self.fire_reset, self.is_training = False, True
        self.action_set = torch.tensor(self.eng.action_codes, dtype=torch.uint8, device=self.device)
        self.action_space = _Space(n=self.eng.n_actions)
        self.observation_space = _Space(shape=(1, 84, 84))
        self.blobs, self.rewards, self.done, self.lives = self.eng.reset_states(num_envs)
        self.stacks = torch.zeros((num_envs, 4, 84, 84), dtype=torch.uint8, device=self.device)
        self.observations1 = torch.zeros((num_envs, 84, 84, 1), dtype=torch.uint8, device=self.device)
        self.observations2 = torch.zeros((num_envs, 84, 84, 1), dtype=torch.uint8, device=self.device)
        self._n = num_envs
        self._frame_idx, self._frozen, self._pending = 0, None, None
        if num_envs > self.eng.max_step_envs:
            self.eng.set_scratch_envs(num_envs)

    # views the reference indexes directly
    @property
    def ram(self):
        return self.blobs[:, :128]

    @property
    def states(self):
        return self.blobs[:, 128:]

    @property
    def frame_states(self):
        return self.blobs[:, 256:]

    def size(self):
        return self._n

    def set_size(self, n):
        assert n <= self.num_envs
        self._n = n
        self._frame_idx = 0

    def sync_other_stream(self):      # everything runs on the caller's stream: nothing to fence
        pass

    sync_this_stream = update_frame_states = sync_other_stream

    def train(self):
        self.is_training = True

    def eval(self):
        self.is_training = False

    # ---- the raw-pointer calls of the reference's CuleBFS.bfs --------------------------------------------------------
    def expand(self, num_envs):
        """Every one of the first size() envs becomes num_envs / size() consecutive children, in place (K1):
        child i <- parent i // fanout (cule_bfs.py:143; the same rule as new_expand's index, :90)."""
        n0 = self._n
        assert num_envs <= self.num_envs and num_envs % n0 == 0
        fanout = num_envs // n0
        if n0 == 1:
            # The reference injects the root's states/ram into row 0 but not its lives (cule_bfs.py:110-112), so the pool's
            # stale count would make episodic_life flag every child "life lost" at once.  As stm_bfs does, the
            # root's lives are read from its RAM when the tree is opened.
            info = self.eng.info
            if info["lives_addr"] != 255:
                self.lives[:1] = (self.blobs[:1, info["lives_addr"] & 0x7F].to(torch.int32) >> info["lives_shift"]) & info["lives_mask"]
        if fanout > 1:
            st, rw, dn, lv = self.eng.expand(self.blobs[:n0], self.rewards[:n0], self.done[:n0], self.lives[:n0], fanout)
            self.blobs[:num_envs], self.rewards[:num_envs], self.done[:num_envs], self.lives[:num_envs] = st, rw, dn, lv
        self._n = num_envs
        self._frame_idx = 0

    def get_data(self, episodic_life, weight, done_ptr, rewards_ptr, lives_ptr):
        """`rewards += weight * r_frame`, done and lives of the frame the preceding base-class step() asked for
        (cule_bfs.py:157-158).  Launches K2 for that one frame (stm_step_frame)."""
        if self._pending is None:
            raise RuntimeError("get_data() without a preceding step()")
        actions_ptr, _ = self._pending
        self._pending = None
        self.eng.step_frame(self.blobs, actions_ptr, self._n, bool(episodic_life), float(weight), int(rewards_ptr),
                            int(done_ptr), int(lives_ptr), self._frozen)

    def generate_frames(self, rescale, last_frame, num_channels, obs_ptr):
        """Grey 84x84 observation of the frame just emulated into size() x 84 x 84 x 1 bytes at obs_ptr
        (cule_bfs.py:162-167); K3 in its single-frame form (stm_frame_obs)."""
        if not rescale or num_channels != 1:
            raise NotImplementedError("rescale=True, num_channels=1 (what main.py:37-38 selects)")
        self.eng.frame_obs(int(obs_ptr), self._n)

    # ---- the tensor-level calls of the env wrappers --------------------------------------------------------------------
    def reset(self, initial_steps=0, verbose=False, **_):
        """Power-on state, then `initial_steps` NOOP agent steps; returns obs (N, 84, 84, 1) u8."""
        self.blobs, self.rewards, self.done, self.lives = self.eng.reset_states(self.num_envs)
        self.stacks.zero_()
        self._frame_idx = 0
        noop = torch.zeros(self.num_envs, dtype=torch.uint8, device=self.device)
        for _ in range(max(1, int(initial_steps))):
            self._raw_step(noop)
        return self.stacks[:, 3].unsqueeze(-1)

    def _raw_step(self, actions):
        self.done.zero_()
        self.rewards.zero_()
        self.eng.step(self.blobs, self.rewards, self.done, self.lives, actions=actions, frameskip=self.frameskip,
                      episodic_life=self.episodic_life, weight=1.0, render=True)
        self.stacks = self.eng.frames(self.stacks, self.num_envs, fanout=1)
        self._frame_idx = 0

    def step(self, actions):
        """actions: policy-action indices.  -> (obs (N,84,84,1) u8, reward (N,) f32, done (N,) bool, info)"""
        a = torch.as_tensor(actions).to(self.device, torch.uint8).reshape(-1)
        self._raw_step(a)
        return (self.stacks[:, 3].unsqueeze(-1), self.rewards.clone(), self.done.bool(),
                {"ale.lives": self.lives.clone()})
