"""ctypes binding of the CPU oracle (oracle/libstm_oracle.so).  TEST INFRASTRUCTURE.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import
this module.  The product package (softtreemax_b200) never does.
"""
import ctypes as C
import json
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(_HERE)
_LIB = None

STATE_BYTES = 256
NATIVE_BYTES = 160 * 210
OBS_BYTES = 84 * 84


class GameDesc(C.Structure):
    _fields_ = [("rom_size", C.c_uint32), ("bank_scheme", C.c_uint8), ("controller", C.c_uint8),
                ("n_actions", C.c_uint8), ("actions", C.c_uint8 * 18), ("score_addr", C.c_uint8 * 3),
                ("score2_addr", C.c_uint8 * 3), ("lives_addr", C.c_uint8), ("lives_shift", C.c_uint8),
                ("lives_mask", C.c_uint8), ("term_addr", C.c_uint8), ("term_mask", C.c_uint8),
                ("term_value", C.c_uint8), ("pad", C.c_uint8 * 3)]


def build(force=False):
    so = os.path.join(_HERE, "libstm_oracle.so")
    src = os.path.join(_HERE, "stm_oracle.c")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-B", "libstm_oracle.so"], stdout=subprocess.DEVNULL)
    return so


def lib():
    global _LIB
    if _LIB is None:
        so = os.path.join(_HERE, "libstm_oracle.so")
        if not os.path.exists(so):
            build()
        L = C.CDLL(so)
        L.orc_game_size.restype = C.c_size_t
        L.orc_game_init.argtypes = [C.c_void_p, C.POINTER(GameDesc), C.c_void_p]
        L.orc_env_reset.argtypes = [C.c_void_p, C.c_void_p]
        L.orc_exec_one.argtypes = [C.c_void_p, C.c_void_p]
        L.orc_frame.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
        L.orc_lives.argtypes = [C.c_void_p, C.c_void_p]
        L.orc_lives.restype = C.c_int
        L.orc_terminal.argtypes = [C.c_void_p, C.c_void_p]
        L.orc_terminal.restype = C.c_int
        L.orc_get_data.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_float, C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_render_obs.argtypes = [C.c_void_p, C.c_void_p]
        L.orc_agent_step.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_float,
                                     C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_bfs.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_float, C.c_int,
                              C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                              C.c_void_p, C.c_int]
        L.orc_bfs.restype = C.c_int
        L.orc_set_palette.argtypes = [C.c_void_p]
        _LIB = L
    return _LIB


def load_game_files(name):
    """(descriptor dict, rom bytes) from softtreemax_b200/roms — plain file reads, no product code."""
    d = os.path.join(_ROOT, "softtreemax_b200", "roms")
    games = json.load(open(os.path.join(d, "games.json")))
    g = games[name]
    rom = open(os.path.join(d, g["file"]), "rb").read()
    return g, rom


def make_desc(g, rom_size):
    d = GameDesc()
    d.rom_size = rom_size
    d.bank_scheme = g["bank_scheme"]
    d.controller = g["controller"]
    d.n_actions = len(g["actions"])
    for i, a in enumerate(g["actions"]):
        d.actions[i] = a
    for i in range(3):
        d.score_addr[i] = g["score_addr"][i]
        d.score2_addr[i] = g["score2_addr"][i]
    d.lives_addr, d.lives_shift, d.lives_mask = g["lives_addr"], g["lives_shift"], g["lives_mask"]
    d.term_addr, d.term_mask, d.term_value = g["term_addr"], g["term_mask"], g["term_value"]
    return d


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


class Game:
    def __init__(self, name=None, desc=None, rom=None):
        L = lib()
        if name is not None:
            g, rom = load_game_files(name)
            desc = make_desc(g, len(rom))
        self.desc = desc
        self.rom = bytes(rom)
        self.n_actions = desc.n_actions
        self._buf = C.create_string_buffer(L.orc_game_size())
        romb = (C.c_uint8 * len(rom)).from_buffer_copy(self.rom)
        if L.orc_game_init(self._buf, C.byref(desc), romb) != 0:
            raise ValueError("bad game descriptor")

    @property
    def h(self):
        return self._buf

    def reset(self):
        s = np.zeros(STATE_BYTES, np.uint8)
        lib().orc_env_reset(self.h, _p(s))
        return s

    def terminal(self, s):
        return bool(lib().orc_terminal(self.h, _p(s)))

    def exec_one(self, s):
        lib().orc_exec_one(self.h, _p(s))

    def frame(self, s, ale_action, fb=None):
        lib().orc_frame(self.h, _p(s), int(ale_action), _p(fb))

    def lives(self, s):
        return lib().orc_lives(self.h, _p(s))

    def agent_step(self, s, action, frameskip=4, episodic_life=True, weight=1.0, done=None,
                   reward=None, lives=None, want_obs=True):
        done = np.zeros(1, np.uint8) if done is None else done
        reward = np.zeros(1, np.float32) if reward is None else reward
        lives = np.array([self.lives(s)], np.int32) if lives is None else lives
        obs = np.zeros(OBS_BYTES, np.uint8) if want_obs else None
        scratch = np.zeros(2 * NATIVE_BYTES, np.uint8)
        lib().orc_agent_step(self.h, _p(s), int(action), frameskip, int(episodic_life), float(weight),
                             _p(done), _p(reward), _p(lives), _p(obs), _p(scratch))
        return (obs.reshape(84, 84) if want_obs else None), float(reward[0]), int(done[0]), int(lives[0])

    def bfs(self, roots, root_stack, depth, gamma=0.99, frameskip=4, episodic_life=True,
            clip_reward=True, want_state=True, nthreads=0):
        roots = np.ascontiguousarray(roots, np.uint8).reshape(-1, STATE_BYTES)
        R = roots.shape[0]
        root_stack = np.ascontiguousarray(root_stack, np.uint8).reshape(R, 4, 84, 84)
        L = self.n_actions ** depth
        stack = np.zeros((R, L, 4, 84, 84), np.uint8)
        rew = np.zeros((R, L), np.float32)
        done = np.zeros((R, L), np.uint8)
        lives = np.zeros((R, L), np.int32)
        st = np.zeros((R, L, STATE_BYTES), np.uint8) if want_state else None
        rc = lib().orc_bfs(self.h, _p(roots), _p(root_stack), R, depth, float(gamma), frameskip,
                           int(episodic_life), int(clip_reward), _p(stack), _p(rew), _p(done),
                           _p(lives), _p(st), nthreads)
        if rc != 0:
            raise MemoryError("orc_bfs")
        return stack, rew, done, lives, st


def set_palette(grey128):
    """Install a colour-index -> grey table in the oracle (process-wide); None restores the bundled default."""
    if grey128 is None:
        import re
        txt = open(os.path.join(_ROOT, "include", "stm_palette.h")).read()
        body = txt[txt.index("STM_GRAY_LUT_INIT {") + len("STM_GRAY_LUT_INIT {"):]
        grey128 = [int(t) for t in re.findall(r"\d+", body[:body.index("}")])]
    g = np.ascontiguousarray(np.asarray(grey128, np.uint8).reshape(128))
    lib().orc_set_palette(_p(g))


def render_obs(fb):
    out = np.zeros(OBS_BYTES, np.uint8)
    lib().orc_render_obs(_p(np.ascontiguousarray(fb, np.uint8)), _p(out))
    return out.reshape(84, 84)


def make_roots(game, R, seed=1234, base_steps=64, spread=192):
    """Roots as BASELINE.md §4 defines them: root i = power-on state advanced by
    base_steps + (i mod spread) agent steps with actions from default_rng(seed)."""
    rng = np.random.default_rng(seed)
    acts = rng.integers(0, game.n_actions, size=base_steps + spread)
    s = game.reset()
    done = np.zeros(1, np.uint8)
    rew = np.zeros(1, np.float32)
    lives = np.array([game.lives(s)], np.int32)
    frames = [np.zeros((84, 84), np.uint8)] * 4
    snaps = {}
    need = sorted(set(base_steps + (i % spread) for i in range(R)))
    t = 0
    for n in need:
        while t < n:
            done[:] = 0      # the root env keeps playing through life losses (CuleEnv resets elsewhere)
            obs, *_ = game.agent_step(s, acts[t], done=done, reward=rew, lives=lives)
            frames = frames[1:] + [obs.copy()]
            t += 1
            if game.terminal(s):     # game over: start a new episode, as the reference's env.reset() would
                s = game.reset()
                frames = [np.zeros((84, 84), np.uint8)] * 4
        snaps[n] = (s.copy(), np.stack(frames))
    roots = np.stack([snaps[base_steps + (i % spread)][0] for i in range(R)])
    stacks = np.stack([snaps[base_steps + (i % spread)][1] for i in range(R)])
    return roots, stacks
