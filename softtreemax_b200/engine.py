"""Engine: the Python face of one stm_ctx (one per GPU).  PyTorch is used for device memory and
streams only; every computation below is a call into libstm_b200.so (include/stm_b200.h).

Tensor conventions (all on cuda:<device>, contiguous):
    states  (n, 256) uint8      rewards (n,) float32     done (n,) uint8     lives (n,) int32
    stacks  (n, 4, 84, 84) uint8
"""
import ctypes as C

import numpy as np
import torch

from . import games
from ._lib import STATE_BYTES, StmError, check, lib


def _ptr(t):
    if t is None:
        return None
    if isinstance(t, int):                       # a raw address, as the reference passes `tensor.data_ptr()`
        return C.c_void_p(t)
    if isinstance(t, np.ndarray):
        assert t.flags["C_CONTIGUOUS"]
        return t.ctypes.data_as(C.c_void_p)
    assert t.is_contiguous(), "stm: tensors must be contiguous"
    return C.c_void_p(t.data_ptr())


class Engine:
    def __init__(self, game, device=0, scratch_envs=None, translate=False):
        """translate: build the block-translated step kernel for this cartridge image first if the library has none
        (translate.build_plugin: g++ + nvcc, about a minute, cached on disk); otherwise such an image is interpreted."""
        if not torch.cuda.is_available():
            raise StmError("softtreemax_b200 needs a CUDA device (sm_100a); there is no CPU fallback")
        self.desc, self.rom, self.info = games.load(game)
        if translate:
            from . import translate as _tr
            _tr.build_plugin(self.rom, self.info.get("bank_scheme", 0))
        self.device = torch.device("cuda", device)
        self._h = C.c_void_p()
        romb = (C.c_uint8 * len(self.rom)).from_buffer_copy(self.rom)
        check(lib().stm_ctx_create(C.byref(self._h), C.byref(self.desc), romb, device))
        self.n_actions = lib().stm_num_actions(self._h)
        self.action_codes = list(self.desc.actions[:self.n_actions])
        if scratch_envs:
            self.set_scratch_envs(scratch_envs)

    def close(self):
        if getattr(self, "_h", None):
            lib().stm_ctx_destroy(self._h)
            self._h = None

    __del__ = close

    # ------------------------------------------------------------------------------------
    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    @property
    def launch_count(self):
        return lib().stm_launch_count(self._h)

    @property
    def translated_core(self):
        """0 = interpreter only; k > 0 = the k-th translated cartridge of the library serves this image (k = their number + 1:
        a plug-in built by translate.py)."""
        return lib().stm_translated_core(self._h)

    @property
    def max_step_envs(self):
        return lib().stm_max_step_envs(self._h)

    def set_scratch_envs(self, n):
        check(lib().stm_set_scratch_envs(self._h, int(n)), self._h)

    def set_palette(self, grey128):
        """Install a colour-index -> grey table (128 uint8 values, e.g. ALE's NTSC palette reduced to luma)."""
        g = np.ascontiguousarray(np.asarray(grey128, dtype=np.uint8).reshape(128))
        check(lib().stm_set_palette(self._h, _ptr(g)), self._h)

    def alloc(self, n):
        d = self.device
        return (torch.empty((n, STATE_BYTES), dtype=torch.uint8, device=d),
                torch.empty((n,), dtype=torch.float32, device=d),
                torch.empty((n,), dtype=torch.uint8, device=d),
                torch.empty((n,), dtype=torch.int32, device=d))

    def reset_states(self, n):
        states, rewards, done, lives = self.alloc(n)
        check(lib().stm_reset_states(self._h, _ptr(states), _ptr(rewards), _ptr(done), _ptr(lives), n,
                                     self._stream()), self._h)
        return states, rewards, done, lives

    def expand(self, states, rewards, done, lives, fanout):
        n = states.shape[0]
        out = self.alloc(n * fanout)
        check(lib().stm_expand(self._h, _ptr(states), _ptr(rewards), _ptr(done), _ptr(lives),
                               _ptr(out[0]), _ptr(out[1]), _ptr(out[2]), _ptr(out[3]), n, fanout,
                               self._stream()), self._h)
        return out

    def step(self, states, rewards, done, lives, actions=None, frameskip=4, episodic_life=True,
             weight=1.0, render=True):
        n = states.shape[0]
        if actions is not None:
            actions = actions.to(torch.uint8).contiguous()
        check(lib().stm_step(self._h, _ptr(states), _ptr(actions), n, frameskip, int(episodic_life),
                             float(weight), _ptr(rewards), _ptr(done), _ptr(lives), int(render),
                             self._stream()), self._h)

    def step_frame(self, states, action_codes, n, episodic_life, weight, rewards, done, lives, frozen=None):
        """One emulated frame + get_data (stm_step_frame).  action_codes / rewards / done / lives may be raw device
        addresses (ints): the torchcule-style calls of the reference hand over `data_ptr()` values."""
        check(lib().stm_step_frame(self._h, _ptr(states), _ptr(action_codes), int(n), int(episodic_life), float(weight),
                                   _ptr(rewards), _ptr(done), _ptr(lives), _ptr(frozen), self._stream()), self._h)

    def frame_obs(self, obs, n):
        """Grey 84x84 observation of the last emulated frame into `obs` (tensor or raw device address), n x 84 x 84 u8."""
        check(lib().stm_frame_obs(self._h, _ptr(obs), int(n), self._stream()), self._h)

    def frames(self, parent_stacks, n, fanout=1, out=None):
        if out is None:
            out = torch.empty((n, 4, 84, 84), dtype=torch.uint8, device=self.device)
        check(lib().stm_frames(self._h, _ptr(parent_stacks), _ptr(out), n, fanout, self._stream()), self._h)
        return out

    def bfs(self, root_states, root_stacks, depth, gamma=0.99, frameskip=4, episodic_life=True,
            clip_reward=True, want_states=False, out=None, stacks_ready=None):
        """stacks_ready: a torch.cuda.Event after which root_stacks is valid (recorded behind its upload on another stream):
        the tree starts at once and only the copy of the older planes into the first level waits (stm_bfs_after)."""
        R = root_states.shape[0]
        L = R * self.n_actions ** depth
        d = self.device
        if out is None:
            out = dict(stacks=torch.empty((L, 4, 84, 84), dtype=torch.uint8, device=d),
                       rewards=torch.empty((L,), dtype=torch.float32, device=d),
                       done=torch.empty((L,), dtype=torch.uint8, device=d),
                       lives=torch.empty((L,), dtype=torch.int32, device=d),
                       states=torch.empty((L, STATE_BYTES), dtype=torch.uint8, device=d) if want_states else None)
        args = (self._h, _ptr(root_states), _ptr(root_stacks), R, depth, float(gamma), frameskip,
                int(episodic_life), int(clip_reward), _ptr(out["stacks"]), _ptr(out["rewards"]),
                _ptr(out["done"]), _ptr(out["lives"]), _ptr(out.get("states")), self._stream())
        if stacks_ready is None:
            check(lib().stm_bfs(*args), self._h)
        else:
            check(lib().stm_bfs_after(*args, C.c_void_p(stacks_ready.cuda_event)), self._h)
        return out

    def bfs_resume(self, node_states, node_stacks, node_rewards, node_done, node_lives, level0, depth_more, gamma=0.99,
                   frameskip=4, episodic_life=True, clip_reward=True, want_states=False):
        """Continue a tree from `n` interior nodes of level `level0` (stm_bfs_resume): the leaf-block partition of
        one tree over several GPUs (shard.py).  node_rewards are the UNCLIPPED accumulated rewards."""
        n = node_states.shape[0]
        L = n * self.n_actions ** depth_more
        d = self.device
        out = dict(stacks=torch.empty((L, 4, 84, 84), dtype=torch.uint8, device=d),
                   rewards=torch.empty((L,), dtype=torch.float32, device=d),
                   done=torch.empty((L,), dtype=torch.uint8, device=d),
                   lives=torch.empty((L,), dtype=torch.int32, device=d),
                   states=torch.empty((L, STATE_BYTES), dtype=torch.uint8, device=d) if want_states else None)
        check(lib().stm_bfs_resume(self._h, _ptr(node_states), _ptr(node_stacks), _ptr(node_rewards), _ptr(node_done),
                                   _ptr(node_lives), n, int(level0), int(depth_more), float(gamma), frameskip,
                                   int(episodic_life), int(clip_reward), _ptr(out["stacks"]), _ptr(out["rewards"]),
                                   _ptr(out["done"]), _ptr(out["lives"]), _ptr(out.get("states")), self._stream()),
              self._h)
        return out

    def bfs_host(self, root_states, root_stacks, depth, gamma=0.99, frameskip=4, episodic_life=True,
                 clip_reward=True, out=None, stacks_on_device=False):
        """Host-buffer entry point: roots come from host memory (numpy arrays or pinned CPU tensors),
        rewards/done/lives go back to host memory; the leaf stacks stay on the GPU when
        stacks_on_device (what the reference's bfs() returns) or are copied to the host too."""
        R = root_states.shape[0]
        L = R * self.n_actions ** depth
        if out is None:
            stacks = (torch.empty((L, 4, 84, 84), dtype=torch.uint8, device=self.device) if stacks_on_device
                      else np.empty((L, 4, 84, 84), np.uint8))
            out = dict(stacks=stacks, rewards=np.empty((L,), np.float32),
                       done=np.empty((L,), np.uint8), lives=np.empty((L,), np.int32))
        on_dev = isinstance(out["stacks"], torch.Tensor) and out["stacks"].is_cuda
        check(lib().stm_bfs_host(self._h, _ptr(root_states), _ptr(root_stacks), R, depth, float(gamma),
                                 frameskip, int(episodic_life), int(clip_reward), _ptr(out["stacks"]), int(on_dev),
                                 _ptr(out["rewards"]), _ptr(out["done"]), _ptr(out["lives"]), self._stream()),
              self._h)
        return out

    # ------------------------------------------------------------------------------------
    PARAM_KEYS = ("cnn.0.weight", "cnn.0.bias", "cnn.2.weight", "cnn.2.bias", "cnn.4.weight", "cnn.4.bias",
                  "linear.0.weight", "linear.0.bias", "action_net.weight", "action_net.bias",
                  "value_net.weight", "value_net.bias")

    def cnn_set_weights(self, params):
        """params: dict of fp32 tensors keyed like PARAM_KEYS (SB3 naming without the
        features_extractor prefix).  Re-tiles them to bf16 on the device."""
        t = [params[k].detach().to(self.device, torch.float32).contiguous() for k in self.PARAM_KEYS]
        A = t[8].shape[0]
        check(lib().stm_cnn_set_weights(self._h, *[_ptr(x) for x in t], A, self._stream()), self._h)
        self._cnn_actions = A

    def cnn_forward(self, obs_u8):
        """obs (n, 4, 84, 84) uint8 on the device -> (action head (n, A), value (n,)) fp32."""
        n = obs_u8.shape[0]
        A = self._cnn_actions
        heads = torch.empty((n, A + 1), dtype=torch.float32, device=self.device)
        check(lib().stm_cnn_forward(self._h, _ptr(obs_u8), n, _ptr(heads), self._stream()), self._h)
        return heads[:, :A], heads[:, A]

    def cnn_forward_train(self, obs_u8):
        """stm_cnn_forward_train: heads (n, A+1) fp32; the activations stay in the ctx for cnn_backward."""
        n = obs_u8.shape[0]
        heads = torch.empty((n, self._cnn_actions + 1), dtype=torch.float32, device=self.device)
        check(lib().stm_cnn_forward_train(self._h, _ptr(obs_u8), n, _ptr(heads), self._stream()), self._h)
        return heads

    def cnn_backward(self, obs_u8, dheads, like):
        """stm_cnn_backward: dL/dheads (n, A+1) -> list of 12 gradient tensors shaped like `like` (PARAM_KEYS order)."""
        grads = [torch.empty_like(p, dtype=torch.float32, device=self.device).contiguous() for p in like]
        check(lib().stm_cnn_backward(self._h, _ptr(obs_u8), obs_u8.shape[0], _ptr(dheads.contiguous().float()),
                                     *[_ptr(g) for g in grads], self._stream()), self._h)
        return grads

    def timing(self, on=True):
        check(lib().stm_timing_enable(self._h, int(on)), self._h)

    def timing_read(self):
        ms = (C.c_float * 4)()
        cnt = (C.c_int * 4)()
        check(lib().stm_timing_read(self._h, ms, cnt), self._h)
        return {k: (float(ms[i]), int(cnt[i])) for i, k in enumerate(("clone", "step", "frames", "shift"))}
