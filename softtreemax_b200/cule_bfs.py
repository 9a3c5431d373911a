"""CuleBFS — drop-in for policies/cule_bfs.py:10-209 (same constructor, attributes and `bfs`
signature), with the whole expansion done by libstm_b200.so on the GPU.

Differences from the reference, all deliberate:
  - no CPU pool and no crossover: every level runs on the GPU;
  - `bfs` also accepts a batch of roots (`bfs_batch`): the reference expands one root per call;
  - leaf observations are uint8 (the reference promotes them to float32 by `th.cat` of a float
    stack with uint8 frames, cule_bfs.py:172); pass `float_obs=True` for the reference dtype;
  - width-limited trees (`max_width != -1`, cule_bfs.py:81-96,145-150,175-196) are not built yet
    (SURVEY §8f row 3): requesting one raises NotImplementedError instead of silently differing.
"""
import torch as th


class CuleBFS:
    def __init__(self, step_env, tree_depth, gamma=0.99, compute_val_func=None, max_width=-1, float_obs=False):
        envs = getattr(step_env, "envs", None)
        first = envs[0] if envs is not None else step_env                  # cule_bfs.py:12-19
        self.multiple_envs = envs is not None
        self.env_kwargs = first.env_kwargs
        self.n_frame_stack = first.n_frame_stack
        self.clip_reward = first.clip_reward
        self.compute_val_func = compute_val_func
        self.gamma = gamma
        self.max_depth = tree_depth
        self.max_width = max_width
        if max_width != -1:
            raise NotImplementedError("max_width != -1 (width-limited SoftTreeMax) is not built yet")
        self.step_env = first.env                                          # the root emulator (cule_bfs.py:43)
        self.eng = self.step_env.eng
        self.device = self.eng.device
        self.min_actions_size = self.eng.n_actions
        self.num_leaves = self.min_actions_size ** tree_depth
        self.float_obs = float_obs
        self.episodic_life = self.env_kwargs.get("episodic_life", True)
        self.frameskip = self.env_kwargs.get("frameskip", 4)

    def bfs_batch(self, root_blobs, root_stacks, tree_depth, want_done=False):
        """root_blobs (R, 256) u8, root_stacks (R, 4, 84, 84) u8 -> dict(stacks, rewards, done, lives)."""
        return self.eng.bfs(root_blobs.contiguous(), root_stacks.contiguous(), tree_depth, gamma=self.gamma,
                            frameskip=self.frameskip, episodic_life=self.episodic_life, clip_reward=self.clip_reward)

    def bfs(self, state, tree_depth):
        """state: the root observation stack (1, 4, 84, 84) (float or uint8), as SB3 hands it to
        policy.forward.  -> (leaves_obs (L,4,84,84), rewards (L,), first_action=None)  (cule_bfs.py:209)"""
        root_stack = th.as_tensor(state).to(self.device).reshape(-1, 4, 84, 84)
        if root_stack.dtype != th.uint8:
            root_stack = root_stack.round().clamp_(0, 255).to(th.uint8)
        root_blob = self.step_env.blobs[:1]                                # cule_bfs.py:110-112 (root injection)
        if tree_depth == 0:
            return (root_stack.float() if self.float_obs else root_stack,
                    th.zeros(1, device=self.device), None)
        out = self.bfs_batch(root_blob, root_stack[:1], tree_depth)
        leaves = out["stacks"].float() if self.float_obs else out["stacks"]
        return leaves, out["rewards"], None
