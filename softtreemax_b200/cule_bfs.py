"""CuleBFS — drop-in for policies/cule_bfs.py:10-209 (same constructor, attributes and `bfs`
signature), with the whole expansion done by libstm_b200.so on the GPU.

Differences from the reference, all deliberate:
  - no CPU pool and no crossover: every level runs on the GPU;
  - `bfs` also accepts a batch of roots (`bfs_batch`): the reference expands one root per call;
  - leaf observations are uint8 (the reference promotes them to float32 by `th.cat` of a float
    stack with uint8 frames, cule_bfs.py:172); pass `float_obs=True` for the reference dtype;
  - width-limited trees (`max_width != -1`, cule_bfs.py:81-96,145-150,175-196) run level by level
    from Python (expand K1, step K2, frames K3, node values K4, survivor sampling in torch), since
    every level needs the CNN's opinion before the next one can be expanded.
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
        self.step_env = first.env                                          # the root emulator (cule_bfs.py:43)
        self.eng = self.step_env.eng
        self.device = self.eng.device
        self.min_actions_size = self.eng.n_actions
        self.num_leaves = (self.min_actions_size ** tree_depth if max_width == -1
                           else min(max_width * self.min_actions_size, self.min_actions_size ** tree_depth))   # :32-33
        self.last_survivors = []        # per pruned level: the survivor indices of the last bfs (tests replay them)
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
        if self.max_width != -1:
            return self._bfs_width(root_blob, root_stack[:1], tree_depth)
        out = self.bfs_batch(root_blob, root_stack[:1], tree_depth)
        leaves = out["stacks"].float() if self.float_obs else out["stacks"]
        return leaves, out["rewards"], None

    def _bfs_width(self, root_blob, root_stack, tree_depth, forced_survivors=None):
        """cule_bfs.py:98-209 with max_width != -1: after every level but the last, if more than
        max_width nodes are alive, max_width survivors are drawn without replacement from
        softmax(reward + gamma^(depth+1) * max_a action_logit(node)) (:176-185) and everything is gathered
        down to them (:186-196); `first_action` tracks the root action of every node (:107,147-148,186)."""
        eng, A, W = self.eng, self.min_actions_size, self.max_width
        info = eng.info
        st = root_blob.contiguous().clone()
        n0 = st.shape[0]
        rw = th.zeros(n0, dtype=th.float32, device=self.device)                       # :115
        dn = th.zeros(n0, dtype=th.uint8, device=self.device)                         # :118
        if info["lives_addr"] == 255:
            lv = th.zeros(n0, dtype=th.int32, device=self.device)
        else:
            lv = ((st[:, info["lives_addr"] & 0x7F].to(th.int32) >> info["lives_shift"]) & info["lives_mask"]).contiguous()
        stacks = root_stack.contiguous()
        first_action = th.arange(A, device=self.device).unsqueeze(1)                  # :107
        self.last_survivors = []
        for depth in range(tree_depth):
            st, rw, dn, lv = eng.expand(st, rw, dn, lv, A)                            # new_expand :81-96
            n = st.shape[0]
            if depth != 0:
                first_action = first_action.repeat(1, A).view(-1, 1)                  # :147-148
            eng.step(st, rw, dn, lv, actions=None, frameskip=self.frameskip, episodic_life=self.episodic_life,
                     weight=self.gamma ** depth, render=True)                         # :152-158, action i % A (:149)
            stacks = eng.frames(stacks, n, fanout=A)                                  # :162-172
            if depth != tree_depth - 1 and n > W:                                     # :175
                if forced_survivors is not None:
                    top = th.as_tensor(forced_survivors[len(self.last_survivors)], device=self.device)
                else:
                    vals = self.compute_val_func(stacks)[0].max(dim=1).values         # :176
                    drew = th.sign(rw) if self.clip_reward else rw                    # :178-179
                    pi_logit = drew + self.gamma ** (depth + 1) * vals.to(rw.dtype)   # :180
                    try:
                        top = th.multinomial(th.softmax(th.clip(pi_logit, -1e6, 1e6), 0), W)   # :182
                    except RuntimeError:
                        top = th.argsort(pi_logit, descending=True)[:W]               # :183-185
                self.last_survivors.append(top.tolist())
                first_action = first_action[top]                                      # :186-195
                stacks, st, rw, dn, lv = (stacks[top].contiguous(), st[top].contiguous(), rw[top].contiguous(),
                                          dn[top].contiguous(), lv[top].contiguous())
        rewards = th.sign(rw) if self.clip_reward else rw                             # :204-206
        leaves = stacks.float() if self.float_obs else stacks
        return leaves, rewards, first_action
