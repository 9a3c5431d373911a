"""The reference's OWN files — main.py, policies/cule_bfs.py, policies/actor_critic_ts.py, environments/cule_env.py,
callbacks.py, byte-identical, imported from /root/reference — running on top of the import-compatible stand-ins under
shims/ (torchcule, stable_baselines3, gym) and the torchcule facade of this package.

/root/reference exists only in the build container and the GPU lives only on the GPU box, so HERE the facade is backed by
a CPU test double of the Engine (tests/oracle_engine.py, the oracle): what is checked is the drop-in plumbing — every
attribute, method, raw-pointer convention and return shape the reference touches — and that the reference's frame-by-frame
CuleBFS.bfs, driven through the facade, reproduces the oracle's whole-tree bfs bit for bit.  The same frame-by-frame calls
against the CUDA library are checked on the GPU box by tests/test_facade.py.  Skipped where /root/reference is absent."""
import os
import runpy
import sys

import numpy as np
import pytest
import torch

REF = "/root/reference"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = pytest.mark.skipif(not os.path.isfile(os.path.join(REF, "main.py")), reason="the reference tree is not on this machine")


@pytest.fixture()
def reference_on_shims(monkeypatch, orc, tmp_path):
    from softtreemax_b200 import atari
    from oracle_engine import OracleEngine
    if not torch.cuda.is_available():                       # with a GPU (and the reference) present this is the real thing, end to end
        monkeypatch.setattr(atari, "_ENGINE_FACTORY", OracleEngine)
    monkeypatch.setattr(atari, "_ENGINES", {})
    monkeypatch.syspath_prepend(REF)
    monkeypatch.syspath_prepend(os.path.join(ROOT, "shims"))
    if not torch.cuda.is_available():                       # the reference calls th.cuda.synchronize() unguarded (cule_bfs.py:78,173,202)
        monkeypatch.setattr(torch.cuda, "synchronize", lambda *a, **k: None)
    monkeypatch.setenv("WANDB_MODE", "disabled")
    monkeypatch.chdir(tmp_path)
    mods = [m for m in sys.modules if m.split(".")[0] in ("policies", "environments", "callbacks", "utils", "torchcule",
                                                          "stable_baselines3", "gym", "clip_reward_env")]
    saved = {m: sys.modules.pop(m) for m in mods}
    yield
    for m in [m for m in sys.modules if m.split(".")[0] in ("policies", "environments", "callbacks", "utils", "torchcule",
                                                            "stable_baselines3", "gym", "clip_reward_env")]:
        del sys.modules[m]
    sys.modules.update(saved)


def test_reference_cule_bfs_through_the_facade_matches_the_oracle_tree(reference_on_shims, orc):
    """policies/cule_bfs.py (unmodified) -> facade expand / step / get_data / generate_frames == orc_bfs, depth 2."""
    from environments.cule_env import CuleEnv                      # the reference's files
    from policies.cule_bfs import CuleBFS
    import policies.cule_bfs as ref_bfs
    assert ref_bfs.__file__.startswith(REF)
    kw = dict(env_name="BreakoutNoFrameskip-v4", color_mode="gray", repeat_prob=0.0, rescale=True, episodic_life=True, frameskip=4)
    np.random.seed(5)
    env = CuleEnv(env_kwargs=kw, device=torch.device("cpu"), clip_reward=True, fire_reset=True)
    obs = env.reset()
    for a in (1, 3, 2, 2, 3, 0, 1, 2, 3, 3, 2, 0):
        obs, *_ = env.step(a)
    tree = CuleBFS(env, tree_depth=2, gamma=0.97)
    root_blob = env.env.blobs[:1].clone()
    leaves, rewards, first_action = tree.bfs(torch.from_numpy(obs)[None], 2)
    assert first_action is None and leaves.shape == (16, 4, 84, 84) and rewards.shape == (16,)
    game = orc.Game("synth_breakout")
    stack, rew, *_ = game.bfs(root_blob.cpu().numpy(), obs.astype(np.uint8)[None], 2, gamma=0.97)
    assert np.array_equal(leaves.cpu().numpy().astype(np.uint8), stack[0]), "leaf observation stacks differ"
    assert np.array_equal(rewards.cpu().numpy(), rew[0]), "leaf rewards differ"
    # the root env itself was not disturbed by the expansion (cule_bfs.py:110-112 copies OUT of it)
    assert torch.equal(env.env.blobs[:1], root_blob)


def test_reference_main_py_runs_unmodified(reference_on_shims, capsys):
    """`python main.py --env_name BreakoutNoFrameskip-v4 --tree_depth 1 --total_timesteps 128` with the reference's main.py,
    policies and environments untouched: one PPO iteration (rollout + update + callback + save), then `--run_type evaluate`
    on the saved agent.  (The depth-0 training arm hard-wires "cuda:0", main.py:43: it cannot run in the GPU-less container;
    its env class is exercised directly below.)"""
    old = sys.argv
    try:
        sys.argv = ["main.py", "--env_name", "BreakoutNoFrameskip-v4", "--tree_depth", "1", "--total_timesteps", "128", "--seed", "3"]
        ns = runpy.run_path(os.path.join(REF, "main.py"), run_name="__main__")
    finally:
        sys.argv = old
    out = capsys.readouterr().out
    assert "Num actions: 4" in out and "Saving model in saved_agents/" in out
    saved = os.listdir("saved_agents")
    assert len(saved) == 1
    assert ns["PPO"].__module__.startswith("stable_baselines3")
    try:
        sys.argv = ["main.py", "--env_name", "BreakoutNoFrameskip-v4", "--tree_depth", "1", "--run_type", "evaluate",
                    "--model_filename", os.path.join("saved_agents", saved[0]), "--n_eval_episodes", "1"]
        runpy.run_path(os.path.join(REF, "main.py"), run_name="__main__")
    finally:
        sys.argv = old
    assert "Average episode score:" in capsys.readouterr().out


def test_reference_depth0_env_and_policy_on_the_shims(reference_on_shims):
    """environments/cule_env_multiple.py + policies/actor_critic_depth0.py (unmodified) under the stand-in PPO: the depth-0 arm
    of main.py:42-45,58-59 with the device main.py hard-wires replaced by "cpu"."""
    from environments.cule_env_multiple import CuleEnvMultiple
    from policies.actor_critic_depth0 import ActorCriticCnnPolicyDepth0
    from stable_baselines3 import PPO
    kw = dict(env_name="PongNoFrameskip-v4", color_mode="gray", repeat_prob=0.0, rescale=True, episodic_life=True, frameskip=4)
    env = CuleEnvMultiple(env_kwargs=kw, device="cpu", clip_reward=True, fire_reset=True, n_envs=3)
    model = PPO(policy=ActorCriticCnnPolicyDepth0, env=env, n_steps=8, batch_size=12, n_epochs=2, verbose=0)
    model.learn(total_timesteps=48, log_interval=None)
    assert model.num_timesteps == 48
    assert len(model.policy.gradients_history["action_net.weight"]) == 2 * 2 * 2 - 1    # 2 iterations x 2 epochs x 2 minibatches, minus the first call
