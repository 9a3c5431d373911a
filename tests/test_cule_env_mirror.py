"""softtreemax_b200.envs.CuleEnv against the behaviours of environments/cule_env.py that VERDICT r01 found missing:
lives forced to 5 after a reset (:67), an all-zero frame (not the no-op's frame) after a life-loss restart (:57-60,69-70),
and the number of no-ops drawn from the GLOBALLY seeded numpy generator (:65; utils.set_seed seeds it from --seed).
Runs on the CPU with the oracle-backed engine double (the env logic is host code)."""
import numpy as np
import pytest
import torch


@pytest.fixture()
def cpu_env_factory(monkeypatch, orc):
    from softtreemax_b200 import atari
    from oracle_engine import OracleEngine
    monkeypatch.setattr(atari, "_ENGINE_FACTORY", OracleEngine)
    monkeypatch.setattr(atari, "_ENGINES", {})
    from softtreemax_b200.envs import CuleEnv
    kw = dict(env_name="BreakoutNoFrameskip-v4", color_mode="gray", repeat_prob=0.0, rescale=True, episodic_life=True, frameskip=4)
    return lambda **k: CuleEnv(device="cpu", env_kwargs=kw, **k)


def test_lives_are_five_after_reset(cpu_env_factory):
    env = cpu_env_factory(fire_reset=False)
    np.random.seed(0)
    env.reset()
    assert int(env.env.lives[0]) == 5 and env.lives == 5


def test_noop_count_follows_the_global_seed(cpu_env_factory):
    env = cpu_env_factory(fire_reset=False)
    frames = {}
    for seed in (1, 2, 1):
        np.random.seed(seed)
        expect = np.random.RandomState(seed).randint(31)
        env.reset()
        frames.setdefault(seed, []).append((expect, int(env.env.blobs[0, 140:144].view(torch.int32)[0])))   # O_FRAME: emulated frames since power-on
    assert frames[1][0] == frames[1][1], "same seed, same number of no-ops"
    for seed, runs in frames.items():
        expect, nframes = runs[0]
        assert nframes == 4 * max(1, expect), f"seed {seed}: {expect} no-op agent steps expected, {nframes} frames emulated"
    assert frames[1][0][1] != frames[2][0][1]


def test_zero_frame_after_a_life_loss_restart(cpu_env_factory):
    env = cpu_env_factory(fire_reset=False)
    np.random.seed(3)
    obs = env.reset()
    lost = False
    for t in range(3000):                                   # never press FIRE-launch: the ball is served and eventually missed
        obs, r, done, info = env.step(1 if t % 50 == 0 else 0)
        if done and env.life_termination:
            lost = True
            break
    assert lost, "no life was lost in 3000 steps"
    before = obs.copy()
    obs2 = env.reset()
    assert obs2.shape == (4, 84, 84)
    assert (obs2[3] == 0).all(), "the frame shown after a life-loss restart is all zero (cule_env.py:57,69-70)"
    assert np.array_equal(obs2[:3], before[1:]), "the older three frames are kept (no buffer reset on a life loss)"
    assert not env.life_termination
