"""GPU parity tests proper: the CUDA path, called through the C ABI (softtreemax_b200.engine is a
thin ctypes wrapper over include/stm_b200.h), against the CPU oracle on identical seeded inputs.
Bar: BIT-EXACT (integer / byte / index work; the only float is the reward accumulator, whose
multiply and add are separately rounded on both sides, so it is compared exactly too)."""
import ctypes as C
import os

import numpy as np
import pytest

from conftest import random_rom_game

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")


@pytest.fixture(scope="module")
def engines():
    from softtreemax_b200.engine import Engine
    cache = {}

    def get(name):
        if name not in cache:
            cache[name] = Engine(name)
        return cache[name]
    yield get
    for e in cache.values():
        e.close()


def dev(a, eng):
    return torch.from_numpy(np.ascontiguousarray(a)).to(eng.device)


def oracle_level(orc, g, states, rewards, done, lives, stacks, A, weight, frameskip=4, episodic_life=True):
    """One tree level on the CPU: clone xA, agent step with action i % A, new stacks."""
    n = states.shape[0] * A
    st = np.repeat(states, A, axis=0).copy()
    rw = np.repeat(rewards, A).copy()
    dn = np.repeat(done, A).copy()
    lv = np.repeat(lives, A).copy()
    ob = np.zeros((n, 4, 84, 84), np.uint8)
    for i in range(n):
        d1, r1, l1 = dn[i:i + 1], rw[i:i + 1], lv[i:i + 1]
        obs, *_ = g.agent_step(st[i], i % A, frameskip=frameskip, episodic_life=episodic_life, weight=weight,
                               done=d1, reward=r1, lives=l1)
        ob[i, :3] = stacks[i // A, 1:]
        ob[i, 3] = obs
    return st, rw, dn, lv, ob


@pytest.mark.parametrize("n,fanout", [(1, 4), (33, 3), (257, 4), (1000, 6), (64, 1)])
def test_k1_clone_bitexact(engines, n, fanout):
    eng = engines("synth_breakout")
    rng = np.random.default_rng(n)
    st = rng.integers(0, 256, (n, 256), dtype=np.uint8)
    rw = rng.standard_normal(n).astype(np.float32)
    dn = rng.integers(0, 2, n).astype(np.uint8)
    lv = rng.integers(0, 9, n).astype(np.int32)
    out = eng.expand(dev(st, eng), dev(rw, eng), dev(dn, eng), dev(lv, eng), fanout)
    torch.cuda.synchronize()
    idx = np.arange(n * fanout) // fanout                      # child i <- parent i // fanout (cule_bfs.py:90)
    assert np.array_equal(out[0].cpu().numpy(), st[idx])
    assert np.array_equal(out[1].cpu().numpy(), rw[idx])
    assert np.array_equal(out[2].cpu().numpy(), dn[idx])
    assert np.array_equal(out[3].cpu().numpy(), lv[idx])


@pytest.mark.parametrize("name", ["synth_breakout", "synth_pong", "synth_pacman", "synth_torture"])
@pytest.mark.parametrize("frameskip", [4, 1, 2])
def test_k2_k3_one_level_bitexact(orc, engines, name, frameskip):
    g = orc.Game(name)
    eng = engines(name)
    A = g.n_actions
    R = 37 if A <= 6 else 9                                     # ragged on purpose (not a multiple of 32)
    roots, stacks = orc.make_roots(g, R, seed=5, base_steps=30, spread=23)
    rew0 = np.zeros(R, np.float32)
    done0 = np.zeros(R, np.uint8)
    done0[3] = 1                                                # a frozen env (and its A children)
    lives0 = np.array([g.lives(r) for r in roots], np.int32)
    w = np.float32(0.99) ** 2
    st, rw, dn, lv, ob = oracle_level(orc, g, roots, rew0, done0, lives0, stacks, A, float(w), frameskip)
    d = eng.expand(dev(roots, eng), dev(rew0, eng), dev(done0, eng), dev(lives0, eng), A)
    eng.step(*d, frameskip=frameskip, weight=float(w), render=True)
    obs = eng.frames(dev(stacks, eng), R * A, fanout=A)
    torch.cuda.synchronize()
    assert np.array_equal(d[0].cpu().numpy(), st), "state blobs (RAM + registers)"
    assert np.array_equal(d[1].cpu().numpy(), rw), "rewards"
    assert np.array_equal(d[2].cpu().numpy(), dn), "done"
    assert np.array_equal(d[3].cpu().numpy(), lv), "lives"
    assert np.array_equal(obs.cpu().numpy(), ob), "observation stacks"
    assert ob[3 * A:4 * A, 3].max() == 0                        # frozen envs produce an all-zero new frame


@pytest.mark.parametrize("wpc,scope,deal", [(3, 0, 0), (5, 0, 1), (16, 0, 0), (6, 1, 1), (4, 2, 0)])
def test_k2_block_shapes_bitexact(orc, monkeypatch, wpc, scope, deal):
    """The step kernel under every block shape the host can choose: groups of the scheduler with 1-4 warps each
    (STM_STEP_WPC = 3, 5, 16 — small launches in the other tests run one warp per block), the block-wide and the
    warp-only lock-step.  Ragged launch, a frozen env, blocks whose last warps carry no env: same bits as the oracle."""
    from softtreemax_b200.engine import Engine
    monkeypatch.setenv("STM_STEP_WPC", str(wpc))
    monkeypatch.setenv("STM_STEP_SCOPE", str(scope))
    monkeypatch.setenv("STM_STEP_DEAL", str(deal))           # warps dealt round-robin to the blocks, or consecutive
    g = orc.Game("synth_breakout")
    eng = Engine("synth_breakout")
    try:
        A, R = g.n_actions, 150                                 # 600 envs: 19 warps, the last one ragged
        roots, stacks = orc.make_roots(g, R, seed=9, base_steps=30, spread=23)
        rew0 = np.zeros(R, np.float32)
        done0 = np.zeros(R, np.uint8)
        done0[[3, 64, 65, 66, 67, 68, 69, 70, 71]] = 1           # frozen envs, among them a whole warp of children (64..71 x 4)
        lives0 = np.array([g.lives(r) for r in roots], np.int32)
        st, rw, dn, lv, ob = oracle_level(orc, g, roots, rew0, done0, lives0, stacks, A, 0.5, 4)
        d = eng.expand(dev(roots, eng), dev(rew0, eng), dev(done0, eng), dev(lives0, eng), A)
        eng.step(*d, frameskip=4, weight=0.5, render=True)
        obs = eng.frames(dev(stacks, eng), R * A, fanout=A)
        torch.cuda.synchronize()
        assert np.array_equal(d[0].cpu().numpy(), st), "state blobs"
        assert np.array_equal(d[1].cpu().numpy(), rw) and np.array_equal(d[2].cpu().numpy(), dn) and np.array_equal(d[3].cpu().numpy(), lv)
        assert np.array_equal(obs.cpu().numpy(), ob), "observation stacks"
    finally:
        eng.close()


def _random_rom_parity(orc, seed, translate):
    from softtreemax_b200.engine import Engine
    d, rom = random_rom_game(orc, seed)
    g = orc.Game(desc=orc.make_desc(d, len(rom)), rom=rom)
    eng = Engine((d, rom), translate=translate)          # an ad-hoc descriptor (no games.json entry)
    if translate:
        from softtreemax_b200 import translate as tr
        assert os.path.exists(tr.plugin_path(rom)), "translate.build_plugin left no plug-in"
        assert eng.translated_core > 0, "the library did not load the plug-in"
    else:
        assert eng.translated_core == 0, "an image nobody translated must run on the interpreter"
    try:
        A = g.n_actions
        n = 70
        s0 = g.reset()
        states = np.repeat(s0[None], n, axis=0)
        rng = np.random.default_rng(seed)
        states[:, 0:128] = rng.integers(0, 256, (n, 128))       # random RAM so that envs diverge
        rw = np.zeros(n, np.float32)
        dn = np.zeros(n, np.uint8)
        lv = np.array([g.lives(s) for s in states], np.int32)
        acts = rng.integers(0, A, n).astype(np.uint8)
        ds, dr, dd, dl = dev(states, eng), dev(rw, eng), dev(dn, eng), dev(lv, eng)
        stacks = np.zeros((n, 4, 84, 84), np.uint8)
        for step in range(3):
            eng.step(ds, dr, dd, dl, actions=dev(acts, eng), frameskip=2, weight=0.5, render=True)
            obs = eng.frames(dev(stacks, eng), n, fanout=1)
            new = np.zeros_like(stacks)
            for i in range(n):
                o, *_ = g.agent_step(states[i], int(acts[i]), frameskip=2, weight=0.5, done=dn[i:i + 1],
                                     reward=rw[i:i + 1], lives=lv[i:i + 1])
                new[i, :3] = stacks[i, 1:]
                new[i, 3] = o
            stacks = new
            torch.cuda.synchronize()
            assert np.array_equal(ds.cpu().numpy(), states), f"state, step {step}"
            assert np.array_equal(obs.cpu().numpy(), stacks), f"obs, step {step}"
            assert np.array_equal(dr.cpu().numpy(), rw) and np.array_equal(dd.cpu().numpy(), dn)
            assert np.array_equal(dl.cpu().numpy(), lv)
    finally:
        eng.close()


@pytest.mark.parametrize("seed", range(9))
def test_random_rom_step_bitexact(orc, seed):
    """Cartridges of random bytes: every opcode incl. undocumented ones, wild TIA/RIOT traffic,
    frames that end on the scanline cap, F8 bank switching."""
    _random_rom_parity(orc, seed, translate=False)


@pytest.mark.parametrize("seed", [10, 11])
def test_translation_plugin_bitexact(orc, seed):
    """A cartridge image the library was NOT built with gets its block-translated step kernel at run time: translate.py runs
    tools/xlate_rom.cpp + nvcc into a plug-in, stm_ctx_create loads it by the image's hash, and the translated kernel matches
    the oracle bit for bit on random-byte cartridges (seed 11: 8 KB, F8 bank switching)."""
    _random_rom_parity(orc, seed, translate=True)


@pytest.mark.parametrize("name,R,depth", [("synth_breakout", 5, 1), ("synth_breakout", 3, 2), ("synth_breakout", 2, 3),
                                          ("synth_pong", 2, 2), ("synth_pacman", 3, 2), ("synth_torture", 1, 1),
                                          ("synth_torture", 3, 2)])
def test_bfs_matches_oracle(orc, engines, name, R, depth):
    g = orc.Game(name)
    eng = engines(name)
    roots, stacks = orc.make_roots(g, R, seed=9, base_steps=40, spread=31)
    want = g.bfs(roots, stacks, depth, gamma=0.99, clip_reward=True)
    got = eng.bfs(dev(roots, eng), dev(stacks, eng), depth, gamma=0.99, clip_reward=True, want_states=True)
    torch.cuda.synchronize()
    L = R * g.n_actions ** depth
    assert np.array_equal(got["stacks"].cpu().numpy(), want[0].reshape(L, 4, 84, 84))
    assert np.array_equal(got["rewards"].cpu().numpy(), want[1].reshape(L))
    assert np.array_equal(got["done"].cpu().numpy(), want[2].reshape(L))
    assert np.array_equal(got["lives"].cpu().numpy(), want[3].reshape(L))
    assert np.array_equal(got["states"].cpu().numpy(), want[4].reshape(L, 256))


@pytest.mark.parametrize("mode", ["translated", "interpreter"])
@pytest.mark.parametrize("name,R,depth", [("synth_breakout", 40, 2), ("synth_pong", 9, 2), ("synth_pacman", 5, 2), ("synth_torture", 7, 2)])
def test_both_step_kernels_match_oracle(orc, name, R, depth, mode, monkeypatch):
    """K2 exists twice for a bundled cartridge: the interpreter and the ahead-of-time translated kernel
    (tools/xlate_rom.cpp -> csrc/gen/xl_*.cu, selected by image hash; by default small launches use it and large
    ones the interpreter).  Forced one way and the other, both are bit-exact."""
    from softtreemax_b200.engine import Engine
    if mode == "translated":
        monkeypatch.setenv("STM_XL_MAX_ENVS", "2000000000")
    else:
        monkeypatch.setenv("STM_XLATE", "0")
    g = orc.Game(name)
    eng = Engine(name)
    assert (eng.translated_core != 0) == (mode == "translated")
    try:
        roots, stacks = orc.make_roots(g, R, seed=21, base_steps=50, spread=60)
        want = g.bfs(roots, stacks, depth, gamma=0.99, clip_reward=False)
        got = eng.bfs(dev(roots, eng), dev(stacks, eng), depth, gamma=0.99, clip_reward=False, want_states=True)
        torch.cuda.synchronize()
        L = R * g.n_actions ** depth
        assert np.array_equal(got["stacks"].cpu().numpy(), want[0].reshape(L, 4, 84, 84))
        assert np.array_equal(got["rewards"].cpu().numpy(), want[1].reshape(L))
        assert np.array_equal(got["done"].cpu().numpy(), want[2].reshape(L))
        assert np.array_equal(got["states"].cpu().numpy(), want[4].reshape(L, 256))
    finally:
        eng.close()


def test_bfs_unclipped_rewards_and_chunking(orc, engines):
    """clip_reward=False returns the raw discounted sums; a scratch smaller than a level forces the
    level to be processed in several K2/K3 chunks — results must not change."""
    from softtreemax_b200.engine import Engine
    g = orc.Game("synth_torture")
    roots, stacks = orc.make_roots(g, 2, seed=2, base_steps=25, spread=3)
    want = g.bfs(roots, stacks, 2, gamma=0.9, clip_reward=False)
    eng = Engine("synth_torture")                           # default scratch: 4096 envs
    try:
        got = eng.bfs(dev(roots, eng), dev(stacks, eng), 2, gamma=0.9, clip_reward=False)
        torch.cuda.synchronize()
        assert np.array_equal(got["rewards"].cpu().numpy(), want[1].reshape(-1))
        assert np.abs(want[1]).max() > 1.0                  # the case really has non-trivial rewards
        assert np.array_equal(got["stacks"].cpu().numpy(), want[0].reshape(-1, 4, 84, 84))
    finally:
        eng.close()
    g = orc.Game("synth_breakout")
    roots, stacks = orc.make_roots(g, 300, seed=4)
    eng = engines("synth_breakout")
    a = eng.bfs(dev(roots, eng), dev(stacks, eng), 2)       # 4800 leaves > 4096-env scratch => 2 chunks
    eng.set_scratch_envs(8192)
    b = eng.bfs(dev(roots, eng), dev(stacks, eng), 2)
    torch.cuda.synchronize()
    for k in ("stacks", "rewards", "done", "lives"):
        assert torch.equal(a[k], b[k]), k
    want = g.bfs(roots[:40], stacks[:40], 2)
    assert np.array_equal(a["stacks"][:640].cpu().numpy(), want[0].reshape(-1, 4, 84, 84))


def test_bfs_host_entry_point(orc, engines):
    g = orc.Game("synth_breakout")
    eng = engines("synth_breakout")
    roots, stacks = orc.make_roots(g, 4, seed=1)
    want = g.bfs(roots, stacks, 2)
    got = eng.bfs_host(roots, stacks, 2)
    assert np.array_equal(got["stacks"], want[0].reshape(-1, 4, 84, 84))
    assert np.array_equal(got["rewards"], want[1].reshape(-1))
    assert np.array_equal(got["done"], want[2].reshape(-1))
    assert np.array_equal(got["lives"], want[3].reshape(-1))


@pytest.mark.parametrize("game,n_roots,depth", [("synth_pong", 5, 2), ("synth_breakout", 7, 3), ("synth_pacman", 3, 2), ("synth_breakout", 300, 2)])
def test_bfs_host_matches_device_path(orc, engines, game, n_roots, depth):
    """stm_bfs_host resolves the frames of level 1 after level 2 has been emulated (the root-stack upload hides under both
    launches): same tree as stm_bfs bit for bit, on ragged level sizes, three levels, and when the scratch is too small
    for the two levels side by side (the in-order path)."""
    g = orc.Game(game)
    eng = engines(game)
    roots, stacks = orc.make_roots(g, n_roots, seed=5)
    ref = eng.bfs(dev(roots, eng), dev(stacks, eng), depth)
    torch.cuda.synchronize()
    got = eng.bfs_host(roots, stacks, depth)
    for k in ("stacks", "rewards", "done", "lives"):
        assert np.array_equal(got[k], ref[k].cpu().numpy()), k
    want = g.bfs(roots, stacks, depth)
    assert np.array_equal(got["stacks"], want[0].reshape(-1, 4, 84, 84))
    assert np.array_equal(got["rewards"], want[1].reshape(-1))


def test_bfs_with_root_stacks_arriving_on_another_stream(orc, engines):
    """stm_bfs_after: the root stacks are uploaded on a side stream while the tree is already being emulated; the event
    recorded behind the upload is all the call needs.  Same tree as stm_bfs on stacks that were there from the start."""
    g = orc.Game("synth_pong")
    eng = engines("synth_pong")
    roots, stacks = orc.make_roots(g, 9, seed=31)
    dr = dev(roots, eng)
    ref = eng.bfs(dr, dev(stacks, eng), 2)
    torch.cuda.synchronize()
    pinned = torch.from_numpy(stacks).pin_memory()
    late = torch.full(stacks.shape, 0xEE, dtype=torch.uint8, device=eng.device)     # garbage until the copy lands
    side = torch.cuda.Stream(device=eng.device)
    ready = torch.cuda.Event()
    torch.cuda.synchronize()
    with torch.cuda.stream(side):
        torch.cuda._sleep(20_000_000)                   # ~10 ms: the copy starts well after the tree was enqueued
        late.copy_(pinned, non_blocking=True)
        ready.record(side)
    got = eng.bfs(dr, late, 2, stacks_ready=ready)
    torch.cuda.synchronize()
    for k in ("stacks", "rewards", "done", "lives"):
        assert torch.equal(got[k], ref[k]), k


def test_bfs_replays_from_a_cuda_graph(orc, engines):
    """stm_bfs forks a second stream for the stack shift and joins it before returning: the whole call captures into one CUDA
    graph, and a replay on NEW roots written into the same buffers reproduces the eager tree bit for bit."""
    g = orc.Game("synth_breakout")
    eng = engines("synth_breakout")
    roots_a, stacks_a = orc.make_roots(g, 6, seed=21)
    roots_b, stacks_b = orc.make_roots(g, 6, seed=22)
    dr, ds = dev(roots_a, eng), dev(stacks_a, eng)
    eng.bfs(dr, ds, 2)                                   # pools, streams and events exist before the capture
    torch.cuda.synchronize()
    out = eng.bfs(dr, ds, 2)
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        eng.bfs(dr, ds, 2, out=out)
    for roots, stacks in ((roots_b, stacks_b), (roots_a, stacks_a)):
        dr.copy_(torch.from_numpy(roots)); ds.copy_(torch.from_numpy(stacks))
        graph.replay()
        torch.cuda.synchronize()
        want = g.bfs(roots, stacks, 2)
        assert np.array_equal(out["stacks"].cpu().numpy(), want[0].reshape(-1, 4, 84, 84))
        assert np.array_equal(out["rewards"].cpu().numpy(), want[1].reshape(-1))
        assert np.array_equal(out["done"].cpu().numpy(), want[2].reshape(-1))


def test_full_size_properties(orc, engines):
    """BASELINE size (Breakout depth 3, 1024 roots = 65 536 leaves): too big for the oracle in a test,
    so check size-independent properties: determinism, root-permutation equivariance, the stack-shift
    identity (leaf frames 0..2 are ancestors' frames), and oracle agreement on a sampled sub-tree."""
    g = orc.Game("synth_breakout")
    eng = engines("synth_breakout")
    R, depth, A = 1024, 3, 4
    eng.set_scratch_envs(65536)
    roots, stacks = orc.make_roots(g, R, seed=1234)
    dr, ds = dev(roots, eng), dev(stacks, eng)
    a = eng.bfs(dr, ds, depth, want_states=True)
    b = eng.bfs(dr, ds, depth, want_states=True)
    torch.cuda.synchronize()
    for k in a:
        assert torch.equal(a[k], b[k]), f"non-deterministic {k}"
    perm = torch.randperm(R, generator=torch.Generator().manual_seed(0)).to(eng.device)
    c = eng.bfs(dr[perm].contiguous(), ds[perm].contiguous(), depth, want_states=True)
    L = A ** depth
    for k in a:
        x = a[k].view(R, L, *a[k].shape[1:])[perm]
        assert torch.equal(x.reshape(c[k].shape), c[k]), f"not equivariant under root permutation: {k}"
    # stack-shift identity: frame 0 of every leaf = frame 3 of the root stack (depth 3 shifts three times)
    leaf = a["stacks"].view(R, L, 4, 84, 84)
    assert torch.equal(leaf[:, :, 0], ds[:, 3].unsqueeze(1).expand(R, L, 84, 84))
    # siblings share frames 0..2
    sib = leaf.view(R, L // A, A, 4, 84, 84)
    assert torch.equal(sib[:, :, 0, :3], sib[:, :, 3, :3])
    # oracle on a sample of roots
    pick = [0, 517, 1023]
    want = g.bfs(roots[pick], stacks[pick], depth)
    for j, r in enumerate(pick):
        assert np.array_equal(leaf[r].cpu().numpy(), want[0][j])
        assert np.array_equal(a["rewards"].view(R, L)[r].cpu().numpy(), want[1][j])
        assert np.array_equal(a["states"].view(R, L, 256)[r].cpu().numpy(), want[4][j])


def test_argument_errors(engines):
    from softtreemax_b200._lib import StmError
    eng = engines("synth_breakout")
    st, rw, dn, lv = eng.reset_states(8)
    with pytest.raises(StmError):
        eng.step(st, rw, dn, lv, frameskip=0)
    with pytest.raises(StmError):
        eng.set_scratch_envs(0)
    big = eng.reset_states(eng.max_step_envs + 32)
    with pytest.raises(StmError, match="scratch"):
        eng.step(*big, render=True)


def test_reset_states_match_oracle(orc, engines):
    for name in ("synth_breakout", "synth_pong"):
        eng = engines(name)
        st, *_ = eng.reset_states(3)
        torch.cuda.synchronize()
        want = orc.Game(name).reset()
        assert np.array_equal(st.cpu().numpy(), np.repeat(want[None], 3, 0))


@pytest.mark.parametrize("name,depth,worlds", [("synth_breakout", 3, (1, 2, 3, 5, 8, 20)), ("synth_pong", 2, (2, 4, 7)),
                                               ("synth_pacman", 2, (2, 9, 12))])
def test_leaf_blocks_of_one_tree_equal_the_whole_tree(orc, engines, name, depth, worlds):
    """Leaf-block partition (shard.expand_leaf_block / stm_bfs_resume): the blocks of all ranks, concatenated,
    are the tree stm_bfs expands in one piece — stacks, rewards (clipped and raw), done, lives — also when the
    blocks resume below nodes that are already done (frozen)."""
    from softtreemax_b200 import shard
    g = orc.Game(name)
    eng = engines(name)
    roots, stacks = orc.make_roots(g, 6, seed=33, base_steps=30, spread=400)      # long traces: some roots near game over
    for r in range(6):
        rs, rk = dev(roots[r:r + 1], eng), dev(stacks[r:r + 1], eng)
        for clip in (True, False):
            want = eng.bfs(rs, rk, depth, clip_reward=clip)
            A = eng.n_actions
            for world in worlds:
                parts = [shard.expand_leaf_block(eng, rs[0], rk[0], depth, rank, world, clip_reward=clip) for rank in range(world)]
                assert parts[0]["leaf_lo"] == 0 and parts[-1]["leaf_hi"] == A ** depth
                assert all(parts[i]["leaf_hi"] == parts[i + 1]["leaf_lo"] for i in range(world - 1))
                for key in ("stacks", "rewards", "done", "lives"):
                    got = torch.cat([p[key] for p in parts])
                    assert torch.equal(got, want[key]), (name, r, clip, world, key)
                fa = torch.cat([p["first_action"] for p in parts])
                assert torch.equal(fa.cpu(), torch.arange(A ** depth) // (A ** (depth - 1)))


def test_palette_is_data_of_the_ctx():
    """stm_set_palette: a different colour-index -> grey table changes the observations exactly as the same table does in
    the oracle, without a rebuild (the default table is a formula: a real NTSC table can be dropped in)."""
    import torch
    from oracle import orc
    from softtreemax_b200.engine import Engine
    g = orc.Game("synth_pacman")
    roots, stacks = orc.make_roots(g, 4, seed=9, base_steps=40, spread=8)
    table = np.random.default_rng(2).integers(0, 256, 128).astype(np.uint8)
    table[0] = 0                                           # colour index 0 stays black: undrawn rows are zero-filled
    eng = Engine("synth_pacman")
    try:
        base = eng.bfs(torch.from_numpy(roots).to(eng.device), torch.from_numpy(stacks).to(eng.device), 2)["stacks"].clone()
        eng.set_palette(table)
        orc.set_palette(table)
        out = eng.bfs(torch.from_numpy(roots).to(eng.device), torch.from_numpy(stacks).to(eng.device), 2)
        o_stack, *_ = g.bfs(roots, stacks, 2)
        assert np.array_equal(out["stacks"].cpu().numpy(), o_stack.reshape(-1, 4, 84, 84))
        assert not torch.equal(out["stacks"], base)
    finally:
        orc.set_palette(None)
        eng.close()
