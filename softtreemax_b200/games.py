"""Game descriptors + cartridge images (what torchcule.atari.Rom supplies in the reference,
policies/cule_bfs.py:29-31; environments/cule_env.py:18-19)."""
import json
import os

from ._lib import GameDesc

ROM_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "roms")

# reference env ids -> the synthetic cartridge that stands for them (no Atari ROM is redistributable)
ENV_ALIASES = {
    "Breakout": "synth_breakout",
    "Pong": "synth_pong",
    "MsPacman": "synth_pacman",
    # the rest of the reference's sweep (yamls/example_sweep.yml:19-26), by size of the minimal action set
    "Asteroids": "synth_joy14", "KungFuMaster": "synth_joy14", "Gopher": "synth_joy8", "Phoenix": "synth_joy8",
    "Krull": "synth_joy18", "NameThisGame": "synth_pong", "VideoPinball": "synth_pacman",
}


def resolve(env_name):
    """'BreakoutNoFrameskip-v4' -> 'synth_breakout'; names of cartridges pass through."""
    games = json.load(open(os.path.join(ROM_DIR, "games.json")))
    if env_name in games:
        return env_name
    for k, v in ENV_ALIASES.items():
        if k in env_name:
            return v
    raise KeyError(f"no cartridge stands for {env_name!r}; have {sorted(games)} and aliases {sorted(ENV_ALIASES)}")


def load(name):
    """-> (GameDesc, rom bytes, descriptor dict).  `name`: a bundled cartridge / an env_name alias, or a
    (descriptor dict, rom bytes) pair for a user-supplied image (same keys as roms/games.json)."""
    if isinstance(name, tuple):
        g, rom = name
        return make_desc(g, len(rom)), bytes(rom), g
    games = json.load(open(os.path.join(ROM_DIR, "games.json")))
    g = games[resolve(name)]
    rom = open(os.path.join(ROM_DIR, g["file"]), "rb").read()
    return make_desc(g, len(rom)), rom, g


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
