"""`from torchcule.atari import Env, Rom` -> the B200 emulator facade."""
from softtreemax_b200.atari import Env, Rom  # noqa: F401
