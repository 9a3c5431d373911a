"""Import-compatible stand-in for the `torchcule` package (NVLabs/cule, branch `bfs`: reference Dockerfile:7-8), which is
not installable here.  Only what the reference imports: `torchcule.atari.Env` and `torchcule.atari.Rom`
(policies/cule_bfs.py:3-4, environments/cule_env.py:7-8, environments/cule_env_multiple.py:4-5).  The emulator behind
them is libstm_b200.so (softtreemax_b200/atari.py)."""
