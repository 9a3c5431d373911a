"""softtreemax_b200 — B200-native tree-expansion hot path of SoftTreeMax behind the reference's
CuleBFS / ActorCriticCnnTSPolicy interface.  Host code is Python (as the reference's is); the
kernels are hand-written sm_100a CUDA reached through the C ABI in include/stm_b200.h."""
__all__ = ["engine", "games"]
