"""dqmc_b200 -- B200-native (sm_100a) implementation of the DQMC sweep hot path of kfkq/DQMC.

capi  : ctypes binding of the C ABI (include/dqmc_b200.h -> libdqmc_b200.so)
host  : Python mirror of the reference's AttractiveHubbard / DQMC surface for W batched walkers
"""
from . import capi, host  # noqa: F401

__all__ = ["capi", "host"]
