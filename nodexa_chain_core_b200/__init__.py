"""nodexa_chain_core_b200 — B200-native KawPow / Ethash proof-of-work engine.

Host-side mirror of the reference's interface for this path (names, argument meaning and error
behaviour follow src/crypto/ethash/include/ethash/{ethash.hpp,progpow.hpp} of
DeonDavisV/Nodexa-Chain-Core); every call goes through the C ABI in include/nx_kawpow.h to
hand-written CUDA for sm_100a.  There is no CPU fallback.
"""
from . import equihash, ethash, node, progpow, sharding  # noqa: F401
from .miner import Miner  # noqa: F401
from ._lib import LIB_PATH, NxError, lib  # noqa: F401
from .ethash import (EpochContext, calculate_epoch_seed, calculate_full_dataset_num_items,  # noqa: F401
                     calculate_light_cache_num_items, create_epoch_context, create_epoch_context_full,
                     find_epoch_number, get_epoch_number)

__version__ = "0.1.0"
