"""cnmf_b200 -- B200-native weighted / masked multiplicative-update NMF
(the hot path of AI-for-Ocean-Science/cnmf)."""
from ._api import Comm, LocalGroup, Plan, init_factors, nmf, prepare_device, shard_rows, solve_virtual_shards  # noqa: F401
from ._lib import CnmfError, launch_count  # noqa: F401
from .extras import NMF, load_result, nmf_restarts, save_result  # noqa: F401

__version__ = "0.1.0"
