"""svoc — B200-native batched robust oracle consensus (drop-in for the consensus hot path of
Ophiase/Stochastic-Vector-Oracle-Consensus).  See DESIGN.md and include/svoc.h."""
from ._lib import (CENTRE_DEFAULT, CENTRE_MEAN, CENTRE_MEDIAN, CENTRE_SMOOTH_MEDIAN_CAIRO, STATUS_NAMES,  # noqa: F401
                   TIE_HIGH_INDEX_KEPT, TIE_LOW_INDEX_KEPT, SvocLibraryError, launch_count)
from .engine import (WSAD, ConsensusResult, consensus_fixed, consensus_fixed_host, consensus_float,  # noqa: F401
                     consensus_float_host, expected_reliability, kernel_variant, rank_array)

from .workload import Generated, RunningStats, generate, montecarlo  # noqa: F401
from .replay import EpochReplay, replay_epochs  # noqa: F401
from . import client, contract, distributed, notebook  # noqa: F401,E402

__version__ = "0.1.0"
