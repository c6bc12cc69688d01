"""genogrove_b200 — B200-native batched interval-overlap engine for genogrove's grove intersect path.

The product is the CUDA library ``libgg_b200.so`` (C ABI: include/gg_b200.h).  This package holds its sources
(``csrc/``), a ctypes binding (``index.DeviceIndex``), the Python flavour of the host-side mirror of the
reference's grove API (``grove.Grove``; the C++20 drop-in facade is include/genogrove_b200/grove.hpp), the
synthetic workload generators of BASELINE.json (``synth``), the multi-GPU handle (``group.DeviceGroup``: gg_group_*) and
the host-side sharding arithmetic the tests restate it with (``shard``).
There is no CPU implementation of the path anywhere in this package.
"""
from .index import DeviceIndex, GGError, NO_CHROM, NO_RANK, TIE_EXPLICIT, TIE_INPUT_ORDER, TIE_REVERSE_INPUT, PinnedArray  # noqa: F401
from .grove import FlankingQueryResult, Grove, Interval, Key, QueryResult  # noqa: F401
from .group import DeviceGroup, SHARD_CHROM, SHARD_REPLICATE, shard_assign  # noqa: F401
