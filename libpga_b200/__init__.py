"""libpga_b200 -- B200-native genetic-algorithm engine behind libpga's C API.

The product is the C-ABI library (libpga.a / libpga_b200.so, headers in include/).
This package is the thin Python mirror of that API used by the tests and the
benchmark; it holds no compute and has no fallback path.
"""
from .api import (PGA, Population, RANDOM_POPULATION, TOURNAMENT, MAX_POPULATIONS, GENES_F32, GENES_BITS,
                  OBJ_HOOK, OBJ_SUM, OBJ_RASTRIGIN, OBJ_SPHERE, OBJ_ONEMAX, TOPO_RING, TOPO_RANDOM,
                  dist_unique_id, migration_count, tournament_of)
from . import islands

__all__ = ["PGA", "Population", "islands", "dist_unique_id", "migration_count"]
