"""b200md -- B200-native short-range MD hot path (bin -> neighbor list -> lj/cut | eam -> nve -> halo)
behind the reference's Pair / Neighbor / Fix / Verlet style API.

The product is the C-ABI library ``libb200md.so`` (include/b200md.h, sources in csrc/); this package is the
thin host side: ``lib`` binds the C ABI with ctypes, ``styles`` mirrors the reference's style classes
(``pair_style lj/cut``, ``pair_style eam``, ``neighbor``/``neigh_modify``, ``fix nve``, ``run``) and
``inputs`` generates the benchmark lattices.  There is no CPU fallback anywhere in this package.
"""
from .build import build, LIB  # noqa: F401
from . import lib, inputs, styles  # noqa: F401
from .styles import B200MD, PairLJCutB200, PairEAMB200, PairEAMAlloyB200, NeighborB200, FixNVEB200, BondFENEB200, FixLangevinB200  # noqa: F401
