"""qmap_b200 - B200-native heuristic router behind QMAP's heuristic-mapper API."""
from ._abi import QmapError  # noqa: F401
from .architecture import Architecture, Properties  # noqa: F401
from .circuit import Circuit, load_qasm, parse_qasm  # noqa: F401
from .routing import make_config  # noqa: F401
from .mapper import HeuristicMapper, MappingResults, verify_mapping  # noqa: F401
from .compile import compile  # noqa: F401,A001,E402
