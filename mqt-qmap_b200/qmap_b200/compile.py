"""`compile()` - the reference's Python entry point for this path
(python/mqt/qmap/compile.py:59-171, method="heuristic"): same argument names,
defaults and meaning; the mapped circuit and the MappingResults come back as
this package's `Circuit` / `MappingResults` (the reference returns a Qiskit
circuit; Qiskit and mqt-core are not part of this library)."""
from __future__ import annotations

import os

from . import routing
from ._abi import QmapError
from .architecture import Architecture, Properties
from .circuit import Circuit, load_qasm, parse_qasm
from .mapper import HeuristicMapper, MappingResults


def load_architecture(arch) -> Architecture:
    """python/mqt/qmap/load_architecture.py: a built-in name, a .arch file or
    an Architecture."""
    if isinstance(arch, Architecture):
        return arch
    a = Architecture()
    if arch is not None:
        a.load_coupling_map(arch)
    return a


def load_calibration(architecture: Architecture, calibration) -> None:
    """python/mqt/qmap/load_calibration.py: a calibration CSV file or Properties."""
    if calibration is not None:
        architecture.load_properties(calibration)


def compile(  # noqa: A001
    circ,
    arch,
    calibration=None,
    method: str = "heuristic",
    heuristic: str = "gate_count_max_distance",
    initial_layout: str = "dynamic",
    iterative_bidirectional_routing_passes: int | None = None,
    layering: str = "individual_gates",
    automatic_layer_splits_node_limit: int | None = 5000,
    early_termination: str = "none",
    early_termination_limit: int = 0,
    lookahead_heuristic: str | None = "gate_count_max_distance",
    lookaheads: int = 15,
    lookahead_factor: float = 0.5,
    encoding: str = "commander",
    commander_grouping: str = "fixed3",
    swap_reduction: str = "coupling_limit",
    swap_limit: int = 0,
    include_WCNF: bool = False,  # noqa: N803
    use_subsets: bool = True,
    subgraph=None,
    pre_mapping_optimizations: bool = True,
    post_mapping_optimizations: bool = True,
    add_measurements_to_mapped_circuit: bool = True,
    add_barriers_between_layers: bool = False,
    verbose: bool = False,
    debug: bool = False,
    visualizer=None,
    *,
    lib=None,
    backend=None,
) -> tuple[Circuit, MappingResults]:
    """Map `circ` (a Circuit, a .qasm file or OpenQASM 2 text) to `arch` with
    the heuristic mapper.  The arguments of the exact mapper (encoding ...
    subgraph) are accepted for signature compatibility and unused; `lib` /
    `backend` select the loaded C ABI and the routing backend (tests)."""
    del encoding, commander_grouping, swap_reduction, swap_limit, include_WCNF, use_subsets, subgraph
    if arch is None and calibration is None:
        raise ValueError("Either arch or calibration must be specified")
    if str(method) != "heuristic":
        raise QmapError(f"method '{method}' is not part of this library (the heuristic mapper is)")
    architecture = load_architecture(arch)
    load_calibration(architecture, calibration)
    if isinstance(circ, Circuit):
        qc = circ
    elif isinstance(circ, (str, os.PathLike)) and os.path.exists(circ):
        qc = load_qasm(os.fspath(circ))
    elif isinstance(circ, str):
        qc = parse_qasm(circ)
    else:
        raise TypeError(f"cannot load a circuit from {type(circ).__name__}")
    cfg = routing.make_config(
        heuristic=heuristic,
        initial_layout=initial_layout,
        iterative_bidirectional_routing_passes=iterative_bidirectional_routing_passes or 0,
        layering=layering,
        automatic_layer_splits=automatic_layer_splits_node_limit is not None,
        automatic_layer_splits_node_limit=(automatic_layer_splits_node_limit
                                           if automatic_layer_splits_node_limit is not None else 5000),
        early_termination=early_termination,
        early_termination_limit=early_termination_limit,
        lookahead_heuristic="none" if lookahead_heuristic is None else lookahead_heuristic,
        nr_lookaheads=0 if lookahead_heuristic is None else lookaheads,
        lookahead_factor=lookahead_factor,
        pre_mapping_optimizations=pre_mapping_optimizations,
        post_mapping_optimizations=post_mapping_optimizations,
        add_measurements_to_mapped_circuit=add_measurements_to_mapped_circuit,
        add_barriers_between_layers=add_barriers_between_layers,
        verbose=verbose,
        debug=debug,
    )
    mapper = HeuristicMapper(qc, architecture, lib=lib, backend=backend)
    if visualizer is not None and getattr(visualizer, "data_logging_path", None) is not None:
        mapper.data_logging_path = visualizer.data_logging_path
    results = mapper.map(cfg)
    mapped = mapper.mapped_circuit()
    mapped.initial_layout = mapper.initial_layout
    mapped.output_permutation = mapper.output_permutation
    return mapped, results
