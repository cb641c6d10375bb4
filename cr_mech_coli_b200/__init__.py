"""cr_mech_coli_b200 — B200-native drop-in for the colony-mechanics path of cr_mech_coli.

Mirrors the names of ``cr_mech_coli`` (reference cr_mech_coli/__init__.py:35-47, src/lib.rs:46-116) for the
hot path: configuration + agent classes, ``generate_agents``, ``run_simulation_with_agents`` ->
``CellContainer``.  Compute lives in ``libcrm_b200.so`` (CUDA, sm_100a) behind ``include/crm_b200.h``."""

from .cell_container import CellContainer, color_to_counter, counter_to_color
from .datatypes import (AgentSettings, CellIdentifier, Configuration, GrowthRateSetter, MiePotentialF32,
                        MorsePotentialF32, PhysicalInteraction, RodAgent, RodMechanicsSettings,
                        SpringLengthThresholdSetter, StorageOption, VoxelPlainIndex)
from .simulation import (SimulationError, generate_agents, generate_positions, params_from_settings,
                         run_simulation_with_agents, run_simulation_with_arrays, run_simulations_batched,
                         sort_cellular_identifiers)

# reference src/lib.rs:38-43,109-111
MINUTE = 1.0
HOUR = 60.0
MICRO_METRE = 1.0

__version__ = "0.1.0"
