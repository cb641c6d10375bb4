"""Import shim: ``import cr_mech_coli as crm`` resolves to the B200-native drop-in.

Callers written against the reference (``crm.Configuration``, ``crm.AgentSettings``, ``crm.generate_agents``,
``crm.RodAgent(pos=..., vel=..., **settings.to_rod_agent_dict())``, ``crm.run_simulation_with_agents``,
``crm.CellContainer.load_from_storage``, ``crm.simulation.StorageOption``; e.g. reference
cr_mech_coli/crm_perf_plots.py:52, cr_mech_coli/crm_multilayer/runner.py:82-115) run unchanged.  Everything
is re-exported from ``cr_mech_coli_b200``; names of the reference that lie outside the mechanics path
(imaging through VTK, plotting, crm_gen) are not provided and raise AttributeError."""

import sys as _sys

import cr_mech_coli_b200 as _impl
from cr_mech_coli_b200 import crm_fit, datatypes, fitting, simulation  # noqa: F401  (sub-modules of the reference)

for _name in dir(_impl):
    if not _name.startswith("_"):
        globals()[_name] = getattr(_impl, _name)
for _sub in ("crm_fit", "datatypes", "fitting", "simulation"):
    _sys.modules[__name__ + "." + _sub] = globals()[_sub]
del _name, _sub
