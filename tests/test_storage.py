"""Storage: JSON save points written and read back bit for bit, container pickle round trip,
the `cr_mech_coli` import shim.  CPU only (the device part is in tests/test_gpu_api.py)."""
import copy
import pickle

import numpy as np
import pytest

from cr_mech_coli_b200 import CellContainer, Configuration, StorageOption, _capi, storage
from cr_mech_coli_b200.cell_container import _Snapshot
from cr_mech_coli_b200.testing import lattice_colony, mie_params, morse_params


def _snapshots():
    rng = np.random.default_rng(0)
    out = []
    for it, n in ((0, 3), (50, 3), (100, 5)):
        pos = rng.normal(50, 20, (n, 8, 3)).astype(np.float32)
        vel = rng.normal(0, 1e-3, (n, 8, 3)).astype(np.float32)
        par = morse_params(n) if it < 100 else mie_params(n, en=0.94, em=0.22)
        par["spring_length"] = rng.uniform(2, 6, n)
        par["n_neighbors"] = rng.integers(0, 5, n)
        ids = np.zeros(n, _capi.IDENT_DTYPE)
        parents = np.zeros(n, _capi.IDENT_DTYPE)
        for k in range(n):
            ids[k] = (_capi.ID_INITIAL, 0, k) if it < 100 or k < 1 else (_capi.ID_DIVISION, 2, k)
            parents[k] = (-1, 0, 0) if it < 100 or k < 1 else (_capi.ID_INITIAL, 0, 1 + (k - 1) // 2)
        out.append((it, pos, vel, par, ids, parents))
    return out


def test_json_round_trip_is_bit_exact(tmp_path):
    snaps = _snapshots()
    for s in snaps:
        storage.write_snapshot(tmp_path, *s)
    assert (tmp_path / "cells" / "json" / f"{100:020d}" / "batch_00000000000000000000.json").exists()
    back = storage.read_snapshots(tmp_path)
    assert [b[0] for b in back] == [0, 50, 100]
    for a, b in zip(snaps, back):
        for x, y in zip(a[1:], b[1:]):
            assert x.dtype == y.dtype and x.tobytes() == y.tobytes()


def test_load_from_storage_container(tmp_path):
    snaps = _snapshots()
    for s in snaps:
        storage.write_snapshot(tmp_path, *s)
    ref = CellContainer(path=tmp_path, _snapshots=[_Snapshot(*s) for s in snaps])
    c = CellContainer.load_from_storage(Configuration(storage_options=[StorageOption.Memory, StorageOption.SerdeJson]),
                                        tmp_path)
    assert c.get_all_iterations() == [0, 50, 100]
    assert c.get_all_identifiers() == ref.get_all_identifiers()
    assert c.get_parent_map() == ref.get_parent_map()
    assert c.cell_to_color == ref.cell_to_color
    cells = c.get_cells_at_iteration(100)
    assert list(cells) == sorted(cells)  # BTreeMap order: Division < Initial
    for ident, (agent, parent) in cells.items():
        ra, rp = ref.get_cells_at_iteration(100)[ident]
        assert np.array_equal(agent.pos, ra.pos) and parent == rp and agent.spring_length == ra.spring_length
    with pytest.raises(OSError):
        CellContainer.load_from_storage(Configuration(), tmp_path / "nothing")


def test_foreign_json_is_rejected(tmp_path):
    d = tmp_path / "cells" / "json" / f"{0:020d}"
    d.mkdir(parents=True)
    (d / "batch_00000000000000000000.json").write_text('{"data": [], "storage_info": "cellular_raza"}')
    with pytest.raises(ValueError, match="not written by cr_mech_coli_b200"):
        storage.read_snapshots(tmp_path)


def test_container_pickle_round_trip():
    c = CellContainer(path=None, _snapshots=[_Snapshot(*s) for s in _snapshots()])
    for d in (CellContainer.deserialize(c.serialize()), pickle.loads(pickle.dumps(c)), copy.deepcopy(c)):
        assert d.get_all_iterations() == c.get_all_iterations()
        assert d.get_parent_map() == c.get_parent_map() and d.get_child_map() == c.get_child_map()
        assert np.array_equal(d.positions_at_iteration(100), c.positions_at_iteration(100))
    with pytest.raises(ValueError, match="serde"):
        CellContainer.deserialize(b"garbage")


def test_unimplemented_backends_say_so():
    storage.check_options([StorageOption.Memory, StorageOption.SerdeJson])
    for o in (StorageOption.Sled, StorageOption.SledTemp, StorageOption.Ron):
        with pytest.raises(NotImplementedError):
            storage.check_options([o])


def test_cr_mech_coli_import_shim():
    import cr_mech_coli as crm
    import cr_mech_coli_b200 as impl

    for name in ("Configuration", "AgentSettings", "RodAgent", "generate_agents", "run_simulation_with_agents",
                 "CellContainer", "CellIdentifier", "MorsePotentialF32", "MiePotentialF32", "PhysicalInteraction",
                 "StorageOption", "sort_cellular_identifiers", "counter_to_color", "color_to_counter", "MINUTE", "HOUR",
                 "MICRO_METRE"):
        assert getattr(crm, name) is getattr(impl, name), name
    assert crm.simulation.StorageOption is impl.StorageOption
    from cr_mech_coli.crm_fit import Settings  # noqa: F401
    from cr_mech_coli.fitting import overlap, parents_diff_mask, render_mask_2d  # noqa: F401
    # the construction pattern of the reference's callers (cr_mech_coli/crm_perf_plots.py:52,
    # cr_mech_coli/crm_multilayer/runner.py:97)
    settings = crm.AgentSettings(growth_rate=0.02)
    p = np.zeros((8, 3), np.float32)
    agent = crm.RodAgent(pos=p, vel=0 * p, **settings.to_rod_agent_dict())
    assert agent.growth_rate == pytest.approx(0.02)
