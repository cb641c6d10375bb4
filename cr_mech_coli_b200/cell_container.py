"""``CellContainer`` — the result of ``run_simulation_with_agents`` (reference src/cell_container.rs).

Snapshots stay packed (numpy arrays straight from the C ABI); ``RodAgent`` objects are only
materialised when a query asks for them, so a 1e6-rod history costs arrays, not objects.
Python-visible values equal what the reference's getters clone out (src/cell_container.rs:114-305)."""

from __future__ import annotations

import pickle

import numpy as np

from . import _capi
from .datatypes import CellIdentifier, RodAgent


def counter_to_color(counter: int):
    """reference src/imaging.rs:28-40"""
    c = (int(counter) * 157 * 163 * 173) % 251**3
    q, m = divmod(c, 251**2)
    q2, m2 = divmod(m, 251)
    return (q, q2, m2)


def color_to_counter(color) -> int:
    """reference src/imaging.rs:56-66"""
    c = int(color[0]) * 251**2 + int(color[1]) * 251 + int(color[2])
    for inv in (12590168, 13775961, 12157008):
        c = (c * inv) % 251**3
    return c


class _Snapshot:
    __slots__ = ("iteration", "pos", "vel", "par", "ids", "parents", "_idents", "_parents")

    def __init__(self, iteration, pos, vel, par, ids, parents):
        self.iteration, self.pos, self.vel, self.par, self.ids, self.parents = iteration, pos, vel, par, ids, parents
        self._idents = None
        self._parents = None

    def idents(self):
        if self._idents is None:
            self._idents = [CellIdentifier._from_record(r) for r in self.ids]
            self._parents = [CellIdentifier._from_record(r) for r in self.parents]
        return self._idents, self._parents

    def agent(self, k) -> RodAgent:
        return RodAgent._from_record(self.pos[k], self.vel[k], self.par[k])

    def as_dict(self):
        """Ordered like the reference's BTreeMap<CellIdentifier, _> (Division < Initial < Inserted): callers
        that stack ``.values()`` into arrays see the reference's row order."""
        idents, parents = self.idents()
        order = sorted(range(len(idents)), key=lambda k: idents[k])
        return {idents[k]: (self.agent(k), parents[k]) for k in order}


class CellContainer:
    """reference src/cell_container.rs:10-378.  Constructor accepts the reference's
    ``{iteration: {ident: (RodAgent, parent)}}`` mapping (``CellContainer::new``, :58-101)."""

    def __init__(self, all_cells=None, path=None, _snapshots=None):
        self.path = path
        if _snapshots is None:
            _snapshots = []
            for it in sorted((all_cells or {}).keys()):
                cells = all_cells[it]
                items = sorted(cells.items(), key=lambda kv: kv[0])
                n = len(items)
                nv = items[0][1][0]._pos.shape[0] if n else 0
                pos = np.zeros((n, nv, 3), np.float32)
                vel = np.zeros((n, nv, 3), np.float32)
                par = np.zeros(n, _capi.ROD_PARAMS_DTYPE)
                ids = np.zeros(n, _capi.IDENT_DTYPE)
                parents = np.zeros(n, _capi.IDENT_DTYPE)
                for k, (ident, (agent, parent)) in enumerate(items):
                    pos[k], vel[k] = agent._pos, agent._vel
                    agent._to_record(par[k])
                    ids[k] = ident._key()
                    parents[k] = parent._key() if parent is not None else (-1, 0, 0)
                _snapshots.append(_Snapshot(int(it), pos, vel, par, ids, parents))
        self._snaps = {s.iteration: s for s in _snapshots}
        # parent map over all iterations (src/cell_container.rs:68-73)
        self.parent_map = {}
        for it in sorted(self._snaps):
            idents, parents = self._snaps[it].idents()
            self.parent_map.update(zip(idents, parents))
        identifiers = sorted(self.parent_map)
        self.cell_to_color = {ident: counter_to_color(n + 1) for n, ident in enumerate(identifiers)}
        self.color_to_cell = {c: i for i, c in self.cell_to_color.items()}
        # child_map as the reference builds it: `or_insert(vec![child]).push(child)` stores the first
        # child twice (src/cell_container.rs:86-92); get_daughter_map is the clean variant.
        self.child_map = {}
        for child in sorted(self.parent_map):
            parent = self.parent_map[child]
            if parent is not None:
                self.child_map.setdefault(parent, [child]).append(child)

    # -- the `cells` attribute of the reference (#[pyo3(get)]) -------------------------------
    @property
    def cells(self):
        return self.get_cells()

    def get_cells(self):
        return {it: self._snaps[it].as_dict() for it in sorted(self._snaps)}

    def get_cells_at_iteration(self, iteration: int):
        s = self._snaps.get(int(iteration))
        return s.as_dict() if s is not None else {}

    def get_cell_history(self, identifier):
        hist, parent = {}, None
        for it in sorted(self._snaps):
            s = self._snaps[it]
            idents, parents = s.idents()
            for k, i in enumerate(idents):
                if i == identifier:
                    hist[it] = s.agent(k)
                    parent = parents[k]
                    break
        return hist, parent

    def get_all_iterations(self):
        return sorted(self._snaps)

    def get_parent(self, identifier):
        if identifier not in self.parent_map:
            raise KeyError(f"No CellIdentifier {identifier!r} in map")
        return self.parent_map[identifier]

    def get_children(self, identifier):
        if identifier not in self.child_map:
            raise KeyError(f"No CellIdentifier {identifier!r} in map")
        return list(self.child_map[identifier])

    def get_color(self, identifier):
        return self.cell_to_color.get(identifier)

    def get_cell_from_color(self, color):
        return self.color_to_cell.get(tuple(int(c) for c in color))

    def cells_are_siblings(self, ident1, ident2) -> bool:
        if ident1 in self.parent_map and ident2 in self.parent_map:
            return self.parent_map[ident1] == self.parent_map[ident2]
        return False

    def get_parent_map(self):
        return dict(self.parent_map)

    def get_daughter_map(self):
        out = {}
        for daughter in sorted(self.parent_map):
            parent = self.parent_map[daughter]
            if parent is not None:
                out.setdefault(parent, []).append(daughter)
        return out

    def get_child_map(self):
        return {k: list(v) for k, v in self.child_map.items()}

    def get_all_identifiers(self):
        return sorted(self.parent_map)

    def get_all_identifiers_unsorted(self):
        return list(self.parent_map)

    def counter_to_cell_identifier(self, counter: int):
        idents = self.get_all_identifiers()
        if not 0 <= counter < len(idents):
            raise KeyError(f"Cannot assign CellIdentifier to counter {counter}")
        return idents[counter].copy()

    def cell_identifier_to_counter(self, identifier) -> int:
        for i, ident in enumerate(self.get_all_identifiers()):
            if ident == identifier:
                return i
        raise KeyError(f"No CellIdentifier {identifier!r} in map")

    def add_ident_divided(self, parent=None):
        """reference src/cell_container.rs:361-377"""
        new_counter = len(self.parent_map) + 1
        ident = CellIdentifier(0, new_counter)
        assert ident not in self.parent_map
        self.parent_map[ident] = parent
        if parent is not None:
            self.child_map.setdefault(parent, []).append(ident)
        color = counter_to_color(new_counter)
        assert ident not in self.cell_to_color and color not in self.color_to_cell
        self.cell_to_color[ident] = color
        self.color_to_cell[color] = ident
        return ident

    # -- packed access (extension; no reference counterpart) ---------------------------------
    def positions_at_iteration(self, iteration: int) -> np.ndarray:
        """(n_rods, n_vertices, 3) f32 array of one snapshot, rows ordered like its identifiers."""
        return self._snaps[int(iteration)].pos

    # -- (de)serialisation: reference uses serde_pickle of {cells, path} (:308-319) ------------
    def serialize(self) -> bytes:
        snaps = [(s.iteration, s.pos, s.vel, s.par, s.ids, s.parents) for s in self._snaps.values()]
        return pickle.dumps({"snapshots": snaps, "path": self.path})

    @staticmethod
    def deserialize(value: bytes) -> "CellContainer":
        try:
            d = pickle.loads(bytes(value))
            return CellContainer(path=d["path"], _snapshots=[_Snapshot(*s) for s in d["snapshots"]])
        except Exception as e:
            raise ValueError(f"serde({e})") from e

    def __reduce__(self):
        return (CellContainer.deserialize, (self.serialize(),))

    def __deepcopy__(self, memo):
        return CellContainer.deserialize(self.serialize())

    @staticmethod
    def load_from_storage(config, path):
        """reference src/cell_container.rs:322-359: the save points below ``path`` (the directory above
        ``cells``) as a container.  Reads the records ``run_simulation_with_agents`` writes under
        ``StorageOption.SerdeJson`` (this repo's JSON schema, see storage.py; cellular_raza's own files are
        rejected with a ValueError)."""
        from . import storage
        from pathlib import Path

        return CellContainer(path=Path(path), _snapshots=[_Snapshot(*s) for s in storage.read_snapshots(path)])
