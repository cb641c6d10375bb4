"""On-disk results of a run: ``StorageOption.SerdeJson`` writer and the reader behind
``CellContainer.load_from_storage`` (reference src/simulation.rs:221-230,307-316, src/cell_container.rs:322-359;
the multilayer driver persists and reloads through them, cr_mech_coli/crm_multilayer/runner.py:30-34,144,189).

Directory shape follows cellular_raza's StorageBuilder as the reference uses it: ``<storage_location>/<date>
[/<storage_suffix>]/cells/json/<iteration, 20 digits>/batch_00000000000000000000.json``; ``CellContainer.path``
is the directory above ``cells`` (reference src/simulation.rs:307-316).  **The JSON inside is this repo's own
schema, not cellular_raza's serde output** (that crate is not vendored, nothing in-tree pins its records):
files written here load here, files written by the Rust reference do not."""

from __future__ import annotations

import datetime
import json
import os
from pathlib import Path

import numpy as np

from . import _capi
from .datatypes import StorageOption

SCHEMA = "cr_mech_coli_b200/cells/1"
_FLOAT_FIELDS = [n for n in _capi.ROD_PARAMS_DTYPE.names if _capi.ROD_PARAMS_DTYPE[n].kind == "f"]
_INT_FIELDS = [n for n in _capi.ROD_PARAMS_DTYPE.names if _capi.ROD_PARAMS_DTYPE[n].kind in "iu"]


def check_options(options):
    """Memory and SerdeJson are implemented; the embedded databases (Sled, SledTemp) and Ron are not."""
    bad = [o for o in options if o not in (StorageOption.Memory, StorageOption.SerdeJson)]
    if bad:
        raise NotImplementedError(f"storage options {[o.name for o in bad]} are not implemented "
                                  "(Memory and SerdeJson are)")


def new_run_directory(config) -> Path:
    """``new_storage_builder`` (reference src/simulation.rs:221-230): location / date [/ suffix]."""
    date = datetime.datetime.now().strftime("%Y-%m-%d-T%H-%M-%S")
    base = Path(str(config.storage_location))
    path = base / date
    k = 0
    while path.exists():  # two runs within one second
        k += 1
        path = base / f"{date}-{k:03d}"
    if config.storage_suffix is not None:
        path = path / str(config.storage_suffix)
    return path


def _ident_json(rec):
    kind = int(rec["kind"])
    if kind < 0:
        return None
    if kind == _capi.ID_INITIAL:
        return {"Initial": int(rec["counter"])}
    if kind == _capi.ID_DIVISION:
        return {"Division": [int(rec["voxel"]), int(rec["counter"])]}
    return {"Inserted": [int(rec["voxel"]), int(rec["counter"])]}


def _ident_record(j):
    if j is None:
        return (-1, 0, 0)
    if "Initial" in j:
        return (_capi.ID_INITIAL, 0, int(j["Initial"]))
    if "Division" in j:
        return (_capi.ID_DIVISION, int(j["Division"][0]), int(j["Division"][1]))
    return (_capi.ID_INSERTED, int(j["Inserted"][0]), int(j["Inserted"][1]))


def write_snapshot(run_dir: Path, iteration: int, pos, vel, par, ids, parents):
    """One save point = one batch file with one record per cell: {identifier, element: {cell, parent}} (the
    logical content of cellular_raza's CellBox).  f32 values are written with repr-exact shortest decimals,
    so a load returns the same bits."""
    d = Path(run_dir) / "cells" / "json" / f"{int(iteration):020d}"
    d.mkdir(parents=True, exist_ok=True)
    cells = []
    for k in range(len(pos)):
        p = par[k]
        cell = {"pos": [[float(np.float32(x)) for x in row] for row in pos[k]],
                "vel": [[float(np.float32(x)) for x in row] for row in vel[k]]}
        for n in _FLOAT_FIELDS:
            cell[n] = float(p[n])
        for n in _INT_FIELDS:
            cell[n] = int(p[n])
        cells.append({"identifier": _ident_json(ids[k]), "element": {"cell": cell, "parent": _ident_json(parents[k])}})
    tmp = d / "batch_00000000000000000000.json.tmp"
    with open(tmp, "w") as f:
        json.dump({"schema": SCHEMA, "iteration": int(iteration), "n_vertices": int(pos.shape[1]) if len(pos) else 0,
                   "data": cells}, f, allow_nan=True)
    os.replace(tmp, d / "batch_00000000000000000000.json")


def read_snapshots(path):
    """All save points below ``path`` (the directory above ``cells``) as (iteration, pos, vel, par, ids, parents)."""
    root = Path(path) / "cells" / "json"
    if not root.is_dir():
        raise OSError(f"no stored cells below {path} (expected {root})")
    out = []
    for it_dir in sorted(p for p in root.iterdir() if p.is_dir()):
        records, nv, iteration = [], 0, int(it_dir.name)
        for batch in sorted(it_dir.glob("batch_*.json")):
            with open(batch) as f:
                j = json.load(f)
            if j.get("schema") != SCHEMA:
                raise ValueError(f"{batch}: not written by cr_mech_coli_b200 (schema {j.get('schema')!r}); cellular_raza's "
                                 "own SerdeJson records are not readable here")
            nv = max(nv, int(j.get("n_vertices", 0)))
            records.extend(j["data"])
        n = len(records)
        pos = np.zeros((n, nv, 3), np.float32)
        vel = np.zeros((n, nv, 3), np.float32)
        par = np.zeros(n, _capi.ROD_PARAMS_DTYPE)
        ids = np.zeros(n, _capi.IDENT_DTYPE)
        parents = np.zeros(n, _capi.IDENT_DTYPE)
        for k, r in enumerate(records):
            cell = r["element"]["cell"]
            pos[k] = np.asarray(cell["pos"], np.float32)
            vel[k] = np.asarray(cell["vel"], np.float32)
            for name in _capi.ROD_PARAMS_DTYPE.names:
                par[k][name] = cell[name]
            ids[k] = _ident_record(r["identifier"])
            parents[k] = _ident_record(r["element"]["parent"])
        out.append((iteration, pos, vel, par, ids, parents))
    return out
