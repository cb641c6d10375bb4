"""Data model of the drop-in boundary: the classes the reference's PyO3 module exports
(reference src/lib.rs:46-116, stub cr_mech_coli/cr_mech_coli.pyi:5-292), with the same names,
constructor signatures, defaults and kwargs fall-through.  They are attribute bags on the host;
everything that executes per step lives behind the C ABI (include/crm_b200.h)."""

from __future__ import annotations

import copy
import enum
import functools
import json

import numpy as np

from . import _capi


class StorageOption(enum.Enum):
    """reference cr_mech_coli/cr_mech_coli.pyi:5-10"""

    Sled = 0
    SledTemp = 1
    SerdeJson = 2
    Ron = 3
    Memory = 4


class VoxelPlainIndex:
    """reference cr_mech_coli/cr_mech_coli.pyi:97-103"""

    __slots__ = ("index",)

    def __init__(self, index):
        self.index = int(index.index if isinstance(index, VoxelPlainIndex) else index)

    def __eq__(self, other):
        return isinstance(other, VoxelPlainIndex) and self.index == other.index

    def __lt__(self, other):
        return self.index < other.index

    def __hash__(self):
        return hash(("VoxelPlainIndex", self.index))

    def __int__(self):
        return self.index

    def __repr__(self):
        return f"VoxelPlainIndex({self.index})"


@functools.total_ordering
class CellIdentifier:
    """cellular_raza ``CellIdentifier`` (reference cr_mech_coli/cr_mech_coli.pyi:81-95):
    ``Division(voxel, counter) < Initial(n) < Inserted(voxel, counter)`` (derive(Ord) variant order;
    `Initial` witnessed at reference src/crm_fit/predict.rs:267, `new_inserted` at src/crm_divide.rs:219)."""

    __slots__ = ("kind", "voxel", "counter")
    DIVISION, INITIAL, INSERTED = _capi.ID_DIVISION, _capi.ID_INITIAL, _capi.ID_INSERTED

    def __init__(self, voxel_index, counter):
        self.kind = self.DIVISION
        self.voxel = int(VoxelPlainIndex(voxel_index).index)
        self.counter = int(counter)

    @staticmethod
    def new_initial(counter: int) -> "CellIdentifier":
        c = CellIdentifier(0, counter)
        c.kind = CellIdentifier.INITIAL
        return c

    @staticmethod
    def new_inserted(voxel_index, counter: int) -> "CellIdentifier":
        c = CellIdentifier(voxel_index, counter)
        c.kind = CellIdentifier.INSERTED
        return c

    @staticmethod
    def _from_record(rec) -> "CellIdentifier | None":
        kind = int(rec["kind"])
        if kind < 0:
            return None
        c = CellIdentifier(int(rec["voxel"]), int(rec["counter"]))
        c.kind = kind
        return c

    def _key(self):
        return (self.kind, self.voxel, self.counter)

    def __eq__(self, other):
        return isinstance(other, CellIdentifier) and self._key() == other._key()

    def __lt__(self, other):
        return self._key() < other._key()

    def __hash__(self):
        return hash(self._key())

    def __getitem__(self, key):
        if self.kind == self.INITIAL:
            if key == 0:
                return self.counter
            raise IndexError("CellIdentifier::Initial has a single entry")
        if key == 0:
            return self.voxel
        if key == 1:
            return self.counter
        raise IndexError("CellIdentifier index out of range")

    def copy(self):
        return copy.copy(self)

    def __deepcopy__(self, memo):
        return copy.copy(self)

    def __repr__(self):
        if self.kind == self.INITIAL:
            return f"CellIdentifier::Initial({self.counter})"
        name = "Division" if self.kind == self.DIVISION else "Inserted"
        return f"CellIdentifier::{name}(VoxelPlainIndex({self.voxel}), {self.counter})"


def _f32(x):
    return float(np.float32(x))


class MorsePotentialF32:
    """reference cr_mech_coli/cr_mech_coli.pyi:35-51 (fields src/agent.rs:333-339)."""

    _fields = ("radius", "potential_stiffness", "cutoff", "strength")

    def __init__(self, radius, potential_stiffness, cutoff, strength):
        self.radius = _f32(radius)
        self.potential_stiffness = _f32(potential_stiffness)
        self.cutoff = _f32(cutoff)
        self.strength = _f32(strength)

    def __eq__(self, other):
        return isinstance(other, MorsePotentialF32) and all(getattr(self, f) == getattr(other, f) for f in self._fields)

    def __repr__(self):
        return "MorsePotentialF32 {\n" + "".join(f"    {f}: {getattr(self, f)},\n" for f in self._fields) + "}"


class MiePotentialF32:
    """reference cr_mech_coli/cr_mech_coli.pyi:14-33 (fields src/agent.rs:323-331)."""

    _fields = ("radius", "strength", "bound", "cutoff", "en", "em")

    def __init__(self, radius, strength, bound, cutoff, en, em):
        self.radius = _f32(radius)
        self.strength = _f32(strength)
        self.bound = _f32(bound)
        self.cutoff = _f32(cutoff)
        self.en = _f32(en)
        self.em = _f32(em)

    def __eq__(self, other):
        return isinstance(other, MiePotentialF32) and all(getattr(self, f) == getattr(other, f) for f in self._fields)

    def __repr__(self):
        return "MiePotentialF32 {\n" + "".join(f"    {f}: {getattr(self, f)},\n" for f in self._fields) + "}"


class PhysicalInteraction:
    """``PhysicalInteraction(PhysInt, n_neighbors)`` (reference src/agent.rs:247,279-378): wraps a Morse
    or Mie potential, proxies attribute access to it, ``ValueError`` for anything else."""

    def __init__(self, obj):
        if isinstance(obj, PhysicalInteraction):
            inner, nn = copy.copy(obj._inner), obj.n_neighbors
        elif isinstance(obj, (MorsePotentialF32, MiePotentialF32)):
            inner, nn = copy.copy(obj), 0
        else:
            raise ValueError(
                f"Could not convert {type(obj)} to any interaction potential. Use one of the provided potentials instead."
            )
        object.__setattr__(self, "_inner", inner)
        object.__setattr__(self, "n_neighbors", nn)

    def __getattr__(self, name):
        inner = object.__getattribute__(self, "_inner")
        if name in inner._fields:
            return getattr(inner, name)
        raise AttributeError(f"Cannot find attribute with name {name}")

    def __setattr__(self, name, value):
        if name in self._inner._fields:
            setattr(self._inner, name, _f32(value))
        elif name == "n_neighbors":
            object.__setattr__(self, name, int(value))
        else:
            raise AttributeError(f"Cannot find attribute with name {name}")

    @property
    def inner(self):
        return self._inner

    def __eq__(self, other):
        return isinstance(other, PhysicalInteraction) and self._inner == other._inner and self.n_neighbors == other.n_neighbors

    def __deepcopy__(self, memo):
        return PhysicalInteraction(self)

    def __repr__(self):
        return repr(self._inner)


class _Setter:
    """Common part of GrowthRateSetter / SpringLengthThresholdSetter (reference src/agent.rs:11-205):
    either ``NormalDistr{mean, std}`` or ``Explicit{a, b}``; built from / converted to a dict."""

    _explicit_keys = ("a", "b")

    def __init__(self, d=None, **kw):
        d = dict(d or {}, **kw)
        k1, k2 = self._explicit_keys
        has_n = "mean" in d and "std" in d
        has_e = k1 in d and k2 in d
        if has_n and not (k1 in d or k2 in d) and len(d) == 2:
            self.kind, self.a, self.b = _capi.SETTER_NORMAL, _f32(d["mean"]), _f32(d["std"])
        elif has_e and not ("mean" in d or "std" in d) and len(d) == 2:
            self.kind, self.a, self.b = _capi.SETTER_EXPLICIT, _f32(d[k1]), _f32(d[k2])
        else:
            raise KeyError(
                f"could not find suitable combination of either ('mean', 'std') or ('{k1}', '{k2}') keys in dict."
            )

    @classmethod
    def from_pydict(cls, d):
        return d if isinstance(d, cls) else cls(d)

    def to_pydict(self):
        if self.kind == _capi.SETTER_NORMAL:
            return {"mean": self.a, "std": self.b}
        k1, k2 = self._explicit_keys
        return {k1: self.a, k2: self.b}

    def __getattr__(self, name):
        d = object.__getattribute__(self, "to_pydict")()
        if name in d:
            return d[name]
        raise AttributeError(name)

    def __eq__(self, other):
        return type(self) is type(other) and (self.kind, self.a, self.b) == (other.kind, other.a, other.b)

    def __repr__(self):
        inner = ", ".join(f"{k}: {v}" for k, v in self.to_pydict().items())
        variant = "NormalDistr" if self.kind == _capi.SETTER_NORMAL else "Explicit"
        return f"{variant} {{ {inner} }}"


class GrowthRateSetter(_Setter):
    """reference src/agent.rs:94-205, stub cr_mech_coli.pyi:59-63; default NormalDistr(0.01, 0)."""

    _explicit_keys = ("g1", "g2")

    @staticmethod
    def default_dict():
        return {"mean": 0.01, "std": 0.0}


class SpringLengthThresholdSetter(_Setter):
    """reference src/agent.rs:11-92, stub cr_mech_coli.pyi:53-57; default NormalDistr(0.01, 0)
    (sic, reference src/agent.rs:28-35)."""

    _explicit_keys = ("l1", "l2")

    @staticmethod
    def default_dict():
        return {"mean": 0.01, "std": 0.0}


def _as_matrix(a, name):
    a = np.array(a, dtype=np.float32)
    if a.ndim != 2 or a.shape[1] != 3:
        raise ValueError(f"{name} must have shape (n_vertices, 3)")
    return a


class RodMechanicsSettings:
    """reference src/config.rs:32-140 (defaults :128-140)."""

    _scalars = ("diffusion_constant", "spring_tension", "rigidity", "spring_length", "damping")

    def __init__(self):
        self._pos = np.zeros((8, 3), dtype=np.float32)
        self._vel = np.zeros((8, 3), dtype=np.float32)
        self.diffusion_constant = 0.0
        self.spring_tension = 1.0
        self.rigidity = 2.0
        self.spring_length = 3.0
        self.damping = 1.0

    pos = property(lambda s: s._pos.copy(), lambda s, v: setattr(s, "_pos", _as_matrix(v, "pos")))
    vel = property(lambda s: s._vel.copy(), lambda s, v: setattr(s, "_vel", _as_matrix(v, "vel")))

    def __setattr__(self, name, value):
        if name in self._scalars:
            value = _f32(value)
        elif name not in ("_pos", "_vel", "pos", "vel"):
            raise AttributeError(f"RodMechanicsSettings has no attribute {name}")
        object.__setattr__(self, name, value)

    def __eq__(self, other):
        return (isinstance(other, RodMechanicsSettings) and all(getattr(self, k) == getattr(other, k) for k in self._scalars)
                and np.array_equal(self._pos, other._pos) and np.array_equal(self._vel, other._vel))

    def __repr__(self):
        return "RodMechanicsSettings {\n" + "".join(f"    {k}: {getattr(self, k)},\n" for k in self._scalars) + "}"


class AgentSettings:
    """reference src/config.rs:143-389.  ``AgentSettings(**kwargs)``: attributes of the class itself are
    set directly, everything else falls through to ``interaction`` and then ``mechanics`` (:275-301)."""

    _own = ("mechanics", "interaction", "growth_rate", "growth_rate_setter", "spring_length_threshold",
            "spring_length_threshold_setter", "neighbor_reduction")

    def __init__(self, **kwargs):
        self.mechanics = RodMechanicsSettings()
        self.interaction = PhysicalInteraction(MorsePotentialF32(3.0, 0.5, 10.0, 0.1))
        self.growth_rate = 0.01
        self.growth_rate_setter = GrowthRateSetter({"mean": 0.01, "std": 0.0})
        self.spring_length_threshold = 6.0
        self.spring_length_threshold_setter = SpringLengthThresholdSetter({"l1": 6.0, "l2": 6.0})
        self.neighbor_reduction = None
        for key, value in kwargs.items():
            if key in self._own:
                setattr(self, key, value)
            else:
                try:
                    getattr(self.interaction, key)
                    setattr(self.interaction, key, value)
                except AttributeError:
                    try:
                        getattr(self.mechanics, key)
                    except AttributeError:
                        raise AttributeError(f"'AgentSettings' object has no attribute '{key}'") from None
                    setattr(self.mechanics, key, value)

    def __setattr__(self, name, value):
        if name == "interaction":
            value = PhysicalInteraction(value)
        elif name == "growth_rate_setter":
            value = GrowthRateSetter.from_pydict(value)
        elif name == "spring_length_threshold_setter":
            value = SpringLengthThresholdSetter.from_pydict(value)
        elif name in ("growth_rate", "spring_length_threshold"):
            value = _f32(value)
        elif name == "neighbor_reduction" and value is not None:
            value = (int(value[0]), _f32(value[1]))
        elif name not in self._own:
            raise AttributeError(f"'AgentSettings' object has no attribute '{name}'")
        object.__setattr__(self, name, value)

    def to_rod_agent_dict(self) -> dict:
        """reference src/config.rs:322-388"""
        m = self.mechanics
        return {
            "diffusion_constant": m.diffusion_constant,
            "spring_tension": m.spring_tension,
            "rigidity": m.rigidity,
            "spring_length": m.spring_length,
            "spring_length_threshold_setter": self.spring_length_threshold_setter.to_pydict(),
            "damping": m.damping,
            "interaction": PhysicalInteraction(self.interaction),
            "growth_rate_setter": self.growth_rate_setter.to_pydict(),
            "growth_rate": self.growth_rate,
            "spring_length_threshold": self.spring_length_threshold,
            "neighbor_reduction": self.neighbor_reduction,
        }

    def __eq__(self, other):
        return isinstance(other, AgentSettings) and all(getattr(self, k) == getattr(other, k) for k in self._own)

    def __deepcopy__(self, memo):
        new = AgentSettings()
        new.mechanics = copy.deepcopy(self.mechanics)
        for k in self._own[1:]:
            setattr(new, k, copy.deepcopy(getattr(self, k)))
        return new

    def __repr__(self):
        return (f"AgentSettings {{\n{self.mechanics!r}\n{self.interaction!r}\ngrowth_rate: {self.growth_rate}\n"
                f"spring_length_threshold: {self.spring_length_threshold}\n}}\n")


class RodAgent:
    """reference src/agent.rs:212-239 (ctor :449-510, accessors :512-570)."""

    def __init__(self, pos, vel, interaction, diffusion_constant=0.0, spring_tension=1.0, rigidity=2.0,
                 spring_length=3.0, damping=1.0, growth_rate=0.01, growth_rate_setter=None,
                 spring_length_threshold=6.0, spring_length_threshold_setter=None, neighbor_reduction=None):
        self._pos = _as_matrix(pos, "pos")
        self._vel = _as_matrix(vel, "vel")
        if self._vel.shape != self._pos.shape:
            raise ValueError("pos and vel must have the same shape")
        self.interaction = PhysicalInteraction(interaction)
        self.diffusion_constant = _f32(diffusion_constant)
        self.spring_tension = _f32(spring_tension)
        self.rigidity = _f32(rigidity)
        self.spring_length = _f32(spring_length)
        self.damping = _f32(damping)
        self.growth_rate = _f32(growth_rate)
        self.growth_rate_setter = GrowthRateSetter.from_pydict(
            GrowthRateSetter.default_dict() if growth_rate_setter is None else growth_rate_setter)
        self.spring_length_threshold = _f32(spring_length_threshold)
        self.spring_length_threshold_setter = SpringLengthThresholdSetter.from_pydict(
            SpringLengthThresholdSetter.default_dict() if spring_length_threshold_setter is None
            else spring_length_threshold_setter)
        self.neighbor_reduction = None if neighbor_reduction is None else (int(neighbor_reduction[0]), _f32(neighbor_reduction[1]))

    pos = property(lambda s: s._pos.copy(), lambda s, v: setattr(s, "_pos", _as_matrix(v, "pos")))
    vel = property(lambda s: s._vel.copy(), lambda s, v: setattr(s, "_vel", _as_matrix(v, "vel")))

    @property
    def radius(self):
        return self.interaction.radius

    @radius.setter
    def radius(self, value):
        self.interaction.radius = value

    @property
    def mechanics(self):
        """Read-only view of the ``RodMechanics`` part (pos, vel, five scalars)."""
        m = RodMechanicsSettings()
        m.pos, m.vel = self._pos, self._vel
        for k in RodMechanicsSettings._scalars:
            setattr(m, k, getattr(self, k))
        return m

    def __deepcopy__(self, memo):
        return self._clone()

    def _clone(self):
        new = object.__new__(RodAgent)
        new.__dict__.update(self.__dict__)
        new._pos, new._vel = self._pos.copy(), self._vel.copy()
        new.interaction = PhysicalInteraction(self.interaction)
        return new

    def __eq__(self, other):
        if not isinstance(other, RodAgent):
            return NotImplemented
        keys = [k for k in self.__dict__ if k not in ("_pos", "_vel")]
        return (np.array_equal(self._pos, other._pos) and np.array_equal(self._vel, other._vel)
                and all(getattr(self, k) == getattr(other, k) for k in keys))

    def __repr__(self):
        return (f"RodAgent {{ pos: {self._pos.tolist()}, spring_length: {self.spring_length}, growth_rate: "
                f"{self.growth_rate}, interaction: {self.interaction!r} }}")

    # -- packing to / from the C ABI record ------------------------------------------------
    def _to_record(self, rec):
        for k in ("diffusion_constant", "spring_tension", "rigidity", "spring_length", "damping", "growth_rate",
                  "spring_length_threshold"):
            rec[k] = getattr(self, k)
        p = self.interaction.inner
        rec["n_neighbors"] = self.interaction.n_neighbors
        rec["radius"], rec["strength"], rec["cutoff"] = p.radius, p.strength, p.cutoff
        if isinstance(p, MorsePotentialF32):
            rec["potential"] = _capi.POT_MORSE
            rec["potential_stiffness"] = p.potential_stiffness
        else:
            rec["potential"] = _capi.POT_MIE
            rec["bound"], rec["en"], rec["em"] = p.bound, p.en, p.em
        g, t = self.growth_rate_setter, self.spring_length_threshold_setter
        rec["growth_setter_kind"], rec["growth_setter_a"], rec["growth_setter_b"] = g.kind, g.a, g.b
        rec["threshold_setter_kind"], rec["threshold_setter_a"], rec["threshold_setter_b"] = t.kind, t.a, t.b
        if self.neighbor_reduction is not None:
            rec["has_neighbor_reduction"] = 1
            rec["neighbor_reduction_max"], rec["neighbor_reduction_exp"] = self.neighbor_reduction

    @staticmethod
    def _from_record(pos, vel, rec) -> "RodAgent":
        if int(rec["potential"]) == _capi.POT_MORSE:
            pot = MorsePotentialF32(rec["radius"], rec["potential_stiffness"], rec["cutoff"], rec["strength"])
        else:
            pot = MiePotentialF32(rec["radius"], rec["strength"], rec["bound"], rec["cutoff"], rec["en"], rec["em"])
        inter = PhysicalInteraction(pot)
        inter.n_neighbors = int(rec["n_neighbors"])
        gk = ("mean", "std") if int(rec["growth_setter_kind"]) == _capi.SETTER_NORMAL else ("g1", "g2")
        tk = ("mean", "std") if int(rec["threshold_setter_kind"]) == _capi.SETTER_NORMAL else ("l1", "l2")
        nr = (int(rec["neighbor_reduction_max"]), float(rec["neighbor_reduction_exp"])) if int(rec["has_neighbor_reduction"]) else None
        return RodAgent(
            pos, vel, inter, rec["diffusion_constant"], rec["spring_tension"], rec["rigidity"], rec["spring_length"],
            rec["damping"], rec["growth_rate"], {gk[0]: rec["growth_setter_a"], gk[1]: rec["growth_setter_b"]},
            rec["spring_length_threshold"], {tk[0]: rec["threshold_setter_a"], tk[1]: rec["threshold_setter_b"]}, nr)


class Configuration:
    """reference src/config.rs:391-563 (defaults :444-465).  ``Configuration(**kwargs)`` sets any attribute
    (:475-486); unknown names raise AttributeError like PyO3's setattr."""

    _defaults = dict(
        n_threads=1, t0=0.0, dt=0.1, t_max=100.0, n_saves=10, progressbar=None, domain_size=(100.0, 100.0),
        domain_height=2.5, n_voxels=(1, 1), rng_seed=0, gel_pressure=0.0, surface_friction=0.0,
        surface_friction_distance=1.0, storage_options=None, storage_location="out", storage_suffix=None)
    _floats = ("t0", "dt", "t_max", "domain_height", "gel_pressure", "surface_friction", "surface_friction_distance")

    def __init__(self, **kwargs):
        for k, v in self._defaults.items():
            setattr(self, k, [StorageOption.Memory] if k == "storage_options" else v)
        for k, v in kwargs.items():
            setattr(self, k, v)

    def __setattr__(self, name, value):
        if name not in self._defaults:
            raise AttributeError(f"'Configuration' object has no attribute '{name}'")
        if name in self._floats:
            value = _f32(value)
        elif name == "domain_size":
            value = [_f32(value[0]), _f32(value[1])]
        elif name == "n_voxels":
            value = [int(value[0]), int(value[1])]
        elif name in ("n_threads", "n_saves", "rng_seed"):
            value = int(value)
            if value < (1 if name == "n_threads" else 0):
                raise ValueError(f"{name} out of range")
        elif name == "storage_options":
            value = [v if isinstance(v, StorageOption) else StorageOption[v] for v in value]
        object.__setattr__(self, name, value)

    def _to_c(self) -> _capi.CrmConfig:
        return _capi.make_config(
            t0=self.t0, dt=self.dt, t_max=self.t_max, n_saves=self.n_saves, domain_size=self.domain_size,
            domain_height=self.domain_height, n_voxels=self.n_voxels, rng_seed=self.rng_seed,
            gel_pressure=self.gel_pressure, surface_friction=self.surface_friction,
            surface_friction_distance=self.surface_friction_distance, n_threads=self.n_threads)

    def _as_dict(self):
        d = {k: getattr(self, k) for k in self._defaults}
        d["storage_options"] = [o.name for o in self.storage_options]
        d["storage_location"] = str(self.storage_location)
        d["storage_suffix"] = None if self.storage_suffix is None else str(self.storage_suffix)
        return d

    def to_json(self) -> str:
        return json.dumps(self._as_dict(), indent=2)

    @staticmethod
    def from_json(json_string: str) -> "Configuration":
        try:
            return Configuration(**json.loads(json_string))
        except Exception as e:  # reference maps serde errors to IOError (src/config.rs:501)
            raise IOError(str(e)) from e

    def to_hash(self) -> int:
        import hashlib

        return int.from_bytes(hashlib.blake2b(self.to_json().encode(), digest_size=8).digest(), "little")

    @staticmethod
    def from_toml_string(toml_string: str) -> "Configuration":
        import tomllib

        try:
            return Configuration(**tomllib.loads(toml_string))
        except Exception as e:  # reference: ValueError (src/config.rs:535)
            raise ValueError(str(e)) from e

    @staticmethod
    def from_toml(filename) -> "Configuration":
        with open(filename) as f:
            return Configuration.from_toml_string(f.read())

    def serialize(self) -> bytes:
        import pickle

        return pickle.dumps(self._as_dict())

    @staticmethod
    def deserialize(data: bytes) -> "Configuration":
        import pickle

        try:
            return Configuration(**pickle.loads(bytes(data)))
        except Exception as e:
            raise ValueError(str(e)) from e

    def __reduce__(self):
        return (Configuration.deserialize, (self.serialize(),))

    def __deepcopy__(self, memo):
        return Configuration(**{k: copy.deepcopy(getattr(self, k)) for k in self._defaults})

    def __eq__(self, other):
        return isinstance(other, Configuration) and self._as_dict() == other._as_dict()

    def __repr__(self):
        return "Configuration {\n" + "".join(f"    {k}: {v!r},\n" for k, v in self._as_dict().items()) + "}"
