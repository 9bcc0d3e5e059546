"""
Minimal unit registry with the slice of the `pint` API that the reference engine uses.

The reference passes `pint.UnitRegistry` / `pint.Quantity` objects through `MDParams`
(/root/reference/swarmrl/engine/espresso.py:51-88) and converts with `.m_as("sim_...")` after
defining its simulation units (espresso.py:211-234).  `pint` is an optional dependency of this
engine: a real `pint.UnitRegistry` works unchanged (only `.define`, `.Quantity`,
`.boltzmann_constant`, attribute access to units and `Quantity.m_as` are used); when `pint` is not
installed this module provides the same calls.

Dimensions tracked: (length, mass, time, temperature).  Magnitudes are stored in SI.
"""
from __future__ import annotations

import ast
import re

import numpy as np

_DIM0 = (0, 0, 0, 0)


def _dim_add(a, b):
    return tuple(x + y for x, y in zip(a, b))


def _dim_scale(a, k):
    return tuple(x * k for x in a)


class Quantity:
    """value * unit; `si` is the magnitude in SI base units, `dim` the dimension exponents."""

    __array_priority__ = 1000  # numpy scalars/arrays defer to our __rmul__ etc.

    def __init__(self, si, dim, ureg, unit_name=None):
        self._si = si if np.isscalar(si) else np.asarray(si, dtype=float)
        self._dim = tuple(dim)
        self._ureg = ureg
        self._unit_name = unit_name

    # -- pint-like accessors -------------------------------------------------
    def m_as(self, unit):
        u = self._ureg._parse(unit) if isinstance(unit, str) else unit
        if u._dim != self._dim:
            raise ValueError(f"cannot convert dimension {self._dim} to '{unit}' ({u._dim})")
        out = self._si / u._si
        return np.array(out) if isinstance(out, np.ndarray) else float(out)

    def to(self, unit):
        u = self._ureg._parse(unit) if isinstance(unit, str) else unit
        return Quantity(self._si, self._dim, self._ureg, unit_name=unit if isinstance(unit, str) else None)

    @property
    def magnitude(self):
        if self._unit_name is not None:
            return self.m_as(self._unit_name)
        return self._si

    m = magnitude

    @property
    def units(self):
        return self._unit_name or f"SI{self._dim}"

    # -- arithmetic ----------------------------------------------------------
    def _coerce(self, other):
        if isinstance(other, Quantity):
            return other
        if hasattr(other, "m_as"):  # a real pint quantity mixed in: not supported
            raise TypeError("cannot mix pint quantities with swarmrl_b200.units quantities")
        return Quantity(other, _DIM0, self._ureg)

    def __mul__(self, other):
        o = self._coerce(other)
        return Quantity(self._si * o._si, _dim_add(self._dim, o._dim), self._ureg)

    __rmul__ = __mul__

    def __truediv__(self, other):
        o = self._coerce(other)
        return Quantity(self._si / o._si, _dim_add(self._dim, _dim_scale(o._dim, -1)), self._ureg)

    def __rtruediv__(self, other):
        o = self._coerce(other)
        return o.__truediv__(self)

    def __pow__(self, k):
        return Quantity(self._si**k, _dim_scale(self._dim, k), self._ureg)

    def _same(self, other):
        o = self._coerce(other)
        if o._dim != self._dim:
            raise ValueError("dimension mismatch")
        return o

    def __add__(self, other):
        return Quantity(self._si + self._same(other)._si, self._dim, self._ureg, self._unit_name)

    __radd__ = __add__

    def __sub__(self, other):
        return Quantity(self._si - self._same(other)._si, self._dim, self._ureg, self._unit_name)

    def __neg__(self):
        return Quantity(-self._si, self._dim, self._ureg, self._unit_name)

    def __lt__(self, other):
        return bool(np.all(self._si < self._same(other)._si))

    def __gt__(self, other):
        return bool(np.all(self._si > self._same(other)._si))

    def __eq__(self, other):
        if not isinstance(other, Quantity):
            return NotImplemented
        return self._dim == other._dim and bool(np.all(self._si == other._si))

    def __hash__(self):
        return hash((self._dim, np.asarray(self._si).tobytes()))

    def __len__(self):
        return len(self._si)

    def __iter__(self):
        for v in np.asarray(self._si):
            yield Quantity(float(v), self._dim, self._ureg, self._unit_name)

    def __getitem__(self, k):
        return Quantity(np.asarray(self._si)[k], self._dim, self._ureg, self._unit_name)

    def __repr__(self):
        return f"<Quantity({self.magnitude}, '{self.units}')>"


class UnitRegistry:
    """`ureg.Quantity(v, "unit")`, `ureg.define("name = expr")`, `ureg.<unit>`."""

    def __init__(self):
        L, M, T, K = (1, 0, 0, 0), (0, 1, 0, 0), (0, 0, 1, 0), (0, 0, 0, 1)
        self._units = {}
        base = {
            "meter": (1.0, L), "m": (1.0, L), "metre": (1.0, L),
            "kilogram": (1.0, M), "kg": (1.0, M), "gram": (1e-3, M),
            "second": (1.0, T), "s": (1.0, T),
            "kelvin": (1.0, K), "K": (1.0, K),
            "dimensionless": (1.0, _DIM0),
        }
        for n, (f, d) in base.items():
            self._units[n] = Quantity(f, d, self)
        prefixes = {"micro": 1e-6, "milli": 1e-3, "nano": 1e-9, "centi": 1e-2, "kilo": 1e3}
        for p, f in prefixes.items():
            for b in ("meter", "second", "gram"):
                q = self._units[b]
                self._units[p + b] = Quantity(q._si * f, q._dim, self)
        self._units["um"] = self._units["micrometer"]
        newton = Quantity(1.0, (1, 1, -2, 0), self)
        self._units.update(
            newton=newton, N=newton,
            pascal=Quantity(1.0, (-1, 1, -2, 0), self), Pa=Quantity(1.0, (-1, 1, -2, 0), self),
            joule=Quantity(1.0, (2, 1, -2, 0), self), J=Quantity(1.0, (2, 1, -2, 0), self),
            boltzmann_constant=Quantity(1.380649e-23, (2, 1, -2, -1), self),
            radian=Quantity(1.0, _DIM0, self), degree=Quantity(np.pi / 180.0, _DIM0, self),
        )

    # -- parsing ---------------------------------------------------------------
    def _parse(self, expr: str) -> Quantity:
        s = expr.strip()
        if s in self._units:
            return self._units[s]
        # pint allows implicit multiplication: "1e-6 meter"
        s = re.sub(r"(?<=[0-9.)])\s+(?=[A-Za-z_(])", "*", s)
        s = s.replace("^", "**")
        tree = ast.parse(s, mode="eval")
        return self._eval(tree.body)

    def _eval(self, node) -> Quantity:
        if isinstance(node, ast.BinOp):
            a, b = self._eval(node.left), self._eval(node.right)
            if isinstance(node.op, ast.Mult):
                return a * b
            if isinstance(node.op, ast.Div):
                return a / b
            if isinstance(node.op, ast.Pow):
                if b._dim != _DIM0:
                    raise ValueError("exponent must be dimensionless")
                return a ** float(b._si)
            raise ValueError("unsupported operator in unit expression")
        if isinstance(node, ast.UnaryOp) and isinstance(node.op, ast.USub):
            return Quantity(-1.0, _DIM0, self) * self._eval(node.operand)
        if isinstance(node, ast.Constant) and isinstance(node.value, (int, float)):
            return Quantity(float(node.value), _DIM0, self)
        if isinstance(node, ast.Name):
            if node.id not in self._units:
                raise ValueError(f"unknown unit '{node.id}'")
            return self._units[node.id]
        raise ValueError("unsupported unit expression")

    # -- public API ------------------------------------------------------------
    def define(self, definition: str):
        name, expr = definition.split("=", 1)
        self._units[name.strip()] = self._parse(expr)

    def Quantity(self, value, unit="dimensionless"):  # noqa: N802  (pint's name)
        u = self._parse(unit) if isinstance(unit, str) else unit
        v = value if np.isscalar(value) else np.asarray(value, dtype=float)
        return Quantity(v * u._si, u._dim, self, unit_name=unit if isinstance(unit, str) else None)

    def __call__(self, unit):
        return self._parse(unit)

    def __getattr__(self, name):
        units = self.__dict__.get("_units", {})
        if name in units:
            q = units[name]
            return Quantity(q._si, q._dim, self, unit_name=name)
        raise AttributeError(name)


def default_registry():
    """A pint registry when pint is installed, else this module's stand-in."""
    try:
        import pint  # noqa: F401

        return pint.UnitRegistry()
    except ModuleNotFoundError:
        return UnitRegistry()
