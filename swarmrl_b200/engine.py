"""
SwarmEngine -- the `swarmrl.engine.Engine` implementation backed by the sm_100a kernels.

Mirrors the public surface of the reference's `EspressoMD`
(/root/reference/swarmrl/engine/espresso.py:132-1347) for the rollout hot path:
`MDParams`, `add_colloids`, `add_colloid_on_point`, `add_confining_walls`,
`add_const_force_to_colloids`, `get_friction_coefficients`, `manage_forces`, `integrate`,
`finalize`, `get_particle_data`, `get_unit_system` and the attributes `colloids`, `params`,
`step_idx`, `slice_idx`, `write_idx`, `traj_holder`, `colloid_radius_register`, `n_dims`, `seed`.
The control flow of `integrate` follows espresso.py:1251-1308 statement by statement (including
the per-chunk `calc_reward`, the kill-switch ordering and the snapshot-before-chunk rule); the
physics that the reference delegates to ESPResSo runs in libswarm_b200.so.

`thermostat_type="langevin"` selects the Langevin thermostat with velocity Verlet (espresso.py:1187-1201,
without a fluid); mass and rinertia are then mandatory as in the reference (:391-400).

Out of scope: rods / virtual sites, Gay-Berne, lattice Boltzmann, flow and
potential fields, user-defined walls.
"""
from __future__ import annotations

import dataclasses
import logging
import pathlib

import numpy as np

from .components import Colloid

logger = logging.getLogger("swarmrl.engine.b200")


class Engine:
    """Abstract engine API (/root/reference/swarmrl/engine/engine.py:8-45)."""

    def integrate(self, n_slices: int, force_model) -> None:
        raise NotImplementedError

    def get_particle_data(self) -> dict:
        raise NotImplementedError

    def finalize(self):
        pass


@dataclasses.dataclass()
class MDParams:
    """Same fields and defaults as the reference's MDParams (espresso.py:30-88)."""

    def __init__(self, ureg, box_length=None, fluid_dyn_viscosity=None, WCA_epsilon=None, temperature=None,
                 time_step=None, time_slice=None, write_interval=None, periodic: bool = True,
                 thermostat_type: str = "brownian"):
        if box_length is None:
            box_length = ureg.Quantity(3 * [1000], "micrometer")
        if fluid_dyn_viscosity is None:
            fluid_dyn_viscosity = ureg.Quantity(1e-3, "pascal*second")
        if WCA_epsilon is None:
            WCA_epsilon = ureg.Quantity(300, "kelvin") * ureg.boltzmann_constant
        if temperature is None:
            temperature = ureg.Quantity(300, "kelvin")
        if time_step is None:
            time_step = ureg.Quantity(1e-3, "second")
        if time_slice is None:
            time_slice = ureg.Quantity(1e-1, "second")
        if write_interval is None:
            write_interval = ureg.Quantity(1, "second")
        self.ureg = ureg
        self.box_length = box_length
        self.fluid_dyn_viscosity = fluid_dyn_viscosity
        self.WCA_epsilon = WCA_epsilon
        self.temperature = temperature
        self.time_step = time_step
        self.time_slice = time_slice
        self.write_interval = write_interval
        self.periodic = periodic
        self.thermostat_type = thermostat_type


def _get_random_start_pos(init_radius, init_center, dim, rng):
    """Draw order of espresso.py:91-105 (r then angle(s))."""
    if dim == 2:
        r = init_radius * np.sqrt(rng.random())
        theta = 2 * np.pi * rng.random()
        pos = r * np.array([np.cos(theta), np.sin(theta), 0])
        assert init_center[2] == 0.0
    elif dim == 3:
        r = init_radius * np.cbrt(rng.random())
        th, ph = np.arccos(2.0 * rng.random() - 1), 2.0 * np.pi * rng.random()
        pos = r * np.array([np.sin(th) * np.cos(ph), np.sin(th) * np.sin(ph), np.cos(th)])
    else:
        raise ValueError("Random position finder only implemented for 2d and 3d")
    return pos + init_center


def _calc_friction_coefficients(dyn_visc, radius):
    """Stokes: gamma_t = 6 pi eta r, gamma_r = 8 pi eta r^3 (espresso.py:108-113)."""
    return 6 * np.pi * dyn_visc * radius, 8 * np.pi * dyn_visc * radius**3


class ColloidHandle:
    """What `engine.colloids[i]` yields: the attributes the reference reads off an ESPResSo
    particle handle (`id, type, pos, v, director`) plus `velocity` so it also passes for a
    `Colloid`.  Values are host copies taken at the engine's current state."""

    __slots__ = ("_view", "_i")

    def __init__(self, view, i):
        self._view, self._i = view, i

    @property
    def id(self):
        return int(self._i)

    @property
    def type(self):
        return int(self._view._engine._types[self._i])

    @property
    def pos(self):
        return self._view._snapshot()[0][self._i].copy()

    @property
    def v(self):
        return self._view._snapshot()[1][self._i].copy()

    velocity = v

    @property
    def director(self):
        return self._view._snapshot()[2][self._i].copy()

    def __eq__(self, other):
        return self.id == other.id

    def __hash__(self):
        return hash(self.id)

    def __repr__(self):
        return f"ColloidHandle(id={self.id}, type={self.type}, pos={self.pos}, director={self.director})"


class ColloidsView:
    """`engine.colloids`: a lazy sequence over the device state.  One device->host transfer per
    engine state version serves every item access; device-aware observables / tasks
    (swarmrl_b200.observables, swarmrl_b200.tasks) recognise the view and never leave HBM."""

    def __init__(self, engine):
        self._engine = engine
        self._cache = None
        self._cache_version = -1

    @property
    def engine(self):
        return self._engine

    def _snapshot(self):
        e = self._engine
        if self._cache is None or self._cache_version != e._state_version:
            self._cache = e._fetch_host_state()
            self._cache_version = e._state_version
        return self._cache

    def __len__(self):
        return self._engine.n_colloids

    def __getitem__(self, i):
        if isinstance(i, slice):
            return [ColloidHandle(self, k) for k in range(*i.indices(len(self)))]
        n = len(self)
        if i < 0:
            i += n
        if not 0 <= i < n:
            raise IndexError(i)
        return ColloidHandle(self, i)

    def __iter__(self):
        for i in range(len(self)):
            yield ColloidHandle(self, i)

    def as_colloids(self):
        """List of frozen `Colloid`s, as manage_forces builds it (espresso.py:1216-1225)."""
        pos, vel, dr = self._snapshot()
        t = self._engine._types
        return [Colloid(pos=pos[i].copy(), velocity=vel[i].copy(), director=dr[i].copy(), id=i, type=int(t[i]))
                for i in range(len(self))]


class _SystemView:
    """The few `espressomd.System` attributes that callers of the reference engine read."""

    class _Cells:
        skin = 0.4

    def __init__(self, engine):
        self._e = engine
        self.cell_system = _SystemView._Cells()

    @property
    def time(self):
        return self._e._time

    @property
    def box_l(self):
        return self._e._box_l.copy()

    @property
    def time_step(self):
        return self._e._time_step

    @property
    def periodicity(self):
        return 3 * [self._e.params.periodic]


def _default_backend(**kw):
    from .device import DeviceSwarm

    return DeviceSwarm(**kw)


class SwarmEngine(Engine):
    """Drop-in for the reference's `EspressoMD` on the colloids + WCA + Brownian-dynamics path."""

    backend_factory = staticmethod(_default_backend)

    def __init__(self, md_params, n_dims=3, seed=42, out_folder=".", write_chunk_size=100, system=None,
                 h5_group_tag=None, skin=None, persistent=True, async_writer=True, domain=None):
        """domain: None (one GPU), or a dict for ONE swarm slab-decomposed over the ranks of a process group -- keys
        world, rank, group (torch.distributed; None = default group), exchange_every, margin, halo_extra (the range of
        the widest per-agent observable, e.g. `vision_range`), device, max_neighbors, capacity_factor: see
        swarmrl_b200/domain_engine.py.  Every rank creates the same engine with the same particles and calls the same
        methods; `get_particle_data` / `colloids` are then collectives, and rank 0 alone writes the trajectory."""
        self._domain = None if domain is None else dict(domain)
        self.params = md_params
        self.out_folder = pathlib.Path(out_folder).resolve()
        self.seed = seed
        self.rng = np.random.default_rng(self.seed)
        if n_dims not in [2, 3]:
            raise ValueError("Only 2d and 3d are allowed")
        self.n_dims = n_dims
        self._init_unit_system()
        self.write_chunk_size = write_chunk_size
        self.h5_group_tag = "colloids" if h5_group_tag is None else h5_group_tag
        self._skin = skin
        self._persistent = persistent
        self._async_writer = async_writer   # False: the synchronous snapshot / chunk writer (what the oracle backend uses)
        self._writer = None
        self._time = 0.0
        self._init_system()
        self.system = _SystemView(self)
        self.colloids = ColloidsView(self)
        self.colloid_radius_register = {}
        self.integration_initialised = False
        self.traj_holder = None
        self.write_idx = 0
        # host-side staging of particles until the first integrate()
        self._pos, self._dir, self._types_l, self._props = [], [], [], []
        self._ext_force = {}
        self._walls = None
        self._types_cache = np.zeros(0, dtype=np.int64)
        self._backend = None
        self._state_version = 0
        self._capacity_proven = False   # first device-path chunk has run with a check (see _run_chunk)
        self._unchecked_steps = 0

    # ------------------------------------------------------------------ setup
    def _init_unit_system(self):
        """espresso.py:211-234."""
        self.ureg = self.params.ureg
        for d in (
            "sim_length = 1e-6 meter",
            "sim_time = 1 second",
            "sim_energy = 293 kelvin * boltzmann_constant",
            "sim_velocity = sim_length / sim_time",
            "sim_angular_velocity = 1 / sim_time",
            "sim_mass = sim_energy / sim_velocity**2",
            "sim_rinertia = sim_length**2 * sim_mass",
            "sim_dyn_viscosity = sim_mass / (sim_length * sim_time)",
            "sim_kin_viscosity = sim_length**2 / sim_time",
            "sim_force = sim_mass * sim_length / sim_time**2",
            "sim_torque = sim_length * sim_force",
        ):
            self.ureg.define(d)

    def _init_system(self):
        """espresso.py:236-288."""
        time_step = self.params.time_step.m_as("sim_time")
        time_slice = self.params.time_slice.m_as("sim_time")
        write_interval = self.params.write_interval.m_as("sim_time")
        box_l = np.array(self.params.box_length.m_as("sim_length"))
        if box_l.ndim == 0:
            raise ValueError("box_length must be a 3d vector (or 2d if you have a 2d system)")
        if self.n_dims == 2 and len(box_l) == 2:
            box_l = np.array([box_l[0], box_l[1], box_l[0]])
        if len(box_l) != 3:
            raise ValueError(f"box_length must be a 3d vector. You gave {self.params.box_length}")
        self._box_l = box_l.astype(float)
        self._time_step = float(time_step)
        steps_per_write_interval = int(round(write_interval / time_step))
        self.params.steps_per_write_interval = steps_per_write_interval
        if abs(steps_per_write_interval - write_interval / time_step) > 1e-10:
            raise ValueError("inconsistent parameters: write_interval must be integer multiple of time_step")
        steps_per_slice = int(round(time_slice / time_step))
        self.params.steps_per_slice = steps_per_slice
        if abs(steps_per_slice - time_slice / time_step) > 1e-10:
            raise ValueError("inconsistent parameters: time_slice must be integer multiple of time_step")

    def _check_already_initialised(self):
        if self.integration_initialised:
            raise RuntimeError("You cannot change the system configuration after the first call to integrate()")

    @property
    def _types(self):
        """Types of the staged particles as an array (rebuilt only when particles were added since the last look)."""
        if len(self._types_cache) != len(self._types_l):
            self._types_cache = np.array(self._types_l, dtype=np.int64)
        return self._types_cache

    @property
    def n_colloids(self):
        return len(self._pos) if self._backend is None else self._backend.n

    def add_colloid_on_point(self, radius_colloid=None, init_position=None, init_direction=np.array([1, 0, 0]),
                             type_colloid=0, gamma_translation=None, gamma_rotation=None, aspect_ratio: float = 1.0,
                             mass=None, rinertia=None):
        """espresso.py:307-457."""
        self._check_already_initialised()
        if aspect_ratio != 1.0:
            raise NotImplementedError("Gay-Berne (aspect_ratio != 1) is outside the B200 hot path")
        if self.params.thermostat_type == "langevin":  # espresso.py:391-400
            if mass is None:
                raise ValueError("If you use the Langevin thermostat, you must set a particle mass")
            if rinertia is None:
                raise ValueError("If you use the Langevin thermostat, you must set a particle rotational inertia")
        if radius_colloid is None:
            radius_colloid = self.ureg.Quantity(1, "micrometer")
        if init_position is None:
            init_position = 0.5 * self.params.box_length
        radius_simunits = float(radius_colloid.m_as("sim_length"))
        if type_colloid in self.colloid_radius_register:
            if self.colloid_radius_register[type_colloid]["radius"] != radius_simunits:
                raise ValueError(
                    f"The chosen type {type_colloid} is already taken and used with a different radius"
                    f" {self.colloid_radius_register[type_colloid]['radius']}. Choose a new combination")
        init_pos = np.array(init_position.m_as("sim_length"), dtype=float)
        init_direction = np.array(init_direction, dtype=float)
        init_direction = init_direction / np.linalg.norm(init_direction)
        gt_s, gr_s = _calc_friction_coefficients(
            self.params.fluid_dyn_viscosity.m_as("sim_dyn_viscosity"), radius_simunits)
        gamma_t = gt_s if gamma_translation is None else gamma_translation.m_as("sim_force/sim_velocity")
        gamma_r = gr_s if gamma_rotation is None else gamma_rotation.m_as("sim_torque/sim_angular_velocity")
        water_dens = self.params.ureg.Quantity(1000, "kg/meter**3")
        if mass is None:
            mass = water_dens * 4.0 / 3.0 * np.pi * radius_colloid**3
        m_sim = float(mass.m_as("sim_mass"))
        if rinertia is None:
            rin_sim = np.full(3, 2.0 / 5.0 * m_sim * radius_simunits**2)
        else:
            rin_sim = np.broadcast_to(np.asarray(rinertia.m_as("sim_rinertia"), dtype=float), (3,)).copy()
        if self.n_dims == 2:
            init_pos[2] = 0
            theta = np.arccos(np.clip(init_direction[2], -1, 1))
            if abs(theta - np.pi / 2) > 10e-6:
                raise ValueError(
                    "It seems like you want to have a 2D simulation with colloids that point some amount in"
                    " Z-direction. Change something in your colloid setup.")
            # the 2-D state is an angle: the director lies exactly in the plane from here on
            init_direction = np.array([init_direction[0], init_direction[1], 0.0])
            init_direction = init_direction / np.linalg.norm(init_direction)
        prop = (
            int(type_colloid),
            tuple(np.broadcast_to(np.asarray(gamma_t, dtype=float), (3,)).tolist()),
            tuple(np.broadcast_to(np.asarray(gamma_r, dtype=float), (3,)).tolist()),
            m_sim,
            tuple(rin_sim.tolist()),
        )
        self._pos.append(init_pos)
        self._dir.append(init_direction)
        self._types_l.append(int(type_colloid))
        self._props.append(prop)
        self.colloid_radius_register.update({type_colloid: {"radius": radius_simunits, "aspect_ratio": aspect_ratio}})
        self._state_version += 1
        return self.colloids[len(self._pos) - 1]

    def add_colloids(self, n_colloids: int, radius_colloid=None, random_placement_center=None,
                     random_placement_radius=None, type_colloid: int = 0, gamma_translation=None,
                     gamma_rotation=None, aspect_ratio: float = 1.0, mass=None, rinertia=None):
        """espresso.py:459-544; RNG draw order per colloid: position, then orientation."""
        self._check_already_initialised()
        if random_placement_center is None:
            init_center = 0.5 * np.array(self.params.box_length.m_as("sim_length"), dtype=float)
            if len(init_center) == 2:
                init_center = np.array([init_center[0], init_center[1], 0.0])
        else:
            init_center = np.array(random_placement_center.m_as("sim_length"), dtype=float)
        if random_placement_radius is None:
            init_rad = 0.5 * float(np.min(self.params.box_length.m_as("sim_length")))
        else:
            init_rad = float(random_placement_radius.m_as("sim_length"))
        for _ in range(n_colloids):
            start_pos = _get_random_start_pos(init_rad, init_center, self.n_dims, self.rng)
            if self.n_dims == 3:
                th, ph = np.arccos(2.0 * self.rng.random() - 1), 2.0 * np.pi * self.rng.random()
            else:
                th, ph = np.pi / 2, 2 * np.pi * self.rng.random()
            init_direction = np.array([np.sin(th) * np.cos(ph), np.sin(th) * np.sin(ph), np.cos(th)])
            self.add_colloid_on_point(
                radius_colloid=radius_colloid, init_position=self.ureg.Quantity(start_pos, "sim_length"),
                init_direction=init_direction, type_colloid=type_colloid, gamma_translation=gamma_translation,
                gamma_rotation=gamma_rotation, aspect_ratio=aspect_ratio, mass=mass, rinertia=rinertia)

    def add_colloids_at(self, positions, directions, radius_colloid=None, type_colloid: int = 0, gamma_translation=None,
                        gamma_rotation=None, mass=None, rinertia=None):
        """Bulk form of `add_colloid_on_point` (espresso.py:307-457) for large swarms -- not in the reference, whose
        per-particle `system.part.add` is fine for its tens of colloids: `positions` is a pint quantity of shape
        (n, 3), `directions` an (n, 3) array; every particle gets the same radius / type / friction, computed and
        checked exactly as for a single particle (the first one goes through `add_colloid_on_point` itself)."""
        positions_sim = np.array(positions.m_as("sim_length"), dtype=float).reshape(-1, 3)
        directions = np.array(directions, dtype=float).reshape(-1, 3)
        if len(positions_sim) != len(directions):
            raise ValueError("positions and directions must have the same length")
        if len(positions_sim) == 0:
            return
        self.add_colloid_on_point(radius_colloid=radius_colloid, init_position=self.ureg.Quantity(positions_sim[0], "sim_length"),
                                  init_direction=directions[0], type_colloid=type_colloid, gamma_translation=gamma_translation,
                                  gamma_rotation=gamma_rotation, mass=mass, rinertia=rinertia)
        prop = self._props[-1]
        rest_p, rest_d = positions_sim[1:].copy(), directions[1:].copy()
        rest_d /= np.linalg.norm(rest_d, axis=1, keepdims=True)
        if self.n_dims == 2:
            if np.any(np.abs(np.arccos(np.clip(rest_d[:, 2], -1, 1)) - np.pi / 2) > 10e-6):
                raise ValueError(
                    "It seems like you want to have a 2D simulation with colloids that point some amount in"
                    " Z-direction. Change something in your colloid setup.")
            rest_p[:, 2] = 0.0
            rest_d[:, 2] = 0.0
            rest_d /= np.linalg.norm(rest_d, axis=1, keepdims=True)
        self._pos.extend(rest_p)
        self._dir.extend(rest_d)
        self._types_l.extend([int(type_colloid)] * len(rest_p))
        self._props.extend([prop] * len(rest_p))
        self._state_version += 1

    def add_confining_walls(self, wall_type: int):
        """WCA walls on the box faces (espresso.py:667-711)."""
        self._check_already_initialised()
        if wall_type in self.colloid_radius_register:
            raise ValueError(f"wall type {wall_type} is already taken by other system component. Choose a new one")
        self._walls = int(wall_type)
        self.colloid_radius_register.update({wall_type: {"radius": 0.0, "aspect_ratio": 1.0}})

    def add_const_force_to_colloids(self, force, type: int):  # noqa: A002 (reference's argument name)
        """espresso.py:834-851."""
        f = np.asarray(force.m_as("sim_force"), dtype=float)
        if not np.any(self._types == type):
            raise ValueError(f"Particles of type {type} not added to engine. "
                             f"You currently have {self.colloid_radius_register.keys()}")
        self._ext_force[int(type)] = f
        if self._backend is not None:
            self._upload_ext_force()

    def get_friction_coefficients(self, type: int):  # noqa: A002
        prop = self.colloid_radius_register.get(type, None)
        if prop is None:
            raise ValueError(f"cannot get friction coefficient for type {type}. Did you actually add that particle type?")
        return _calc_friction_coefficients(self.params.fluid_dyn_viscosity.m_as("sim_dyn_viscosity"), prop["radius"])

    def get_unit_system(self):
        return self.ureg

    # ---------------------------------------------------------- first integrate
    def _setup_interactions(self):
        """WCA table per type pair (espresso.py:802-832): sigma = (r_a + r_b) 2^(-1/6), cut = r_a + r_b."""
        ars = [d["aspect_ratio"] for d in self.colloid_radius_register.values()]
        if len(np.unique(ars)) > 1:
            raise ValueError("All particles in the system must have the same aspect ratio.")
        tmax = max(self.colloid_radius_register.keys(), default=0)
        if tmax >= 8 or min(self.colloid_radius_register.keys(), default=0) < 0:
            raise ValueError("particle types must be in 0..7")
        T = tmax + 1
        eps = float(self.params.WCA_epsilon.m_as("sim_energy"))
        sig, ep, cut = np.zeros((T, T)), np.zeros((T, T)), np.zeros((T, T))
        for a, pa in self.colloid_radius_register.items():
            for b, pb in self.colloid_radius_register.items():
                s = (pa["radius"] + pb["radius"]) * 2 ** (-1 / 6)
                sig[a, b], ep[a, b], cut[a, b] = s, eps, s * 2 ** (1 / 6)
        walls = None
        if self._walls is not None:
            w = self._walls
            walls = dict(sigma=sig[:, w].copy(), eps=ep[:, w].copy(), cut=cut[:, w].copy())
            # the wall type never appears as a particle: no pair interaction rows needed for it
        return dict(sigma=sig, eps=ep, cut=cut), walls

    def _create_backend(self):
        wca, walls = self._setup_interactions()
        uniq, species = {}, []
        for p in self._props:
            species.append(uniq.setdefault(p, len(uniq)))
        if len(uniq) > 16:
            raise ValueError("at most 16 distinct (type, friction, mass, inertia) property sets are supported")
        table = [dict(type=p[0], gamma_t=p[1], gamma_r=p[2], mass=p[3], rinertia=p[4]) for p in uniq]
        if not table:
            table = [dict(type=0, gamma_t=(1.0,) * 3, gamma_r=(1.0,) * 3, mass=1.0, rinertia=(1.0,) * 3)]
        kT = float((self.params.temperature * self.ureg.boltzmann_constant).m_as("sim_energy"))
        self._kT = kT
        allowed = ["brownian", "langevin"]
        if self.params.thermostat_type not in allowed:
            raise ValueError(f"integrator_type must be one of {allowed}")
        kw = dict(n_dims=self.n_dims, box=self._box_l, periodic=self.params.periodic, dt=self._time_step, kT=kT,
                  seed=int(self.seed), species=table, wca=wca, walls=walls,
                  skin=0.4 if self._skin is None else float(self._skin), persistent=self._persistent)
        if self._domain is not None:
            from .domain_engine import DomainBackend

            if self._ext_force:
                raise NotImplementedError("constant external forces are not supported on the domain-decomposed swarm")
            self._backend = DomainBackend(**kw, velocities=(self.params.thermostat_type == "langevin"), **self._domain)
        else:
            self._backend = type(self).backend_factory(**kw)
        pos = np.array(self._pos, dtype=float).reshape(-1, 3)
        dirs = np.array(self._dir, dtype=float).reshape(-1, 3)
        if self.n_dims == 2:
            theta = np.arctan2(dirs[:, 1], dirs[:, 0])
            self._backend.set_state(pos, np.array(species, dtype=np.int32), theta=theta)
        else:
            from .device import director_to_quat

            self._backend.set_state(pos, np.array(species, dtype=np.int32), quat=director_to_quat(dirs))
        if self._ext_force:
            self._upload_ext_force()

    def _upload_ext_force(self):
        F = np.zeros((self.n_colloids, 3))
        for t, f in self._ext_force.items():
            F[self._types == t] = f
        self._backend.set_ext_force_host(F)

    def _remove_overlap(self):
        """espresso.py:1161-1201: 1000 steepest-descent iterations, time restored, then the thermostat and the
        integrator of `thermostat_type` (Brownian dynamics, or Langevin + velocity Verlet without a fluid)."""
        self._backend.relax_checked(1000, 0.1, 0.1)
        self._backend.set_thermostat(self._kT, int(self.seed))
        self._backend.set_integrator(self.params.thermostat_type)
        self._state_version += 1

    # -------------------------------------------------------------- trajectory
    def _init_h5_output(self):
        """espresso.py:1054-1108 (dataset layout); h5py is optional, see `_write_traj_chunk_to_file`."""
        self.h5_filename = self.out_folder / "trajectory.hdf5"
        self.out_folder.mkdir(parents=True, exist_ok=True)
        self.traj_holder = {k: list() for k in
                            ("Times", "Ids", "Types", "Unwrapped_Positions", "Velocities", "Directors")}
        self.write_idx = 0
        self.h5_time_steps_written = 0
        self._writer = None
        be = self._backend
        self._writes_files = getattr(be, "rank", 0) == 0   # decomposed swarm: every rank takes the (collective) snapshot, rank 0 writes
        if self._async_writer and be is not None and hasattr(be, "lib") and hasattr(be, "h") and self.n_colloids > 0 \
                and not getattr(be, "is_domain", False):
            # CUDA backend: device snapshot ring -> pinned host -> writer thread (swarmrl_b200/trajectory.py); the hot
            # loop only enqueues.  `traj_holder` is the writer's pending chunk (filled by its thread).
            from .trajectory import AsyncTrajectoryWriter

            self._writer = AsyncTrajectoryWriter(be, self._types, self.out_folder, self.h5_group_tag, self.write_chunk_size,
                                                 h5_filename=self.h5_filename)
            self.traj_holder = self._writer.pending
            self._h5py = self._writer._h5py
            return
        try:
            import h5py
        except ModuleNotFoundError:
            self._h5py = None
            self._npz_chunks = 0
            return
        self._h5py = h5py
        if not self._writes_files:
            return
        n = self.n_colloids
        with h5py.File(self.h5_filename.as_posix(), "a") as f:
            g = f.require_group(self.h5_group_tag)
            kw = dict(compression="gzip")
            L = self.write_chunk_size
            g.require_dataset("Times", shape=(L, 1, 1), maxshape=(None, 1, 1), dtype=float, **kw)
            for name in ["Ids", "Types"]:
                g.require_dataset(name, shape=(L, n, 1), maxshape=(None, n, 1), dtype=int, **kw)
            for name in ["Unwrapped_Positions", "Velocities", "Directors"]:
                g.require_dataset(name, shape=(L, n, 3), maxshape=(None, n, 3), dtype=float, **kw)

    def _update_traj_holder(self):
        """Snapshot taken BEFORE the chunk is integrated (espresso.py:1110-1130, 1278-1280)."""
        if self.n_colloids == 0:
            logger.warning("No colloids in the system. Not writing to hdf5")
            return
        if self._writer is not None:
            self._writer.snapshot(self._time)
            return
        pos, vel, dr = self.colloids._snapshot()
        self.traj_holder["Times"].append(np.array([self._time])[:, np.newaxis])
        self.traj_holder["Ids"].append(np.arange(self.n_colloids)[:, np.newaxis])
        self.traj_holder["Types"].append(self._types.copy()[:, np.newaxis])
        self.traj_holder["Unwrapped_Positions"].append(pos.copy())
        self.traj_holder["Velocities"].append(vel.copy())
        self.traj_holder["Directors"].append(dr.copy())

    def _write_traj_chunk_to_file(self):
        """espresso.py:1132-1159.  Without h5py the same arrays go to numbered .npz chunks
        `trajectory.<group>.<k>.npz` next to where the HDF5 file would be."""
        if self.traj_holder is None:
            return
        n_new = len(self.traj_holder["Times"])
        if n_new == 0:
            return
        if not getattr(self, "_writes_files", True):
            self.h5_time_steps_written += n_new
            return
        if self._h5py is not None:
            with self._h5py.File(self.h5_filename, "a") as f:
                g = f[self.h5_group_tag]
                for key, vals in self.traj_holder.items():
                    ds = g[key]
                    ds.resize(self.h5_time_steps_written + n_new, axis=0)
                    ds[self.h5_time_steps_written: self.h5_time_steps_written + n_new, ...] = np.stack(vals, axis=0)
        else:
            fn = self.out_folder / f"trajectory.{self.h5_group_tag}.{self._npz_chunks:05d}.npz"
            np.savez_compressed(fn, **{k: np.stack(v, axis=0) for k, v in self.traj_holder.items()})
            self._npz_chunks += 1
        self.h5_time_steps_written += n_new

    # ------------------------------------------------------------------ forces
    def manage_forces(self, force_model=None):
        """espresso.py:1203-1249."""
        if force_model is None:
            return
        if self._backend is None:
            self._create_backend()
        if getattr(force_model, "has_device_path", False):
            from .device_view import SwarmDeviceView

            force_model.calc_action_device(SwarmDeviceView(self))
            self._state_version += 1
            return
        actions = force_model.calc_action(self.colloids.as_colloids())
        n = self.n_colloids
        f = np.zeros(n)
        tq = np.zeros((n, 3))
        nd = np.zeros((n, 3))
        has = np.zeros(n, dtype=np.int32)
        for i, a in enumerate(actions[:n]):
            f[i] = a.force
            if a.torque is not None:
                tq[i] = a.torque
            if a.new_direction is not None:
                nd[i] = a.new_direction
                has[i] = 1
        self._backend.set_actions_host(f, tq)
        if has.any():
            if self.n_dims == 2:
                # espresso.py:1240-1248: an exactly antiparallel new_direction has old x new = 0, the rotation axis is
                # 0/0 and `round(nan)` raises ValueError there; same here instead of an arbitrary turn
                old = self.colloids._snapshot()[2]
                sel = has.astype(bool)
                ang = np.arccos(np.einsum("ij,ij->i", nd[sel], old[sel]))
                cr = np.cross(old[sel], nd[sel])
                if np.any((ang > 1e-6) & (np.linalg.norm(cr, axis=1) == 0.0)):
                    raise ValueError("cannot convert float NaN to integer")
            self._backend.set_new_directions_host(nd, has)
            self._backend.apply_new_direction()
            self._state_version += 1

    # --------------------------------------------------------------- integrate
    def integrate(self, n_slices, force_model=None):
        """The rollout hot loop, statement for statement espresso.py:1251-1308."""
        if not self.integration_initialised:
            self.slice_idx = 0
            self.step_idx = 0
            if self._backend is None:
                self._create_backend()
            self._remove_overlap()
            self._init_h5_output()
            self.integration_initialised = True

        old_slice_idx = self.slice_idx
        device_path = getattr(force_model, "has_device_path", False)

        while self.step_idx < self.params.steps_per_slice * (old_slice_idx + n_slices):
            if self.step_idx == self.params.steps_per_write_interval * self.write_idx:
                self._update_traj_holder()
                self.write_idx += 1
                if self._writer is None and len(self.traj_holder["Times"]) >= self.write_chunk_size:
                    self._write_traj_chunk_to_file()
                    for val in self.traj_holder.values():
                        val.clear()

            if force_model is not None:
                if force_model.kill_switch:
                    break

            if self.step_idx == self.params.steps_per_slice * self.slice_idx:
                self.slice_idx += 1
                self.manage_forces(force_model)

            steps_to_next_write = self.params.steps_per_write_interval * self.write_idx - self.step_idx
            steps_to_next_slice = self.params.steps_per_slice * self.slice_idx - self.step_idx
            steps_to_next = min(steps_to_next_write, steps_to_next_slice)

            self._run_chunk(steps_to_next, deferred_check=device_path)
            # the reference calls this unguarded (espresso.py:1307): force_model=None raises there too
            if device_path:
                from .device_view import SwarmDeviceView

                force_model.calc_reward_device(SwarmDeviceView(self))
            else:
                force_model.calc_reward(self.colloids)
            self.step_idx += steps_to_next
        self._check_deferred_runs()

    def _run_chunk(self, n_steps, deferred_check=False):
        """system.integrator.run(n_steps) (espresso.py:1304-1306).

        On the device path (nothing of a slice visits the host) only the FIRST chunk of this engine is run with a
        synchronising check -- it grows the Verlet capacity to twice what the swarm needs -- and the later chunks are
        only enqueued; `integrate` reads the counters once, after its loop.  A list overflow in an unchecked chunk
        halts the integrator on the device (later launches are no-ops) and is raised there as OverflowError."""
        be = self._backend
        if deferred_check and self._capacity_proven and hasattr(be, "run") and hasattr(be, "stats"):
            be.run(int(n_steps), int(self.step_idx))
            self._unchecked_steps += int(n_steps)
        else:
            st = be.run_checked(int(n_steps), int(self.step_idx))
            if deferred_check and hasattr(be, "set_max_neighbors") and st is not None:
                if 2 * int(st.max_neighbors_seen) > int(st.capacity):
                    be.set_max_neighbors(min(128, max(2 * int(st.max_neighbors_seen), int(st.capacity))))
                self._capacity_proven = True
        self._time += n_steps * self._time_step
        self._state_version += 1

    def _check_deferred_runs(self):
        if not self._unchecked_steps:
            return
        self._unchecked_steps = 0
        st = self._backend.stats()
        if st.overflow:
            self._capacity_proven = False
            raise OverflowError(
                f"a Verlet list overflowed in a chunk that was enqueued without a check (capacity {st.capacity}, "
                f"needed >= {st.max_neighbors_seen}); the integrator halted on the device. Re-create the engine with a "
                "larger max_neighbors, or drive it through the object path, which grows the lists chunk by chunk")

    def finalize(self):
        """espresso.py:1310-1318: flush, no snapshot."""
        if self.traj_holder is None:
            return
        if self._writer is not None:
            self._writer.flush()
            self.h5_time_steps_written = self._writer.frames_written
            return
        self._write_traj_chunk_to_file()
        for val in self.traj_holder.values():
            val.clear()

    def reward_means(self, agent):
        """Device tensor (n_evaluations, 2) of the [sum, mean] of every reward evaluation a `DeviceAgent(reward_mean=True)`
        collected (trainers/trainer.py:92-95 takes the mean on the host).  On a decomposed swarm the per-rank sums are
        reduced over the ranks here, in ONE collective for the whole rollout -- every rank must call it."""
        import torch

        means = list(agent.reward_means)
        be = self._backend
        if getattr(be, "is_domain", False):
            return be.all_reduce_means(means)
        if not means:
            return torch.zeros((0, 2), dtype=torch.float64)
        return torch.stack(means)

    # ---------------------------------------------------------------- read-out
    def _fetch_host_state(self):
        if self._backend is None:
            pos = np.array(self._pos, dtype=float).reshape(-1, 3)
            dr = np.array(self._dir, dtype=float).reshape(-1, 3)
            return pos, np.zeros_like(pos), dr
        return self._backend.particle_data_host()

    def get_particle_data(self):
        """espresso.py:1320-1336."""
        pos, vel, dr = self.colloids._snapshot()
        return {
            "Id": np.arange(self.n_colloids),
            "Type": self._types.copy(),
            "Unwrapped_Positions": pos.copy(),
            "Velocities": vel.copy(),
            "Directors": dr.copy(),
        }


# the name the reference's scripts import
EspressoMD = SwarmEngine
