"""Shared builders for the parity tests: the same seeded inputs for the oracle and the CUDA path."""
import numpy as np

ETA_SIM = 0.2200  # water, 8.9e-4 Pa s in simulation units
KT_300 = 300.0 / 293.0


def stokes(radius, eta=ETA_SIM):
    """gamma_t = 6 pi eta r, gamma_r = 8 pi eta r^3 (/root/reference/swarmrl/engine/espresso.py:108-113)."""
    return 6 * np.pi * eta * radius, 8 * np.pi * eta * radius**3


def species_for(radii_by_type, eta=ETA_SIM, aniso=None):
    """One species per particle type with Stokes friction and the reference's default mass / inertia
    (espresso.py:406-413: water density, solid sphere)."""
    table = []
    for t, r in enumerate(radii_by_type):
        gt, gr = stokes(r, eta)
        rho = 1000.0 / 4.0453e-9 * 1e-18  # kg/m^3 -> sim_mass / um^3
        m = rho * 4.0 / 3.0 * np.pi * r**3
        gt3 = (gt, gt, gt) if aniso is None else tuple(aniso[0])
        gr3 = (gr, gr, gr) if aniso is None else tuple(aniso[1])
        table.append(dict(type=t, gamma_t=gt3, gamma_r=gr3, mass=m, rinertia=(0.4 * m * r * r,) * 3))
    return table


def wca_for(radii_by_type, eps=297.0 / 293.0):
    """espresso.py:810-819: sigma = (r_a + r_b) 2^(-1/6), cut = sigma 2^(1/6)."""
    T = len(radii_by_type)
    sig = np.zeros((T, T)); ep = np.zeros((T, T)); cut = np.zeros((T, T))
    for a in range(T):
        for b in range(T):
            s = (radii_by_type[a] + radii_by_type[b]) * 2 ** (-1 / 6)
            sig[a, b], ep[a, b], cut[a, b] = s, eps, s * 2 ** (1 / 6)
    return dict(sigma=sig, eps=ep, cut=cut)


def jittered_lattice(n, L, n_dims, rng, jitter=0.25, types=1):
    """n particles on a jittered square / cubic lattice in a periodic box of edge L; no overlaps for
    spacing - 2*jitter*spacing > 2 r (caller's business).  Positions may be unfolded by +-L."""
    m = int(np.ceil(n ** (1.0 / n_dims)))
    a = L / m
    idx = np.arange(m**n_dims)[:n]
    pos = np.zeros((n, 3))
    for d in range(n_dims):
        pos[:, d] = ((idx // m**d) % m + 0.5) * a
    pos[:, :n_dims] += rng.uniform(-jitter * a, jitter * a, (n, n_dims))
    return pos


def random_quats(n, rng):
    q = rng.normal(size=(n, 4))
    return q / np.linalg.norm(q, axis=1, keepdims=True)


def make_pair(n_dims, box, periodic, dt, kT, seed, species, wca, pos, sp_idx, theta=None, quat=None, walls=None,
              device_kw=None):
    """(OracleSwarm, DeviceSwarm) on identical inputs."""
    from oracle.oracle import OracleSwarm
    from swarmrl_b200.device import DeviceSwarm

    o = OracleSwarm(n_dims, box, periodic, dt, kT, seed, species, wca, walls=walls)
    o.set_state(pos, sp_idx, theta=theta, quat=quat)
    d = DeviceSwarm(n_dims, box, periodic, dt, kT, seed, species, wca, walls=walls, **(device_kw or {}))
    d.set_state(pos, sp_idx, theta=theta, quat=quat)
    return o, d


def set_actions(o, d, f_swim, torque):
    """Write the same swim forces / torques into both."""
    f_swim = np.asarray(f_swim, dtype=float)
    torque = np.asarray(torque, dtype=float).reshape(-1, 3)
    if o is not None:
        o.f_swim[:] = f_swim
        for a in range(3):
            o.torque[a][:] = torque[:, a]
    if d is not None:
        d.set_actions_host(f_swim, torque)


def rel_err(a, b, floor=1e-300):
    a, b = np.asarray(a, dtype=float), np.asarray(b, dtype=float)
    scale = max(np.max(np.abs(b)), floor)
    return float(np.max(np.abs(a - b)) / scale) if a.size else 0.0
