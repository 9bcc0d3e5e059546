"""
TEST INFRASTRUCTURE: an oracle-backed stand-in for `swarmrl_b200.device.DeviceSwarm`, so the host
logic of `SwarmEngine` (the integrate loop of /root/reference/swarmrl/engine/espresso.py:1251-1308,
manage_forces, trajectory holder, setup errors) can be tested on a box without a GPU.  It is
plugged in through `SwarmEngine.backend_factory` by the tests only; the product never imports it.
"""
import numpy as np

from oracle.oracle import OracleSwarm


class OracleBackend:
    def __init__(self, n_dims, box, periodic, dt, kT, seed, species, wca, walls=None, skin=0.4, persistent=True,
                 **_):
        self._o = OracleSwarm(n_dims, box, periodic, dt, 0.0, seed, species, wca, walls=walls)
        self.n_dims = n_dims
        self.n = 0
        self.kT, self.seed = kT, seed
        self.runs = []
        self.thermostat_type = "brownian"

    def set_state(self, pos, species, theta=None, quat=None):
        self._o.set_state(pos, species, theta=theta, quat=quat)
        self.n = self._o.n
        self.ptype_host = self._o.ptype
        self._nd = np.zeros((self.n, 3))
        self._has = np.zeros(self.n, dtype=np.int32)

    def set_thermostat(self, kT, seed):
        self._o.params.kT, self._o.params.seed = float(kT), int(seed)

    def set_integrator(self, thermostat_type):
        if thermostat_type not in ("brownian", "langevin"):
            raise ValueError("integrator_type must be one of ['brownian', 'langevin']")
        self.thermostat_type = thermostat_type

    def relax_checked(self, n_iter=1000, gamma=0.1, max_disp=0.1):
        self._o.sd_steps(n_iter, gamma, max_disp, mode=0)

    def run_checked(self, n_steps, step0, grow=True):
        self.runs.append((int(n_steps), int(step0)))
        if self.thermostat_type == "langevin":
            self._o.lv_steps(step0, n_steps, mode=0)
        else:
            self._o.bd_steps(step0, n_steps, mode=0)

    def particle_data_host(self):
        return self._o.pos.copy(), self._o.velocities.copy(), self._o.directors.copy()

    def set_actions_host(self, f_swim, torque):
        self._o.f_swim[:] = f_swim
        t = np.asarray(torque, dtype=float).reshape(-1, 3)
        for a in range(3):
            self._o.torque[a][:] = t[:, a]

    def set_ext_force_host(self, F):
        F = np.asarray(F, dtype=float).reshape(-1, 3)
        for a in range(3):
            self._o.ext_force[a][:] = F[:, a]

    def set_new_directions_host(self, nd, has):
        self._nd, self._has = np.asarray(nd, dtype=float).reshape(-1, 3), np.asarray(has)

    def apply_new_direction(self):
        """espresso.py:1236-1249 (2-D branch: rotate in plane by acos(new.old) when > 1e-6)."""
        o = self._o
        for i in np.nonzero(self._has)[0]:
            new = self._nd[i]
            if self.n_dims == 3:
                from swarmrl_b200.device import director_to_quat

                q = director_to_quat(new)[0]
                for a in range(4):
                    o.q[a][i] = q[a]
            else:
                old = o.directors[i]
                ang = np.arccos(np.dot(new, old))
                if ang > 1e-6:
                    cr = np.cross(old, new)
                    o.theta[i] += np.round(cr[2] / np.linalg.norm(cr)) * ang
