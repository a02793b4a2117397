"""Parameter bag of a run; call-compatible with the reference's
simulation/simulation_parameters.py:6-87 (same keywords, defaults, attributes and clock logic)."""
import math

import torch


class SimulationParameters(object):
    _PHYSICS = ("density_gas", "density_fluid", "gravity_strength", "axisymmetric_type", "contact_angle", "Q", "tau",
                "rho_gas", "rho_fluid", "kappa", "tau_g", "tau_f", "k")

    def __init__(self, dtype=torch.float32, dim=2, dt=1.0, device=torch.device("cpu"), simulation_size=(1, 1, 256, 256),
                 density_gas=0.0, density_fluid=1.0, gravity_strength=0.0, axisymmetric_type=0,
                 contact_angle=0.75 * math.pi, Q=9, tau=1.0, rho_gas=0.038, rho_fluid=0.265, kappa=0.08, tau_g=0.7,
                 tau_f=0.7, k=1.0):
        if dim == 2 and len(simulation_size) != 4:
            raise ValueError("For 2d simulation simulation size should have 4 parameters B x C x H x W")
        given = locals()
        self.dim, self.dtype, self.dt, self.device = dim, dtype, dt, device
        self.simulation_size = simulation_size
        for name in self._PHYSICS:
            setattr(self, name, given[name])
        # frame clock (simulation_parameters.py:33-36)
        self.frame = 0
        self.time_per_frame = 0
        self.frame_length = 1.0
        self.time_total = 0

    def step(self):
        """Advance the clock by dt; a frame closes every `frame_length` (simulation_parameters.py:61-71)."""
        self.time_per_frame += self.dt
        self.time_total += self.dt
        if self.time_per_frame >= self.frame_length:
            self.frame += 1
            self.time_total = self.frame * self.frame_length  # re-derived to prevent drift
            self.time_per_frame = 0

    def get_dx(self):
        return 1.0 / max(self.simulation_size)

    def is_2d(self):
        return self.dim == 2

    def is_3d(self):
        return self.dim == 3

    def set_device(self, device="cuda"):
        if device not in ["cuda", "cpu"]:
            raise ValueError("Set_device: device {} must be either cuda or cpu.".format(device))
        self.device = torch.device(device)
