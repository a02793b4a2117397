"""Test infrastructure (oracle/): the numpy oracle behind the reference's solver-object API.

Lets jrl-lbm_b200/jrl_scenarios.py (written once, against the reference's class and method
names) run on the oracle: tensors in and out are CPU torch tensors, the arithmetic is
oracle/lbm_oracle.py.  Method signatures follow SURVEY.md section 8b; each method cites the
reference method it stands in for.  Not product code.
"""
from __future__ import annotations

import enum
import math
from types import SimpleNamespace

import numpy as np
import torch

import lbm_oracle as O


class CellType(enum.IntEnum):  # utils/types.py:5-15
    NOTHING = 0
    FLUID = 1
    OBSTACLE = 2
    EMPTY = 4
    INFLOW = 8
    OUTFLOW = 16
    INFLOW_2 = 32


class KBCType(enum.IntEnum):  # utils/types.py:18-23
    LBGK = 0
    KBC_A = 0b10000101
    KBC_B = 0b10000110
    KBC_C = 0b10001001
    KBC_D = 0b10001010


class AxiSymmetricType(enum.IntEnum):  # utils/types.py:69-73
    NOT = 0
    LINE_X_EQ_0 = 1
    LINE_Y_EQ_0 = 2
    LINE_Z_EQ_0 = 3


def _n(t):
    if t is None:
        return None
    if isinstance(t, torch.Tensor):
        return np.ascontiguousarray(t.detach().cpu().numpy())
    return t


def _t(a):
    return torch.from_numpy(np.ascontiguousarray(a))


def create_2d_meshgrid_tensor(size, device=None, dtype=torch.float32):
    B, _, H, W = size
    return _t(O.meshgrid(B, H, W)).to(dtype)


def UnionPhiObs(p1, p2, alpha=0.0):
    return _t(O.union_phi_obs(_n(p1), _n(p2), alpha))


class SimulationParameters:  # simulation_parameters.py:6-59 (only what the solver objects read)
    def __init__(self, dtype=torch.float32, dim=2, dt=1.0, device=None, simulation_size=(1, 1, 256, 256),
                 density_gas=0.0, density_fluid=1.0, gravity_strength=0.0, axisymmetric_type=0,
                 contact_angle=0.75 * math.pi, Q=9, tau=1.0, rho_gas=0.038, rho_fluid=0.265, kappa=0.08,
                 tau_g=0.7, tau_f=0.7, k=1.0):
        self.__dict__.update(locals())

    def step(self):
        pass


class _Prop:
    def __init__(self, p):
        self.axisymmetric_type = p.axisymmetric_type

    def propagation(self, f, flags, phi_obs=None, rho=None, vel=None, inflow_vel=None, inflow_density=None,
                    is_convection=False):
        # LBM_propagation_2d.py:35-92 incl. the return-arity rule (:52-59, 89-92)
        inflow_info = inflow_density is not None or inflow_vel is not None
        if inflow_info:
            assert rho is not None
        if bool((flags == int(CellType.INFLOW_2)).any()):
            inflow_info = True
        rho_n, vel_n = _n(rho), _n(vel)
        f_new = O.stream(_n(flags), _n(f), _n(phi_obs), rho_n, vel_n, _n(inflow_vel),
                         -1.0 if inflow_density is None else float(inflow_density), self.axisymmetric_type)
        if rho is not None:
            rho.copy_(_t(rho_n))   # in-place mutation semantics, propagation.cu:300, 362
        if vel is not None:
            vel.copy_(_t(vel_n))
        if inflow_info:
            return (rho, vel, _t(f_new))
        return _t(f_new)

    def rebounce_obstacle(self, f, flags):
        return _t(O.rebounce_obstacle(_n(f), _n(flags)))


class _Macro:
    def __init__(self, p):
        self.axisymmetric_type = p.axisymmetric_type
        ca = p.contact_angle
        self._contact_angle = None if ca is None else (np.float32(ca.item()) if isinstance(ca, torch.Tensor) else ca)

    def macro_compute(self, dx, dt, f, rho, vel, flags, g=None, pressure=None, density=None):
        r, v = O.macro_compute(dx, dt, _n(f), _n(rho), _n(vel), _n(flags), self._contact_angle)
        return [_t(r), _t(v)]

    def macro_compute_C(self, dx, dt, h, C, flags, vel, mesh_grid):
        return _t(O.macro_compute_C(_n(h), _n(C), _n(flags)))

    def ApplyCompute_fluid(self, dx, dt, f, rho, vel, flags, g=None, pressure=None, density=None):
        r, v = O.macro_compute_kernel(dx, dt, _n(f), _n(rho), _n(vel), _n(flags))
        return [_t(r), _t(v)]

    def ApplyCompute_C(self, dx, dt, h, C, flags, vel, mesh_grid):
        return _t(O.macro_compute_C_kernel(dx, dt, _n(h), _n(C), _n(flags), _n(vel), _n(mesh_grid),
                                           self.axisymmetric_type))

    def get_vort(self, vel, dx):
        return _t(O.get_vort(_n(vel), dx))


class _Collision:
    kind = "bgk"

    def __init__(self, p):
        self.axisymmetric_type = p.axisymmetric_type
        self._tau = p.tau
        self._tau_D = None
        self._gravity = torch.Tensor([0.0, -p.gravity_strength]).reshape(1, 2, 1, 1).to(torch.float32)
        self._mat = None

    def preset_KBC(self, dx, dt, tau, tau_D=None):
        self._tau = tau
        if tau_D is not None:
            self._tau_D = tau_D
        if self.kind == "mrt":
            self._mat = O.mrt_matrix(tau)

    def get_feq_(self, dx, dt, rho, vel, tau=None, force=None, is_convection=False):
        return _t(O.get_feq(dx, dt, _n(rho), _n(vel), _n(force)))

    def get_grad(self, rho, dx, flags):
        return _t(O.grad(_n(rho), _n(flags), dx, self.axisymmetric_type))

    def get_div(self, vel, dx, flags):
        return _t(O.div(_n(vel), _n(flags), dx, self.axisymmetric_type))

    def collision(self, dx, dt, f, rho, vel, flags, force=None, mesh_grid=None, is_convection=False,
                  KBC_type=None, **_):
        a = (dx, dt, _n(f), _n(rho), _n(vel), _n(flags))
        if self.kind == "bgk":
            out = O.collide_bgk(*a, self._tau, _n(force), _n(mesh_grid), is_convection, self._tau_D,
                                self.axisymmetric_type)
        elif self.kind == "mrt":
            assert not is_convection
            out = O.collide_mrt(*a, self._tau, _n(force), _n(mesh_grid), self.axisymmetric_type, self._mat)
        else:
            out = O.collide_kbc(*a, self._tau, KBC_type, _n(force), _n(mesh_grid), self.axisymmetric_type)
        return _t(out)


class _CollisionMRT(_Collision):
    kind = "mrt"


class _CollisionKBC(_Collision):
    kind = "kbc"


class _Boundary:
    def __init__(self, p):
        self.axisymmetric_type = p.axisymmetric_type
        self._tau = p.tau

    def solve_boundary(self, dt, dx, rho, vel, flags, f, phi_obs, obs_vel, tau=None, is_convection=False):
        tau = self._tau if tau is None else tau
        return _t(O.wall(dt, dx, tau, _n(vel), _n(flags), _n(f), _n(phi_obs), _n(obs_vel),
                         self.axisymmetric_type))


class SimulationRunner:  # simulation_runner.py:16-126
    def __init__(self, parameters):
        self.parameters = parameters

    def create_propagation(self):
        return _Prop(self.parameters)

    def create_macro_compute(self):
        return _Macro(self.parameters)

    def create_collision(self):
        return _Collision(self.parameters)

    def create_collision_MRT(self):
        return _CollisionMRT(self.parameters)

    def create_collision_KBC(self):
        return _CollisionKBC(self.parameters)

    def create_LBM_fluid_solid_coupling(self):
        return _Boundary(self.parameters)

    def step(self):
        self.parameters.step()


def api():
    return SimpleNamespace(SimulationParameters=SimulationParameters, SimulationRunner=SimulationRunner,
                           CellType=CellType, KBCType=KBCType, AxiSymmetricType=AxiSymmetricType,
                           create_2d_meshgrid_tensor=create_2d_meshgrid_tensor, UnionPhiObs=UnionPhiObs)
