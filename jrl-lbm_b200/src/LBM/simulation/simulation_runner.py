"""Factory for the solver objects; call-compatible with the reference's
simulation/simulation_runner.py:16-126.  3-D is not implemented there either (:31-32)."""
from src.LBM.LBM_collision import LBMCollision2d, LBMCollisionKBC2d, LBMCollisionMRT2d
from src.LBM.LBM_macro_compute import LBMMacroCompute2d
from src.LBM.LBM_propagation import LBMPropagation2d
from src.LBM.LBM_solid_fluid_coupling import LBMBoundary2d


class SimulationRunner(object):
    def __init__(self, parameters):
        self.parameters = parameters

    def _need_2d(self, what, exc=RuntimeError):
        if not self.parameters.is_2d():
            raise exc("Not Implemented for 3D " + what)

    def _common(self):
        p = self.parameters
        return dict(tau=p.tau, axisymmetric_type=p.axisymmetric_type, dtype=p.dtype, device=p.device)

    def _phase(self):
        p = self.parameters
        return dict(Q=p.Q, density_liquid=p.density_fluid, density_gas=p.density_gas, rho_liquid=p.rho_fluid,
                    rho_gas=p.rho_gas, contact_angle=p.contact_angle)

    def _collision(self, cls, what):
        self._need_2d(what)
        p = self.parameters
        return cls(gravity_strength=p.gravity_strength, kappa=p.kappa, tau_f=p.tau_f, tau_g=p.tau_g,
                   **self._phase(), **self._common())

    def create_propagation(self):
        self._need_2d("propagation")
        return LBMPropagation2d(**self._common())

    def create_macro_compute(self):
        self._need_2d("macro computation")
        return LBMMacroCompute2d(**self._phase(), **self._common())

    def create_collision(self):
        return self._collision(LBMCollision2d, "collision")

    def create_collision_KBC(self):
        return self._collision(LBMCollisionKBC2d, "KBC collision")

    def create_collision_MRT(self):
        return self._collision(LBMCollisionMRT2d, "MRT collision")

    def create_LBM_fluid_solid_coupling(self):
        if not self.parameters.is_2d():
            raise NotImplementedError("3D Immersed boundary not implemented")
        return LBMBoundary2d(**self._common())

    def create_fused_step(self, **kw):
        """Extension: the single-launch lattice update (jrl_step.FusedStep)."""
        from jrl_step import FusedStep
        return FusedStep(self.parameters, **kw)

    def step(self):
        self.parameters.step()
