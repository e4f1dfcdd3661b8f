"""Lattice-unit <-> physical-unit conversion; same method names and formulas as
`lettuce.UnitConversion` (reference/lettuce/_unit.py:13-145).  Host-side scalars only."""
import numpy as np

__all__ = ['UnitConversion']


class UnitConversion:
    def __init__(self, reynolds_number, mach_number=0.05, characteristic_length_pu=1,
                 characteristic_velocity_pu=1, characteristic_length_lu=1,
                 characteristic_density_lu=1, characteristic_density_pu=1, cs=1 / np.sqrt(3.0)):
        self.cs = cs
        self.reynolds_number = reynolds_number
        self.mach_number = mach_number
        self.characteristic_length_pu = characteristic_length_pu
        self.characteristic_velocity_pu = characteristic_velocity_pu
        self.characteristic_length_lu = characteristic_length_lu
        self.characteristic_density_lu = characteristic_density_lu
        self.characteristic_density_pu = characteristic_density_pu

    # derived scales (_unit.py:34-60)
    @property
    def characteristic_velocity_lu(self):
        return self.cs * self.mach_number

    @property
    def characteristic_pressure_pu(self):
        return self.characteristic_density_pu * self.characteristic_velocity_pu ** 2

    @property
    def characteristic_pressure_lu(self):
        return self.characteristic_density_lu * self.characteristic_velocity_lu ** 2

    @property
    def viscosity_lu(self):
        return self.characteristic_length_lu * self.characteristic_velocity_lu / self.reynolds_number

    @property
    def viscosity_pu(self):
        return self.characteristic_length_pu * self.characteristic_velocity_pu / self.reynolds_number

    @property
    def relaxation_parameter_lu(self):
        return self.viscosity_lu / self.cs ** 2 + 0.5

    # every conversion is "divide by the source scale, multiply by the target scale",
    # in that order (_unit.py:62-145) -- the order matters for bit-level parity of boundary values
    @staticmethod
    def _scale(value, src, dst):
        return value / src * dst

    def convert_velocity_to_pu(self, v):
        return self._scale(v, self.characteristic_velocity_lu, self.characteristic_velocity_pu)

    def convert_velocity_to_lu(self, v):
        return self._scale(v, self.characteristic_velocity_pu, self.characteristic_velocity_lu)

    def _accel_scales(self):
        return (self.characteristic_velocity_lu ** 2 / self.characteristic_length_lu,
                self.characteristic_velocity_pu ** 2 / self.characteristic_length_pu)

    def convert_acceleration_to_pu(self, a):
        x, y = self._accel_scales()
        return self._scale(a, x, y)

    def convert_acceleration_to_lu(self, a):
        x, y = self._accel_scales()
        return self._scale(a, y, x)

    def _time_scales(self):
        return (self.characteristic_length_lu / self.characteristic_velocity_lu,
                self.characteristic_length_pu / self.characteristic_velocity_pu)

    def convert_time_to_pu(self, t):
        x, y = self._time_scales()
        return self._scale(t, x, y)

    def convert_time_to_lu(self, t):
        x, y = self._time_scales()
        return self._scale(t, y, x)

    def convert_density_lu_to_pressure_pu(self, density_lu):
        return self.convert_pressure_to_pu((density_lu - self.characteristic_density_lu) * self.cs ** 2)

    def convert_pressure_pu_to_density_lu(self, pressure_pu):
        return self.convert_pressure_to_lu(pressure_pu) / self.cs ** 2 + self.characteristic_density_lu

    def convert_density_to_pu(self, rho):
        return self._scale(rho, self.characteristic_density_lu, self.characteristic_density_pu)

    def convert_density_to_lu(self, rho):
        return self._scale(rho, self.characteristic_density_pu, self.characteristic_density_lu)

    def convert_pressure_to_pu(self, p):
        return self._scale(p, self.characteristic_pressure_lu, self.characteristic_pressure_pu)

    def convert_pressure_to_lu(self, p):
        return self._scale(p, self.characteristic_pressure_pu, self.characteristic_pressure_lu)

    def convert_length_to_pu(self, length):
        return length * self.characteristic_length_pu / self.characteristic_length_lu

    def convert_length_to_lu(self, length):
        return length * self.characteristic_length_lu / self.characteristic_length_pu

    def convert_energy_to_pu(self, e):
        """energy in units of [density * velocity**2]"""
        return e * self.characteristic_pressure_pu / self.characteristic_pressure_lu

    def convert_energy_to_lu(self, e):
        return e * self.characteristic_pressure_lu / self.characteristic_pressure_pu

    def convert_incompressible_energy_to_pu(self, e):
        """energy in units of [velocity**2]"""
        return e * (self.characteristic_velocity_pu ** 2) / (self.characteristic_velocity_lu ** 2)

    def convert_incompressible_energy_to_lu(self, e):
        return e * (self.characteristic_velocity_lu ** 2) / (self.characteristic_velocity_pu ** 2)
