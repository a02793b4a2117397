"""Enumerations whose VALUES are baked into the kernels (csrc/lattice.cuh) -- same names and
values as the reference's utils/types.py:5-79."""
from enum import Enum


class _IntLike(Enum):
    def __int__(self):
        return self.value


class CellType(_IntLike):  # compared with ==, never as bit sets (propagation.cu)
    NOTHING = 0
    FLUID = 1
    OBSTACLE = 2
    EMPTY = 4
    INFLOW = 8
    OUTFLOW = 16
    INFLOW_2 = 32


class KBCType(_IntLike):
    LBGK = 0b00000000
    KBC_A = 0b10000101
    KBC_B = 0b10000110
    KBC_C = 0b10001001
    KBC_D = 0b10001010

    @staticmethod
    def _has(code, mask):
        return code is not None and (code & mask) > 0

    @staticmethod
    def is_KBC(input):
        return KBCType._has(input, 0b10000000)

    @staticmethod
    def is_KBC_AC(input):
        return KBCType._has(input, 0b10000001)

    @staticmethod
    def is_KBC_BD(input):
        return KBCType._has(input, 0b10000010)

    @staticmethod
    def is_KBC_AB(input):
        return KBCType._has(input, 0b10000100)

    @staticmethod
    def is_KBC_CD(input):
        return KBCType._has(input, 0b10001000)


class ObsType(_IntLike):
    BOX = 0
    SPHERE = 1


class AxiSymmetricType(_IntLike):
    NOT = 0
    LINE_X_EQ_0 = 1
    LINE_Y_EQ_0 = 2
    LINE_Z_EQ_0 = 3
