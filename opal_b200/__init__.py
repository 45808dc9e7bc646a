"""opal_b200 -- B200-native hot path of the opal 1d3v PIC code (host-side mirror of the
reference's loop interface above the C ABI in include/opal_b200.h)."""
from . import abi, constants  # noqa: F401
from .abi import (BDY_ABSORBING, BDY_CONDUCTING, BDY_INTERNAL, BDY_LASER, ELECTRON, ION, PHOTON, LEFT, RIGHT,  # noqa: F401
                  FLAG_LEAN_STATE, FLAG_STRICT_FP, OpalABI, OpalError, Slab)
