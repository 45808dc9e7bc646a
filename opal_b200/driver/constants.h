// constants.h -- the physical constants of src/constants.rs:4-28 (7-digit roundings, kept literally: parity quirk Q6)
#pragma once

namespace opal_driver {
namespace constants {

constexpr double SPEED_OF_LIGHT = 2.997925e8;                 // constants.rs:4
constexpr double SPEED_OF_LIGHT_SQD = 89875517873681764.0;    // constants.rs:6 (not the square of the line above)
constexpr double VACUUM_PERMITTIVITY = 8.854188e-12;          // constants.rs:8
constexpr double VACUUM_PERMEABILITY = 1.256637e-6;           // constants.rs:10
constexpr double ELECTRON_CHARGE = -1.602177e-19;             // constants.rs:12
constexpr double ELEMENTARY_CHARGE = -ELECTRON_CHARGE;        // constants.rs:14
constexpr double ELECTRON_MASS = 9.109383e-31;                // constants.rs:16
constexpr double PROTON_MASS = 1.672622e-27;                  // constants.rs:18
constexpr double ELECTRON_MASS_MEV = 0.510999;                // constants.rs:20

}  // namespace constants
}  // namespace opal_driver
