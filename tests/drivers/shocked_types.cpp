// The type composition of demo/laminar/shocked.hpp:43-111 against include/mini: EnergyBasedViscosity over the
// Navier-Stokes Riemann type selects MCFD_NAVIER_STOKES_CELL_VISCOSITY_DDG.  Compile-only (tests/test_host_logic.py).
using Scalar = double;
#include "mini/integrator/lobatto.hpp"
#include "mini/polynomial/hexahedron.hpp"
#include "mini/mesh/part.hpp"
#include "mini/riemann/euler/types.hpp"
#include "mini/riemann/euler/hllc.hpp"
#include "mini/riemann/rotated/euler.hpp"
#include "mini/riemann/concept.hpp"
#include "mini/riemann/diffusive/navier_stokes.hpp"
#include "mini/riemann/diffusive/linear.hpp"
#include "mini/riemann/diffusive/direct.hpp"
#include "mini/spatial/fr/lobatto.hpp"
#include "mini/spatial/viscosity.hpp"
#include "mini/spatial/with_viscosity.hpp"

using Gx = mini::integrator::Lobatto<Scalar, 3>;
using Polynomial = mini::polynomial::Hexahedron<Gx, Gx, Gx, 5, false>;
using Part = mini::mesh::part::Part<int64_t, Polynomial>;
using Gas = mini::riemann::euler::IdealGas<Scalar, 1.4>;
using Convection = mini::riemann::rotated::Euler<mini::riemann::euler::HartenLaxLeerContact<Gas, 3>>;
using Diffusion = mini::riemann::diffusive::Direct<mini::riemann::diffusive::NavierStokes<Gas>>;
using Riemann = mini::riemann::ConvectionDiffusion<Convection, Diffusion>;
using RiemannWithViscosity = mini::spatial::EnergyBasedViscosity<Part, Riemann>;
using General = mini::spatial::fr::Lobatto<Part, RiemannWithViscosity>;
using Spatial = mini::spatial::WithViscosity<General>;
static_assert(RiemannWithViscosity::kDiffusion == MCFD_NAVIER_STOKES_CELL_VISCOSITY_DDG);
using Plain = mini::spatial::EnergyBasedViscosity<Part, mini::riemann::ConvectionDiffusion<Convection,
    mini::riemann::diffusive::Direct<mini::riemann::diffusive::Isotropic<Scalar, 5>>>>;
static_assert(Plain::kDiffusion == MCFD_CELL_VISCOSITY_DDG);
Spatial *Make(Part *part) { return new Spatial(part); }
