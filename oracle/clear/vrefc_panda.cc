// TEST INFRASTRUCTURE — NOT PRODUCT CODE.  Reference panda with instrumented primitives (see vrefc_instr.hh).
#include "vrefc_instr.hh"
#include <vamp/robots/panda_grid.hh>
#include "vrefc_impl.hh"
namespace
{
    using namespace vamp::robots;
    vrefc::Registrar r00("panda_0_0", vrefc::Impl<Panda_0_0>::vt());
    vrefc::Registrar r10("panda_1_0", vrefc::Impl<Panda_1_0>::vt());
    vrefc::Registrar r22("panda_2_2", vrefc::Impl<Panda_2_2>::vt());
}  // namespace
