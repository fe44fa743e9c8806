// TEST INFRASTRUCTURE — NOT PRODUCT CODE.  Reference ur5 with instrumented primitives (see vrefc_instr.hh).
#include "vrefc_instr.hh"
#include <vamp/robots/ur5.hh>
#include "vrefc_impl.hh"
namespace
{
    using namespace vamp::robots;
    vrefc::Registrar reg("ur5", vrefc::Impl<vamp::robots::UR5>::vt());
}  // namespace
