// TEST INFRASTRUCTURE — NOT PRODUCT CODE.  Reference sphere with instrumented primitives (see vrefc_instr.hh).
#include "vrefc_instr.hh"
#include <vamp/robots/sphere.hh>
#include "vrefc_impl.hh"
namespace
{
    using namespace vamp::robots;
    vrefc::Registrar reg("sphere", vrefc::Impl<vamp::robots::Sphere>::vt());
}  // namespace
