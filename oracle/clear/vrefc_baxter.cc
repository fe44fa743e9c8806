// TEST INFRASTRUCTURE — NOT PRODUCT CODE.  Reference baxter with instrumented primitives (see vrefc_instr.hh).
#include "vrefc_instr.hh"
#include <vamp/robots/baxter.hh>
#include "vrefc_impl.hh"
namespace
{
    using namespace vamp::robots;
    vrefc::Registrar reg("baxter", vrefc::Impl<vamp::robots::Baxter>::vt());
}  // namespace
