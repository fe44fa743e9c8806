// TEST INFRASTRUCTURE — NOT PRODUCT CODE.  Reference fetch with instrumented primitives (see vrefc_instr.hh).
#include "vrefc_instr.hh"
#include <vamp/robots/fetch.hh>
#include "vrefc_impl.hh"
namespace
{
    using namespace vamp::robots;
    vrefc::Registrar reg("fetch", vrefc::Impl<vamp::robots::Fetch>::vt());
}  // namespace
