// TEST INFRASTRUCTURE — NOT PRODUCT CODE.  Reference Baxter (robots/baxter.hh) behind vref_api.h.
#include <vamp/robots/baxter.hh>
#include "vref_impl.hh"
namespace
{
    vref::Registrar reg("baxter", vref::Impl<vamp::robots::Baxter>::vt());
}  // namespace
