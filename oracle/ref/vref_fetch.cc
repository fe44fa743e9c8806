// TEST INFRASTRUCTURE — NOT PRODUCT CODE.  Reference Fetch (robots/fetch.hh) behind vref_api.h.
#include <vamp/robots/fetch.hh>
#include "vref_impl.hh"
namespace
{
    vref::Registrar reg("fetch", vref::Impl<vamp::robots::Fetch>::vt());
}  // namespace
