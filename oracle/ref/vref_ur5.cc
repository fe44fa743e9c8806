// TEST INFRASTRUCTURE — NOT PRODUCT CODE.  Reference UR5 (robots/ur5.hh) behind vref_api.h.
#include <vamp/robots/ur5.hh>
#include "vref_impl.hh"
namespace
{
    vref::Registrar reg("ur5", vref::Impl<vamp::robots::UR5>::vt());
}  // namespace
