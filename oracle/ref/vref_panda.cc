// TEST INFRASTRUCTURE — NOT PRODUCT CODE.  Reference Panda (robots/panda_grid.hh) behind vref_api.h.
#include <vamp/robots/panda_grid.hh>
#include "vref_impl.hh"
namespace
{
    using namespace vamp::robots;
    vref::Registrar r00("panda_0_0", vref::Impl<Panda_0_0>::vt());
    vref::Registrar r10("panda_1_0", vref::Impl<Panda_1_0>::vt());
    vref::Registrar r22("panda_2_2", vref::Impl<Panda_2_2>::vt());
}  // namespace
