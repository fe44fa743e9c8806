// TEST INFRASTRUCTURE — NOT PRODUCT CODE.  Reference sphere robot (robots/sphere.hh) behind vref_api.h.
#include <vamp/robots/sphere.hh>
#include "vref_impl.hh"
namespace
{
    vref::Registrar reg("sphere", vref::Impl<vamp::robots::Sphere>::vt());
}  // namespace
extern "C" void vref_sphere_robot_set_radius(float r)
{
    vamp::robots::sphere::set_radius(r);
}
