// TEST-ONLY STUB (not nigh): declarations that let the reference's planning/nn.hh parse (see ../Eigen/Dense).
#pragma once
#include <cstddef>
namespace unc::robotics::nigh
{
    namespace metric
    {
        template <int P>
        struct LP
        {
        };
        template <typename T, typename M>
        struct Space;
    }  // namespace metric
    struct NoThreadSafety
    {
    };
    template <std::size_t B>
    struct KDTreeBatch
    {
    };
    template <typename T, typename S, typename K, typename C, typename A>
    struct Nigh;
}  // namespace unc::robotics::nigh
