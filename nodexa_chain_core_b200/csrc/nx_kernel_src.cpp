// The text of kawpow_kernel.cuh, embedded at build time (see Makefile: kawpow_kernel_src.inc) so
// that NVRTC can compile it behind the per-period program at run time.
#include "nx_internal.hpp"

namespace nx
{
const char* kawpow_kernel_skeleton()
{
    static const char text[] =
#include "kawpow_kernel_src.inc"
        ;
    return text;
}
}  // namespace nx
