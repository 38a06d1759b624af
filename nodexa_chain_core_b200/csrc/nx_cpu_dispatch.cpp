// Run-time CPU feature test, kept in its own translation unit compiled WITHOUT -mavx512f so that asking the
// question can never itself execute an AVX-512 instruction.
#include <cstdlib>

namespace nx
{
bool cpu_has_avx512f()
{
    static const bool yes = [] {
        if (const char* e = std::getenv("NX_NO_AVX512"))
            if (*e && *e != '0') return false;
#if defined(__x86_64__) || defined(__i386__)
        __builtin_cpu_init();
        return __builtin_cpu_supports("avx512f") != 0;
#else
        return false;
#endif
    }();
    return yes;
}
}  // namespace nx
