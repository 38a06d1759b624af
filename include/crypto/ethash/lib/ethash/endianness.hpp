// forwards the reference include path to the nx drop-in (ethash::le lives in ethash_compat.hpp)
#include "../../../../nx/ethash_compat.hpp"
