// forwards the reference include path to the nx drop-in
#include "../../../../nx/ethash_compat.hpp"
