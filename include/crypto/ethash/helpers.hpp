// forwards the reference include path to the nx drop-in
#include "../../nx/helpers_compat.hpp"
