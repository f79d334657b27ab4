// c++/include/celerite2/core.hpp:4-6 under the reference's name: forward + interface + reverse operators.
#ifndef CELERITE2_B200_SHIM_CORE_HPP_
#define CELERITE2_B200_SHIM_CORE_HPP_
#include "forward.hpp"
#include "interface.hpp"
#include "reverse.hpp"
#endif
