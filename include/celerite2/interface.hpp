// c++/include/celerite2/interface.hpp:37-220 under the reference's name: the seven overloads without a workspace.
#ifndef CELERITE2_B200_SHIM_INTERFACE_HPP_
#define CELERITE2_B200_SHIM_INTERFACE_HPP_
#include "../celerite2_b200/core.hpp"
#endif
