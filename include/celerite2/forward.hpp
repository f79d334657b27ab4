// c++/include/celerite2/forward.hpp under the reference's name: to_dense, factor, solve_lower/upper, matmul_lower/upper,
// general_matmul_lower/upper with the workspace argument.  All of celerite2::core::* lives in one header here.
#ifndef CELERITE2_B200_SHIM_FORWARD_HPP_
#define CELERITE2_B200_SHIM_FORWARD_HPP_
#include "../celerite2_b200/core.hpp"
#endif
