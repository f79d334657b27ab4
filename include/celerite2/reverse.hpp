// c++/include/celerite2/reverse.hpp:10-217 under the reference's name: factor_rev, solve_*_rev, matmul_*_rev.
#ifndef CELERITE2_B200_SHIM_REVERSE_HPP_
#define CELERITE2_B200_SHIM_REVERSE_HPP_
#include "../celerite2_b200/core.hpp"
#endif
