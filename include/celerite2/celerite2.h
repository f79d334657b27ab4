// Umbrella header under the reference's name (c++/include/celerite2/celerite2.h:4-6): core operators + terms, served by
// libcelerite2_b200.so.  Put <repo>/include ahead of the reference's include directory and link -lcelerite2_b200.
#ifndef CELERITE2_B200_SHIM_CELERITE2_H_
#define CELERITE2_B200_SHIM_CELERITE2_H_
#include "core.hpp"
#include "terms.hpp"
#endif
