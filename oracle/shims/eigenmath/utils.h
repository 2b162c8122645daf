// TEST INFRASTRUCTURE ONLY.  Stand-in for eigenmath/utils.h (nothing from it is used on the hot path).
#ifndef ORACLE_SHIMS_EIGENMATH_UTILS_H_
#define ORACLE_SHIMS_EIGENMATH_UTILS_H_
#include "eigenmath/types.h"
#endif
