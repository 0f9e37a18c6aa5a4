#include "sevenn_common.cuh"
#include "../../include/sevenn_b200.h"
extern "C" int sevenn_abi_version(void) { return 1; }
