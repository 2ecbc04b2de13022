// abi.cu -- version / build information of libcollision_b200.so
#include "common.cuh"

COL_API int col_abi_version(void) { return 2; }
COL_API int col_built_arch(void) { return 100; }
