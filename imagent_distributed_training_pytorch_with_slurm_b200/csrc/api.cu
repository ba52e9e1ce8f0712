// Library identity and error text for the C ABI.
#include "common.cuh"

extern "C" int b200_abi_version(void) { return B200_ABI_VERSION; }

extern "C" const char *b200_error_string(int code) { return cudaGetErrorString(static_cast<cudaError_t>(code)); }
