// Library-level entry points of the C ABI (include/te_b200.h).
#include "common.cuh"

extern "C" int te_abi_version(void) { return TE_B200_ABI_VERSION; }

extern "C" const char* te_status_string(int status) {
  switch (status) {
    case TE_OK:
      return "ok";
    case TE_ERR_INVALID_ARGUMENT:
      return "invalid argument";
    case TE_ERR_SCRATCH_TOO_SMALL:
      return "scratch buffer missing or too small (ask te_*_scratch_bytes)";
    case TE_ERR_UNSUPPORTED:
      return "unsupported configuration";
    case TE_ERR_STATIC_SCRATCH_TOO_SMALL:
      return "problem size exceeds the built-in static scratch of this no-allocation entry point";
    default:
      return status > 0 ? cudaGetErrorString((cudaError_t)status) : "unknown status";
  }
}
