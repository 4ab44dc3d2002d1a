// ABI version and error strings.
#include "../../include/gp_b200.h"
#include <cuda_runtime.h>

extern "C" int gp_abi_version(void) { return 1; }

extern "C" const char* gp_error_string(int code) {
  switch (code) {
    case GP_OK: return "ok";
    case GP_ERR_BAD_ARG: return "gp_b200: bad argument";
    case GP_ERR_WORKSPACE_TOO_SMALL: return "gp_b200: workspace too small";
    case GP_ERR_UNSUPPORTED: return "gp_b200: unsupported configuration";
    case GP_ERR_LIMIT: return "gp_b200: size beyond a fixed limit of this entry point";
    default: break;
  }
  if (code > 0) return cudaGetErrorString((cudaError_t)code);
  return "gp_b200: unknown error";
}
