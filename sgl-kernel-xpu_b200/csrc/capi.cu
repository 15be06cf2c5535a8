// C-ABI housekeeping for libsglk_sm100a.so (see include/sglk_fp8.h).
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/sglk_fp8.h"

extern "C" int sglk_abi_version(void) { return SGLK_ABI_VERSION; }

extern "C" const char* sglk_error_string(int code) {
  switch (code) {
    case SGLK_OK: return "ok";
    case SGLK_E_NULL: return "null pointer argument";
    case SGLK_E_DTYPE: return "unsupported dtype";
    case SGLK_E_GROUP: return "unsupported group_size (supported: 16, 32, 64, 128, 256, 512)";
    case SGLK_E_SHAPE: return "inconsistent or unsupported shape";
    case SGLK_E_STRIDE: return "unsupported stride or alignment";
    case SGLK_E_LAYOUT: return "unknown or inconsistent scale layout";
    case SGLK_E_ARG: return "invalid argument";
    case SGLK_E_DRIVER: return "cuTensorMapEncodeTiled unavailable or failed (needs an sm_100 CUDA driver)";
    case SGLK_E_VARIANT: return "unknown GEMM kernel variant";
    case SGLK_E_CAPACITY: return "problem exceeds the fixed capacity of the library's built-in scratch array (use the entry point that takes a workspace)";
    default: break;
  }
  if (code > 0) return cudaGetErrorString(static_cast<cudaError_t>(code));
  return "unknown sglk error";
}
