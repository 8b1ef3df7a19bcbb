// Library-level entry points of libmagicsoup_b200.so (see include/magicsoup_b200.h).
#include <stdio.h>
#include <string.h>

#include "ms_common.cuh"

static thread_local char g_cuda_err[256] = "";

extern "C" void ms_set_cuda_error_(cudaError_t e) {
  const char *name = cudaGetErrorName(e), *msg = cudaGetErrorString(e);
  snprintf(g_cuda_err, sizeof(g_cuda_err), "%s: %s", name ? name : "?", msg ? msg : "?");
  cudaGetLastError();  // clear the non-sticky error state
}

extern "C" int ms_version(void) { return 100; }

extern "C" const char *ms_last_cuda_error(void) { return g_cuda_err; }

extern "C" const char *ms_error_string(int code) {
  switch (code) {
    case MS_OK: return "ok";
    case MS_ERR_INVALID_ARGUMENT: return "invalid argument";
    case MS_ERR_CUDA: return "CUDA error (see ms_last_cuda_error)";
    case MS_ERR_SCRATCH_TOO_SMALL: return "scratch buffer missing or too small";
    case MS_ERR_NO_DEVICE: return "no CUDA device";
    case MS_ERR_TOO_LARGE: return "problem size beyond the supported maximum";
    default: return "unknown error";
  }
}

extern "C" int ms_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  return n;
}
