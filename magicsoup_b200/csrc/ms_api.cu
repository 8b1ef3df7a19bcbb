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

// The host's reading of the early-exit flags (magicsoup_b200/kinetics.py:_settle_early_exit, reference
// kinetics.py:846-847) on the device: gate = 1 if the launch that wrote `flags` with eq_iters
// iterations per part already equals the reference (and met no cell outside the map), else 0.
__global__ void early_exit_verdict_kernel(const unsigned *flags, int e0, int e1, int e2, int n_parts, unsigned *gate) {
  const int eq[3] = {e0, e1, e2};
  unsigned ok = flags[MS_FLAG_BAD_POSITION] ? 0u : 1u;
  for (int q = 0; q < n_parts && ok; q++)
    for (int k = 0; k < eq[q]; k++)
      if (flags[q * 4 + k] == 0u) {  // no high-impact protein adjusting anywhere: the reference stops here ...
        if (flags[12 + q * 4 + k] != 0u) ok = 0u;  // ... but the launch went on adjusting low-impact ones
        break;
      }
  *gate = ok;
}

extern "C" int ms_early_exit_verdict(const uint32_t *flags, const int32_t *eq_iters_host, int32_t n_parts,
                                     uint32_t *gate, void *stream) {
  if (!flags || !gate || n_parts < 1 || n_parts > 3) return MS_ERR_INVALID_ARGUMENT;
  int eq[3] = {4, 4, 4};
  for (int q = 0; q < n_parts; q++) {
    if (eq_iters_host) eq[q] = eq_iters_host[q];
    if (eq[q] < 0 || eq[q] > 4) return MS_ERR_INVALID_ARGUMENT;
  }
  early_exit_verdict_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(flags, eq[0], eq[1], eq[2], n_parts, gate);
  MS_CUDA_CHECK(cudaGetLastError());
  return MS_OK;
}
