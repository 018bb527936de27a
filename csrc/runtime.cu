// Small runtime helpers shared by all entry points.
#include "common.cuh"

DGB_API const char* dgb_error_string(int code) { return cudaGetErrorName((cudaError_t)code); }

DGB_API int dgb_version() { return 100; }

// Number of SMs of the current device (grid sizing for persistent kernels).
DGB_API int dgb_sm_count() {
  int dev = 0, n = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return -1;
  if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return -1;
  return n;
}
