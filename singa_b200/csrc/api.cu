// Library-wide C-ABI helpers: version and last-error string.
#include "efl_common.cuh"
#include <string.h>

static thread_local char g_efl_error[512] = "";

void efl_set_error(const char* msg) {
  strncpy(g_efl_error, msg ? msg : "", sizeof(g_efl_error) - 1);
  g_efl_error[sizeof(g_efl_error) - 1] = 0;
}

extern "C" const char* efl_last_error_string(void) { return g_efl_error; }
extern "C" int efl_version(void) { return 100; }

// sm_100a only: report the device's compute capability so the host side can fail loudly elsewhere
extern "C" int efl_device_cc(int device, int* major, int* minor) {
  cudaDeviceProp p;
  cudaError_t e = cudaGetDeviceProperties(&p, device);
  if (e != cudaSuccess) {
    efl_set_error(cudaGetErrorString(e));
    return (int)e;
  }
  *major = p.major;
  *minor = p.minor;
  return EFL_OK;
}
