// Error plumbing, version and device queries of libgenseg_b200.
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>

#include "common.cuh"

namespace gsb {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

static unsigned long long g_launches = 0;
unsigned long long launch_count() { return g_launches; }

int check_launch(const char* what, int n) {
  g_launches += (unsigned long long)n;
  cudaError_t e = cudaPeekAtLastError();
  if (e != cudaSuccess) {
    cudaGetLastError();
    set_error("%s: CUDA error: %s", what, cudaGetErrorString(e));
    return -3;
  }
  return 0;
}

bool pdl_enabled() {
  static int on = -1;
  if (on < 0) {
    const char* e = getenv("GSB_PDL");
    on = (e && e[0] == '1') ? 1 : 0;   // off by default: measured neutral (profiles/README.md, r01r)
  }
  return on == 1;
}

int num_sms() {
  static int sms = 0;
  if (sms == 0) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess ||
        sms <= 0)
      sms = 148;
  }
  return sms;
}

// SMs the persistent kernels size their grids for: all of them, minus the ones set aside for a collective that runs
// concurrently on another stream (gsb_set_sm_reserve; see parallel.py -- the gradient all-reduce overlapped with the
// encoder backward).  A persistent grid that does not fit beside the collective's CTAs would run its tail as a second wave.
static int g_sm_reserve = 0;
int num_sms_avail() {
  const int n = num_sms() - g_sm_reserve;
  return n < 8 ? 8 : n;
}

}  // namespace gsb

extern "C" unsigned long long gsb_launch_count(void) { return gsb::launch_count(); }
extern "C" int gsb_version(void) { return 100; }
extern "C" const char* gsb_last_error(void) { return gsb::g_err; }
extern "C" int gsb_tc_available(void) {
  int dev = 0, major = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 0;
  if (cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev) != cudaSuccess) return 0;
  return major == 10 ? 1 : 0;
}

extern "C" int gsb_set_sm_reserve(int n) {
  const int prev = gsb::g_sm_reserve;
  gsb::g_sm_reserve = n < 0 ? 0 : (n > 64 ? 64 : n);
  return prev;
}
