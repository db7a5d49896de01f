// core.cu -- error reporting, launch accounting (bench.py's gpu_launches), version.
#include <cstdarg>
#include <cstdio>
#include <atomic>
#include <string>

#include "common.cuh"

namespace b200lab {
static thread_local char g_err[512] = "";
static std::atomic<int64_t> g_launches{0};

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

int check_launch(const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("%s: CUDA launch failed: %s", what, cudaGetErrorString(e));
    return B200LAB_ERR_CUDA;
  }
  return B200LAB_OK;
}
}  // namespace b200lab

extern "C" {
const char* b200lab_last_error(void) { return b200lab::g_err; }
int b200lab_version(void) { return B200LAB_VERSION; }
int64_t b200lab_launch_count(void) { return b200lab::g_launches.load(); }
}

// struct sizes, checked by the Python binding (which generates its ctypes mirrors from include/b200lab.h)
extern "C" {
int64_t b200lab_sizeof(const char* name) {
  std::string n(name);
  if (n == "B200LabTerm") return sizeof(B200LabTerm);
  if (n == "B200LabActuatorGroup") return sizeof(B200LabActuatorGroup);
  if (n == "B200LabSimView") return sizeof(B200LabSimView);
  if (n == "B200LabEnvState") return sizeof(B200LabEnvState);
  if (n == "B200LabTransition") return sizeof(B200LabTransition);
  if (n == "B200LabGather") return sizeof(B200LabGather);
  if (n == "B200LabPpoLoss") return sizeof(B200LabPpoLoss);
  return -1;
}
}
