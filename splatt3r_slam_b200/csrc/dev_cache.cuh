// b200slam — per-DEVICE caches of launch-time facts (SM count, cooperative-launch support, "this
// kernel's shared-memory opt-in has been set").  Function attributes and device properties belong
// to a device, not to the process: a process that calls the library on cuda:1 after cuda:0 must
// repeat the opt-in there, so every cache below is indexed by the current device ordinal.
#pragma once
#include <cuda_runtime.h>

namespace b200dev {

constexpr int MAX_DEVICES = 64;

inline int current() {
  int d = 0;
  cudaGetDevice(&d);
  return (d >= 0 && d < MAX_DEVICES) ? d : 0;
}

struct Info {
  int num_sms;  // 0 = not queried yet
  int coop;
};

// SM count (>= 1) and cooperative-launch flag of the current device.
inline Info info() {
  static Info table[MAX_DEVICES] = {};
  const int d = current();
  if (table[d].num_sms == 0) {
    int sms = 0, coop = 0;
    if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, d) != cudaSuccess || sms < 1) sms = 148;
    cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, d);
    table[d].coop = coop;
    table[d].num_sms = sms;
  }
  return table[d];
}

// One flag per (call site, device): `static b200dev::Once once; if (once.first()) cudaFuncSetAttribute(...)`.
struct Once {
  bool done[MAX_DEVICES] = {};
  bool first() {
    const int d = current();
    if (done[d]) return false;
    done[d] = true;
    return true;
  }
  void retry() { done[current()] = false; }  // the guarded call failed: try again next time
};

// One small integer per (call site, device), 0 = unset.
struct PerDeviceInt {
  int v[MAX_DEVICES] = {};
  int &ref() { return v[current()]; }
};

}  // namespace b200dev
