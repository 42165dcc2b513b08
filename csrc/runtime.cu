// Host-side runtime helpers: topology discovery (what Open MPI/Horovod learn at MPI_Init,
// horovod/tensorflow_mnist.py:90) — GPUs, SM counts, memory, and the peer-access (NVLink/NVSwitch) matrix.
#include <stdlib.h>

#include <string>

#include "../include/b200dp.h"
#include "common.cuh"

extern "C" int b200_topology_json(char* buf, int buflen) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) n = 0;
  std::string s = "{\"gpus\":[";
  for (int d = 0; d < n; ++d) {
    cudaDeviceProp p;
    if (cudaGetDeviceProperties(&p, d) != cudaSuccess) continue;
    char pci[32];
    snprintf(pci, sizeof pci, "%04x:%02x:%02x.0", p.pciDomainID, p.pciBusID, p.pciDeviceID);
    char tmp[512];
    snprintf(tmp, sizeof tmp,
             "%s{\"index\":%d,\"name\":\"%s\",\"cc\":\"%d.%d\",\"sms\":%d,\"mem_bytes\":%zu,\"l2_bytes\":%d,"
             "\"smem_per_block_optin\":%zu,\"pci\":\"%s\"}",
             d ? "," : "", d, p.name, p.major, p.minor, p.multiProcessorCount, p.totalGlobalMem, p.l2CacheSize,
             p.sharedMemPerBlockOptin, pci);
    s += tmp;
  }
  s += "],\"peer_access\":[";
  for (int a = 0; a < n; ++a) {
    s += a ? ",[" : "[";
    for (int b = 0; b < n; ++b) {
      int can = a == b;
      if (a != b) cudaDeviceCanAccessPeer(&can, a, b);
      s += (b ? "," : "");
      s += can ? "1" : "0";
    }
    s += "]";
  }
  s += "],\"p2p_native_atomics\":[";
  for (int a = 0; a < n; ++a) {
    s += a ? ",[" : "[";
    for (int b = 0; b < n; ++b) {
      int v = a == b;
      if (a != b) cudaDeviceGetP2PAttribute(&v, cudaDevP2PAttrNativeAtomicSupported, a, b);
      s += (b ? "," : "");
      s += v ? "1" : "0";
    }
    s += "]";
  }
  s += "]}";
  if ((int)s.size() + 1 > buflen) {
    b200_set_error("topology buffer too small: need %zu", s.size() + 1);
    return -(int)(s.size() + 1);
  }
  memcpy(buf, s.c_str(), s.size() + 1);
  return (int)s.size();
}

// PDL switch (default on; B200DP_PDL=0 or b200_set_pdl(0) turns every launch back into an ordinary one).
static int g_pdl = -1;
int b200_pdl_enabled() {
  if (g_pdl < 0) {
    const char* e = getenv("B200DP_PDL");
    g_pdl = (e && e[0] == '0') ? 0 : 1;
  }
  return g_pdl;
}
extern "C" int b200_set_pdl(int on) { g_pdl = on ? 1 : 0; return 0; }

// "The weights a kernel reads were not written by the kernel launched just before it" — true for every launch inside a
// training step (weights change only in the previous step's update).  Kernels launched while it is set may load their
// weights ahead of the programmatic dependency wait.  Off by default: a caller that writes weights and then launches
// a convolution right behind that write must not set it.
static int g_stable_w = 0;
int b200_stable_weights() { return g_stable_w; }
extern "C" int b200_set_stable_weights(int on) { g_stable_w = on ? 1 : 0; return 0; }

// Pin [base, base+bytes) in L2 for every kernel launched into `stream` afterwards (also the kernel nodes a capture of that
// stream creates): hits inside the window are kept with the persisting property, misses stream through.  Experiment
// switch B200DP_L2_PERSIST of the engine: the Adam state (fp32 master/m/v, 39 MB for the MNIST CNN) is swept once per
// step and is otherwise evicted by the step's other 40 MB of traffic.  bytes == 0 clears the window.
extern "C" int b200_stream_l2_persist(void* base, size_t bytes, float hit_ratio, void* stream) {
  int dev = 0, max_persist = 0, max_window = 0;
  B200_CHECK(cudaGetDevice(&dev));
  B200_CHECK(cudaDeviceGetAttribute(&max_persist, cudaDevAttrMaxPersistingL2CacheSize, dev));
  B200_CHECK(cudaDeviceGetAttribute(&max_window, cudaDevAttrMaxAccessPolicyWindowSize, dev));
  cudaStreamAttrValue v;
  memset(&v, 0, sizeof v);
  if (bytes == 0) {
    v.accessPolicyWindow.num_bytes = 0;
    B200_CHECK(cudaStreamSetAttribute((cudaStream_t)stream, cudaStreamAttributeAccessPolicyWindow, &v));
    return B200_OK;
  }
  B200_REQUIRE(max_persist > 0 && max_window > 0, "this device has no persisting L2 (max %d, window %d)", max_persist, max_window);
  B200_REQUIRE(hit_ratio > 0.f && hit_ratio <= 1.f, "hit_ratio %f outside (0,1]", hit_ratio);
  const size_t carve = bytes < (size_t)max_persist ? bytes : (size_t)max_persist;
  B200_CHECK(cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, carve));
  v.accessPolicyWindow.base_ptr = base;
  v.accessPolicyWindow.num_bytes = bytes < (size_t)max_window ? bytes : (size_t)max_window;
  // a window larger than the carve-out would thrash the persisting lines against each other: scale the hit ratio down
  float hr = hit_ratio;
  if (v.accessPolicyWindow.num_bytes > carve) hr *= (float)carve / (float)v.accessPolicyWindow.num_bytes;
  v.accessPolicyWindow.hitRatio = hr;
  v.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
  v.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
  B200_CHECK(cudaStreamSetAttribute((cudaStream_t)stream, cudaStreamAttributeAccessPolicyWindow, &v));
  return B200_OK;
}
