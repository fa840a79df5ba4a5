// Host-side plumbing shared by every entry point of libtm_b200: error text, cached device properties.
#include <mutex>
#include <string>

#include "common.cuh"

namespace tmb {

static thread_local std::string g_last_error;

void set_cuda_error(cudaError_t e, const char* where) {
  g_last_error = std::string(cudaGetErrorName(e)) + ": " + cudaGetErrorString(e) + " at " + where;
  cudaGetLastError();  // clear the sticky-less error so the next call starts clean
}

void set_error_text(const char* text) { g_last_error = text; }

struct DevProps {
  int sms = 0;
  int smem_optin = 0;
};
static DevProps g_props[64];
static std::mutex g_mu;

static const DevProps& props() {
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64) dev = 0;
  std::lock_guard<std::mutex> lk(g_mu);
  DevProps& p = g_props[dev];
  if (p.sms == 0) {
    cudaDeviceGetAttribute(&p.sms, cudaDevAttrMultiProcessorCount, dev);
    cudaDeviceGetAttribute(&p.smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
    if (p.sms <= 0) p.sms = 148;
    if (p.smem_optin <= 0) p.smem_optin = 227 * 1024;
  }
  return p;
}

int sm_count() { return props().sms; }
int max_optin_smem() { return props().smem_optin; }

}  // namespace tmb

extern "C" int tmb_abi_version(void) { return TMB_ABI_VERSION; }
extern "C" const char* tmb_last_cuda_error(void) { return tmb::g_last_error.c_str(); }
extern "C" int tmb_sm_count(void) { return tmb::sm_count(); }
