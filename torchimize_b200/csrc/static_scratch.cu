// static_scratch.cu — nlls_model_normal_eq_static: the fused normal-equations kernel with its scratch (the two [B,N]
// model-value ping-pong buffers) in library-owned static device memory, for callers that cannot allocate: code under CUDA graph
// capture, serving loops, harnesses that forbid cudaMalloc.  No counterpart in the reference.
// Allocates nothing: no cudaMalloc / cudaFree, no allocating library routine.
//
// The array below is a HOST-SIDE STAND-IN in host code, not a device address: the address the kernels get comes from
// cudaGetSymbolAddress, looked up once per device with that device current.
#include <mutex>
#include "common.cuh"

namespace nlls {

constexpr long long kStaticElems = NLLS_STATIC_SCRATCH_ELEMS;            // B*N limit of one call (either dtype)
constexpr size_t kStaticHalf = (size_t)kStaticElems * sizeof(double);    // bytes of one ping-pong buffer (8 MiB)
constexpr int kMaxDevices = 64;

// 16 MiB per GPU that loads the library, held until the process exits.  Shared by every call; keeps the last call's contents.
static __device__ __align__(256) unsigned char g_m_scratch[2 * kStaticHalf];

static std::mutex g_scratch_mutex;
static void* g_scratch_addr[kMaxDevices] = {};

// device address of g_m_scratch on the CURRENT device (cached per device ordinal); nullptr + set_error on failure
static unsigned char* scratch_on_current_device() {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) { set_error("nlls_model_normal_eq_static: cudaGetDevice: %s", cudaGetErrorString(e)); return nullptr; }
  if (dev < 0 || dev >= kMaxDevices) { set_error("nlls_model_normal_eq_static: device ordinal %d >= %d", dev, kMaxDevices); return nullptr; }
  std::lock_guard<std::mutex> lock(g_scratch_mutex);
  if (!g_scratch_addr[dev]) {
    void* ptr = nullptr;
    e = cudaGetSymbolAddress((void**)&ptr, g_m_scratch);
    if (e != cudaSuccess || !ptr) {
      set_error("nlls_model_normal_eq_static: cudaGetSymbolAddress on device %d: %s", dev, cudaGetErrorString(e));
      cudaGetLastError();
      return nullptr;
    }
    g_scratch_addr[dev] = ptr;
  }
  return (unsigned char*)g_scratch_addr[dev];
}

}  // namespace nlls

using namespace nlls;

extern "C" int nlls_model_normal_eq_static(int model, int dtype, const void* p, int P, const void* x, const void* y,
                                           long long y_bstride, long long y_cstride, const void* wvec, int wlen, long long B,
                                           int C, int N, void* g, void* H, double* stats, const nlls_ctrl* ctrl,
                                           const nlls_track* track, void* stream) {
  if (B < 0 || N < 1) { set_error("nlls_model_normal_eq_static: B=%lld N=%d", B, N); return NLLS_E_SHAPE; }
  if (B > kStaticElems / N) {                       // B*N > limit, without overflow
    set_error("nlls_model_normal_eq_static: B*N = %lld x %d exceeds the static scratch of NLLS_STATIC_SCRATCH_ELEMS = %lld "
              "model values per buffer; use nlls_model_normal_eq with caller-owned m_pp buffers", B, N, kStaticElems);
    return NLLS_E_SHAPE;
  }
  unsigned char* base = scratch_on_current_device();
  if (!base) return NLLS_E_NULL;
  return nlls_model_normal_eq(model, dtype, p, P, x, y, y_bstride, y_cstride, wvec, wlen, B, C, N, base, base + kStaticHalf, g, H,
                              stats, ctrl, track, stream);
}
