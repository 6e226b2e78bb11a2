// noalloc.cu — gp_pattern_verify_noalloc: the pattern verify for callers that must not
// allocate.  It has no counterpart in the reference.
//
// Who cannot allocate: a health check that re-grades a GPU while tenant pods own it (free
// device memory can be a few MiB; a cudaMalloc there fails, or takes memory a pod was
// promised), and callers that sit where the CUDA allocator must not be entered at all —
// cudaMalloc and cudaFree synchronise the whole device, so code on a latency-critical stream
// or between two kernels that must stay back to back keeps them out.  Such a caller already
// holds the memory to be graded and a place for the 48-byte answer; the only thing the verify
// kernel needs beyond that is its scratch (the accumulator slots described in pattern.cu).
// Here that scratch is a fixed array in the library's own device image, so nothing is
// allocated, on the device or through any allocating library routine, by this entry point.
//
// Cost: every GPU that loads this library carries the kVerifyScratchBytes (33,024 byte) array
// until the process exits.
#include <mutex>

#include "common.cuh"
#include "engine.h"
#include "kernels.h"

namespace {

constexpr int kMaxOrdinals = 64;

// The scratch itself.  In HOST code this name is only a stand-in for the symbol, not a device
// address: handing it to a kernel would fault, or on a GPU that can reach host memory quietly
// use the wrong bytes.  Host code must go through cudaGetSymbolAddress (below).
static __device__ __align__(256) unsigned char g_noalloc_verify_scratch[gp::kVerifyScratchBytes];

std::mutex g_setup_mu;
void* g_scratch_addr[kMaxOrdinals] = {};  // per device, filled on that device's first call

bool kind_pass_ok(int kind, int pass) {
  if (kind == GP_PAT_WALKING_ONES || kind == GP_PAT_WALKING_ZEROS) return pass >= 0 && pass < 64;
  if (kind == GP_PAT_ADDRESS) return pass == 0 || pass == 1;
  return false;
}

}  // namespace

// A cudaDeviceReset destroys the device's copy of the array; the address cached for it is then stale.
// gp_device_reset calls this; a caller that resets the device itself cannot, see include/gpuprobe.h.
void gp::noalloc_forget_device(int ordinal) {
  std::lock_guard<std::mutex> g(g_setup_mu);
  if (ordinal >= 0 && ordinal < kMaxOrdinals) g_scratch_addr[ordinal] = nullptr;
}

extern "C" int gp_pattern_verify_noalloc(int device_ordinal, const uint64_t* d_words, uint64_t nwords,
                                         int kind, int pass, gp_pattern_result* d_result, void* stream) {
  using namespace gp;
  if (device_ordinal < 0 || device_ordinal >= kMaxOrdinals) {
    set_error("gp_pattern_verify_noalloc: device ordinal %d outside [0, %d)", device_ordinal, kMaxOrdinals);
    return GP_ERR_NO_DEVICE;
  }
  if (!d_result || (!d_words && nwords) || !kind_pass_ok(kind, pass) || (nwords % 64)) {
    set_error("gp_pattern_verify_noalloc: null pointer, bad kind/pass, or nwords not a multiple of 64");
    return GP_ERR_INVALID;
  }
  // The kernels read 256 bits at a time and the result is written as six 64-bit words.  A misaligned
  // pointer would not fail here but inside the kernel, as a sticky fault that takes the caller's whole
  // context with it, so it is refused up front.
  if ((reinterpret_cast<uintptr_t>(d_words) & 31u) || (reinterpret_cast<uintptr_t>(d_result) & 7u)) {
    set_error("gp_pattern_verify_noalloc: d_words must be 32-byte aligned and d_result 8-byte aligned");
    return GP_ERR_INVALID;
  }
  if (nwords > GP_NOALLOC_MAX_WORDS) {
    set_error("gp_pattern_verify_noalloc: %llu words exceeds the limit of %llu words (%llu GiB) per call; "
              "split the range and combine the results",
              (unsigned long long)nwords, (unsigned long long)GP_NOALLOC_MAX_WORDS,
              (unsigned long long)(GP_NOALLOC_MAX_WORDS >> 27));
    return GP_ERR_TOO_LARGE;
  }
  cudaError_t e = cudaSetDevice(device_ordinal);
  if (e != cudaSuccess) return cuda_fail(e, "cudaSetDevice", nullptr);

  void* scratch = nullptr;
  {
    // Set-up, once per device, with that device current: ask the runtime where this device's
    // copy of the array lives.  (First call on a device; keep it out of stream capture.)
    std::lock_guard<std::mutex> g(g_setup_mu);
    if (!g_scratch_addr[device_ordinal]) {
      void* p = nullptr;
      e = cudaGetSymbolAddress(&p, g_noalloc_verify_scratch);
      if (e != cudaSuccess) return cuda_fail(e, "cudaGetSymbolAddress(g_noalloc_verify_scratch)", nullptr);
      g_scratch_addr[device_ordinal] = p;
    }
    scratch = g_scratch_addr[device_ordinal];
  }
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  // The array is shared by every call and keeps whatever the last one left (a launch that was
  // aborted half-way leaves partial sums behind), so each call first puts the slots back into
  // their rest state: sums zero, first-bad-index all-ones.
  e = launch_verify_scratch_init(scratch, s);
  if (e != cudaSuccess) return cuda_fail(e, "verify scratch init launch", nullptr);
  e = launch_pattern_verify(d_words, nwords, kind, pass, scratch, d_result, 0, s);
  if (e != cudaSuccess) return cuda_fail(e, "verify launch", nullptr);
  return GP_OK;
}
