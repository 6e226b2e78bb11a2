// kernels.h — host-callable launchers for the probe kernels.  Every launcher
// only enqueues on `stream`; none allocates, synchronises or touches the host.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/gpuprobe.h"

namespace gp {

// internal pattern kind: checksum only, no expected value (mismatches stay 0)
constexpr int kPatNone = 3;

// accumulator slot of a verify launch (scratch memory; see pattern.cu).  CTAs add their
// partials into slot (blockIdx & (kVerifySlots-1)); 64-byte stride keeps slots in separate
// L2 sectors so the reductions spread over the L2 slices.
struct VerifySlot {
  uint64_t sum, xor_, wsum, mismatches;
  uint64_t first_bad_index, first_bad_xor;  // one 16-byte pair at offset 32: updated together by a 128-bit CAS
  uint64_t pad[2];
};
static_assert(sizeof(VerifySlot) == 64, "one slot per 64 bytes");
constexpr int kVerifySlots = 512;
// scratch layout: [kVerifySlots] VerifySlot.  Rest state: sums 0, first_bad_index all-ones
// (set once by launch_verify_scratch_init, restored by every verify launch's finalize CTA).
constexpr size_t kVerifyScratchBytes = kVerifySlots * sizeof(VerifySlot) + 256;

cudaError_t launch_pattern_fill(uint64_t* words, uint64_t nwords, int kind, int pass, int sm_count,
                                cudaStream_t stream);
cudaError_t launch_verify_scratch_init(void* scratch, cudaStream_t stream);
// two launches: the chunk grid, then a one-CTA finalize.  result is fully overwritten by plain
// stores of the finalize CTA; scratch must be
// kVerifyScratchBytes in its rest state (it is left in that state on completion).
cudaError_t launch_pattern_verify(const uint64_t* words, uint64_t nwords, int kind, int pass,
                                  void* scratch, gp_pattern_result* d_result, int sm_count,
                                  cudaStream_t stream);
// fold the accumulator slots of `scratch` into *d_result (sum, xor, wsum as accumulated; mismatches as
// accumulated) and return the slots to their rest state: the one-CTA finalize alone, for kernels that fill
// the slots themselves (the P2P read kernels)
cudaError_t launch_verify_fold(void* scratch, gp_pattern_result* d_result, cudaStream_t stream);
cudaError_t launch_bitflip(uint64_t* words, uint64_t word_index, int bit, cudaStream_t stream);

// HBM stream kernels.  bytes is a multiple of 1 MiB; variant is GP_STREAM_VEC or GP_STREAM_TMA.
cudaError_t launch_copy(void* dst, const void* src, uint64_t bytes, int variant, int sm_count,
                        cudaStream_t stream);
cudaError_t launch_triad(float* a, const float* b, const float* c, float s, uint64_t bytes,
                         int variant, int sm_count, cudaStream_t stream);
// operand generators (GP_K_INIT)
cudaError_t launch_gen_f32(float* dst, uint64_t n, uint64_t seed, int sm_count, cudaStream_t stream);
cudaError_t launch_gen_bf16(uint16_t* dst, uint64_t n, uint64_t seed, int sm_count,
                            cudaStream_t stream);
// counts fp32 elements of a that differ bitwise from fmaf(s, gen(c), gen(b)); *d_count += count
cudaError_t launch_triad_check(const float* a, uint64_t n, uint64_t seed, float s,
                               unsigned long long* d_count, int sm_count, cudaStream_t stream);
cudaError_t stream_kernels_configure();  // opt-in dynamic smem; call once per device
cudaError_t gemm_kernels_configure();    // same, for the tcgen05 kernels (sm_100 devices only)

}  // namespace gp
