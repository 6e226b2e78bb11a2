// p2p.cu — the only cross-device probes (K4a / K4b of SURVEY.md §8).
//
//   gp_probe_p2p       : NVLink bandwidth matrix with IN-KERNEL peer loads / stores — the
//                        kernel runs on GPU `src` and dereferences a pointer into GPU `dst`'s
//                        memory (LDG / STG on a peer-mapped address), no copy engine.
//   gp_probe_allreduce : ncclAllReduce(sum, fp32, in place) bus bandwidth, single process,
//                        one communicator per device (ncclCommInitAll), group launch.
//
// No reference counterpart: the reference's only topology input is go-gpuallocator's NVML
// link query (plugin/plugin.go:259-276); it moves no bytes.  Algorithmic bytes:
// P2P S bytes per direction per transfer; all-reduce algbw S/t, busbw S/t * 2(n-1)/n.
//
// Pairs are scheduled as the n-1 rounds of a 1-factorisation (circle method): in a round every
// GPU is in exactly one pair, so all GPUs are busy and no GPU serves two peers at once.  A
// unidirectional round runs in two phases (a->b for every pair, then b->a); the bidirectional
// mode runs both directions of every pair at once and reports each direction.
#include <dlfcn.h>
#include <nccl.h>

#include <algorithm>
#include <chrono>
#include <cstdlib>
#include <cstring>
#include <map>
#include <vector>

#include "common.cuh"
#include "engine.h"

namespace gp {

// accumulate what a peer read loaded: sum and xor of the u64 words, folded per warp into the verify
// scratch slots (pattern.cu) with fire-and-forget reductions; launch_verify_fold turns the slots into a
// gp_pattern_result the host compares with the closed form of the pattern the peer holds
__device__ __forceinline__ void p2p_fold(VerifySlot* __restrict__ slots, uint64_t sum, uint64_t xr) {
  sum = warp_sum_u64(sum);
  xr = warp_xor_u64(xr);
  if ((threadIdx.x & 31) == 0) {
    VerifySlot* s = slots + ((blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)) & (kVerifySlots - 1));
    atomicAdd(reinterpret_cast<unsigned long long*>(&s->sum), (unsigned long long)sum);
    atomicXor(reinterpret_cast<unsigned long long*>(&s->xor_), (unsigned long long)xr);
  }
}
__device__ __forceinline__ void p2p_acc(const u64x4& x, uint64_t& sum, uint64_t& xr) {
  sum += (x.v[0] + x.v[1]) + (x.v[2] + x.v[3]);
  xr ^= (x.v[0] ^ x.v[1]) ^ (x.v[2] ^ x.v[3]);
}

#define CU_OR_RETURN(call)                                               \
  do {                                                                   \
    cudaError_t e__ = (call);                                            \
    if (e__ != cudaSuccess) return ::gp::cuda_fail(e__, #call, nullptr); \
  } while (0)

constexpr int kP2PThreads = 512;
constexpr int kP2PUnroll = 4;  // 4 x 32 B in flight per thread: peer latency is ~2-4k cycles

// kernel on `src`: read every vector of the peer buffer and fold it (sum, xor) for the host to check
__global__ void __launch_bounds__(kP2PThreads)
p2p_read_kernel(const u64x4* __restrict__ peer, uint64_t nvec, VerifySlot* __restrict__ slots) {
  const uint64_t stride = (uint64_t)gridDim.x * kP2PThreads;
  uint64_t g = (uint64_t)blockIdx.x * kP2PThreads + threadIdx.x;
  uint64_t sum = 0, xr = 0;
  for (; g + (kP2PUnroll - 1) * stride < nvec; g += kP2PUnroll * stride) {
    u64x4 v[kP2PUnroll];
#pragma unroll
    for (int u = 0; u < kP2PUnroll; ++u) v[u] = ld256_stream(peer + g + u * stride);
#pragma unroll
    for (int u = 0; u < kP2PUnroll; ++u) p2p_acc(v[u], sum, xr);
  }
  for (; g < nvec; g += stride) {
    const u64x4 v = ld256_stream(peer + g);
    p2p_acc(v, sum, xr);
  }
  p2p_fold(slots, sum, xr);
}

// kernel on `src`: store the address pattern into the peer buffer
__global__ void __launch_bounds__(kP2PThreads)
p2p_write_kernel(u64x4* __restrict__ peer, uint64_t nvec, uint64_t tag) {
  const uint64_t stride = (uint64_t)gridDim.x * kP2PThreads;
  for (uint64_t g = (uint64_t)blockIdx.x * kP2PThreads + threadIdx.x; g < nvec; g += stride) {
    u64x4 v;
    v.v[0] = tag; v.v[1] = g; v.v[2] = ~g; v.v[3] = tag ^ g;
    st256_stream(peer + g, v);
  }
}

// non-persistent shapes: one CTA per contiguous chunk (kP2PThreads x U vectors), no loop
template <int U>
__global__ void __launch_bounds__(kP2PThreads)
p2p_read_np_kernel(const u64x4* __restrict__ peer, uint64_t nvec, VerifySlot* __restrict__ slots) {
  const uint64_t g0 = (uint64_t)blockIdx.x * (kP2PThreads * U) + threadIdx.x;
  uint64_t sum = 0, xr = 0;
  u64x4 v[U];
#pragma unroll
  for (int u = 0; u < U; ++u) {
    const uint64_t g = g0 + (uint64_t)u * kP2PThreads;
    if (g < nvec) v[u] = ld256_stream(peer + g); else v[u] = u64x4{{0, 0, 0, 0}};
  }
#pragma unroll
  for (int u = 0; u < U; ++u) p2p_acc(v[u], sum, xr);
  p2p_fold(slots, sum, xr);
}
template <int U>
__global__ void __launch_bounds__(kP2PThreads)
p2p_write_np_kernel(u64x4* __restrict__ peer, uint64_t nvec, uint64_t tag) {
  const uint64_t g0 = (uint64_t)blockIdx.x * (kP2PThreads * U) + threadIdx.x;
#pragma unroll
  for (int u = 0; u < U; ++u) {
    const uint64_t g = g0 + (uint64_t)u * kP2PThreads;
    if (g < nvec) {
      u64x4 v;
      v.v[0] = tag; v.v[1] = g; v.v[2] = ~g; v.v[3] = tag ^ g;
      st256_stream(peer + g, v);
    }
  }
}

// all-pairs-at-once: one kernel on `src` reads from EVERY peer concurrently (CTA b -> peer b % npeers),
// the switch-saturation round of SURVEY.md §8(e)
struct PeerTable {
  const u64x4* p[8];
  int n;
};
__global__ void __launch_bounds__(kP2PThreads)
p2p_read_all_kernel(PeerTable t, uint64_t nvec_per_peer, VerifySlot* __restrict__ slots) {
  const int peer = blockIdx.x % t.n;
  const uint64_t cta = blockIdx.x / t.n, ctas = gridDim.x / t.n;
  const u64x4* __restrict__ base = t.p[peer];
  const uint64_t stride = ctas * kP2PThreads;
  uint64_t sum = 0, xr = 0;
  uint64_t g = cta * kP2PThreads + threadIdx.x;
  for (; g + (kP2PUnroll - 1) * stride < nvec_per_peer; g += kP2PUnroll * stride) {
    u64x4 v[kP2PUnroll];
#pragma unroll
    for (int u = 0; u < kP2PUnroll; ++u) v[u] = ld256_stream(base + g + u * stride);
#pragma unroll
    for (int u = 0; u < kP2PUnroll; ++u) p2p_acc(v[u], sum, xr);
  }
  for (; g < nvec_per_peer; g += stride) {
    const u64x4 v = ld256_stream(base + g);
    p2p_acc(v, sum, xr);
  }
  p2p_fold(slots, sum, xr);
}

// peer write through TMA: threads build a 16 KiB tile of the pattern in shared memory, one
// thread bulk-stores it to the peer (cp.async.bulk shared -> peer global)
constexpr uint32_t kP2PTile = 16 * 1024;
__global__ void __launch_bounds__(128)
p2p_write_tma_kernel(uint8_t* __restrict__ peer, uint64_t ntiles, uint64_t tag) {
  __shared__ __align__(128) uint8_t tile[2][kP2PTile];
  int buf = 0;
  for (uint64_t t = blockIdx.x; t < ntiles; t += gridDim.x, buf ^= 1) {
    u64x4* sv = reinterpret_cast<u64x4*>(tile[buf]);
    const uint64_t g0 = t * (kP2PTile / 32);
    for (uint32_t i = threadIdx.x; i < kP2PTile / 32; i += 128) {
      const uint64_t g = g0 + i;
      u64x4 v;
      v.v[0] = tag; v.v[1] = g; v.v[2] = ~g; v.v[3] = tag ^ g;
      sv[i] = v;
    }
    fence_proxy_async_smem();
    __syncthreads();
    if (threadIdx.x == 0) {
      bulk_s2g(peer + t * kP2PTile, tile[buf], kP2PTile);
      bulk_commit();
      bulk_wait_read<1>();  // the other buffer's store has finished reading: safe to refill it next trip
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) bulk_wait_all<0>();
}

__global__ void p2p_check_kernel(const u64x4* __restrict__ buf, uint64_t nvec, uint64_t tag,
                                 unsigned long long* __restrict__ bad) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  unsigned long long b = 0;
  for (uint64_t g = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; g < nvec; g += stride) {
    const u64x4 v = buf[g];
    b += (v.v[0] != tag) | (v.v[1] != g) | (v.v[2] != ~g) | (v.v[3] != (tag ^ g));
  }
  if (b) atomicAdd(bad, b);
}

static int p2p_grid(uint64_t nvec, int sm_count) {
  const uint64_t want = (nvec + kP2PThreads - 1) / kP2PThreads;
  const uint64_t cap = (uint64_t)sm_count * 4;  // 2048 threads per SM
  return (int)std::max<uint64_t>(1, std::min(want, cap));
}

// 1-factorisation of K_n (n even) by the circle method; odd n gets a bye.
static std::vector<std::vector<std::pair<int, int>>> rounds_for(int n) {
  std::vector<std::vector<std::pair<int, int>>> out;
  const int m = (n % 2) ? n + 1 : n;
  for (int r = 0; r < m - 1; ++r) {
    std::vector<std::pair<int, int>> round;
    for (int i = 0; i < m / 2; ++i) {
      int a = (r + i) % (m - 1);
      int b = (i == 0) ? m - 1 : (r + m - 1 - i) % (m - 1);
      if (a >= n || b >= n) continue;
      if (a > b) std::swap(a, b);
      round.push_back({a, b});
    }
    out.push_back(round);
  }
  return out;
}

struct Xfer {
  int src, dst;
};

enum XferMode { kXferRead = 0, kXferWrite = 1, kXferCopyEngine = 2 };

// SUM and XOR of the address pattern (pass 0) over n words: SURVEY.md §8(c) closed forms
static void address_closed_form(uint64_t n, uint64_t* sum, uint64_t* xr) {
  *sum = 4 * n * (n - 1);  // mod 2^64 by construction
  const uint64_t m = n - 1;
  const uint64_t X[4] = {m, 1, m + 1, 0};
  *xr = n ? 8 * X[m & 3] : 0;
}

struct P2PState {
  bool data_bad = false;  // a read returned the wrong bytes, or a write / copy did not land intact
};

// run a set of transfers concurrently (one per src device), return best-of-iters GB/s of each.
// Reads fold what they loaded into the src device's verify slots; the fold lands in d_results[it] and
// is compared with the closed form after the sync.
static int run_xfers(Ctx* ctx, const std::vector<Xfer>& xs, size_t bytes, XferMode mode, int iters,
                     std::vector<double>* gbps_out, P2PState* st) {
  const uint64_t nvec = bytes / 32;
  gbps_out->assign(xs.size(), 0.0);
  std::vector<float> best(xs.size(), 1e30f);
  // launch shape: measured on 2 and 8 x B200 (tools/p2p_shapes.py, profiles/multi_gpu_8_r01.json): read
  // 776-785 GB/s and write 713 GB/s per direction at 1 GiB for EVERY shape tried — the link protocol, not
  // the kernel, is the limit — so the default is the non-persistent shape the local kernels use
  static const int shape = GP_DEV_SWITCH("GPUPROBE_P2P_SHAPE", 1);
  for (int it = 0; it < iters + 1; ++it) {  // first pass is the warm-up
    for (size_t i = 0; i < xs.size(); ++i) {
      Device* s = ctx->devs[(size_t)xs[i].src].get();
      Device* d = ctx->devs[(size_t)xs[i].dst].get();
      CU_OR_RETURN(cudaSetDevice(s->ordinal));
      CU_OR_RETURN(cudaEventRecord(s->events[0], s->stream));
      const uint64_t tag = 0xC0DE0000ull + (uint64_t)xs[i].src;
      u64x4* pa = reinterpret_cast<u64x4*>(d->arena);
      VerifySlot* slots = reinterpret_cast<VerifySlot*>(s->scratch);
      auto npgrid = [&](int U) { return (unsigned)((nvec + (uint64_t)kP2PThreads * U - 1) / ((uint64_t)kP2PThreads * U)); };
      if (mode == kXferCopyEngine) {
        // the yardstick: the driver's own peer copy, src's source half -> dst's landing half
        CU_OR_RETURN(cudaMemcpyPeerAsync(d->arena + bytes, d->ordinal, s->arena, s->ordinal, bytes, s->stream));
      } else if (mode == kXferWrite) {
        if (shape == 1) p2p_write_np_kernel<1><<<npgrid(1), kP2PThreads, 0, s->stream>>>(pa, nvec, tag);
        else if (shape == 2) p2p_write_np_kernel<4><<<npgrid(4), kP2PThreads, 0, s->stream>>>(pa, nvec, tag);
        else if (shape == 3) p2p_write_kernel<<<s->sm_count * 2, kP2PThreads, 0, s->stream>>>(pa, nvec, tag);
        else if (shape == 4) p2p_write_kernel<<<s->sm_count, kP2PThreads, 0, s->stream>>>(pa, nvec, tag);
        else if (shape == 5 && bytes % kP2PTile == 0)
          p2p_write_tma_kernel<<<s->sm_count * 4, 128, 0, s->stream>>>(reinterpret_cast<uint8_t*>(pa), bytes / kP2PTile, tag);
        else p2p_write_kernel<<<p2p_grid(nvec, s->sm_count), kP2PThreads, 0, s->stream>>>(pa, nvec, tag);
      } else {
        if (shape == 1) p2p_read_np_kernel<2><<<npgrid(2), kP2PThreads, 0, s->stream>>>(pa, nvec, slots);
        else if (shape == 2) p2p_read_np_kernel<8><<<npgrid(8), kP2PThreads, 0, s->stream>>>(pa, nvec, slots);
        else if (shape == 3) p2p_read_kernel<<<s->sm_count * 2, kP2PThreads, 0, s->stream>>>(pa, nvec, slots);
        else if (shape == 4) p2p_read_kernel<<<s->sm_count, kP2PThreads, 0, s->stream>>>(pa, nvec, slots);
        else p2p_read_kernel<<<p2p_grid(nvec, s->sm_count), kP2PThreads, 0, s->stream>>>(pa, nvec, slots);
      }
      CU_OR_RETURN(cudaGetLastError());
      CU_OR_RETURN(cudaEventRecord(s->events[1], s->stream));
      if (mode == kXferRead) CU_OR_RETURN(launch_verify_fold(s->scratch, s->d_results + it, s->stream));
    }
    for (size_t i = 0; i < xs.size(); ++i) {
      Device* s = ctx->devs[(size_t)xs[i].src].get();
      CU_OR_RETURN(cudaSetDevice(s->ordinal));
      CU_OR_RETURN(cudaStreamSynchronize(s->stream));
      float ms = 0;
      CU_OR_RETURN(cudaEventElapsedTime(&ms, s->events[0], s->events[1]));
      if (it > 0) {
        best[i] = std::min(best[i], ms);
        s->stats[GP_K_P2P].launches += 1;
        s->stats[GP_K_P2P].total_ms += ms;
        s->stats[GP_K_P2P].algo_bytes += (double)bytes;
      }
    }
  }
  if (mode == kXferRead) {
    uint64_t esum, exor;
    address_closed_form(bytes / 8, &esum, &exor);
    for (size_t i = 0; i < xs.size(); ++i) {
      Device* s = ctx->devs[(size_t)xs[i].src].get();
      CU_OR_RETURN(cudaSetDevice(s->ordinal));
      CU_OR_RETURN(cudaMemcpy(s->h_results, s->d_results, (size_t)(iters + 1) * sizeof(gp_pattern_result), cudaMemcpyDeviceToHost));
      for (int it = 0; it < iters + 1; ++it)
        if (s->h_results[it].sum != esum || s->h_results[it].xor_ != exor) {
          set_error("device %d read the wrong bytes from device %d: sum %016llx xor %016llx, expected %016llx %016llx",
                    xs[i].src, xs[i].dst, (unsigned long long)s->h_results[it].sum, (unsigned long long)s->h_results[it].xor_,
                    (unsigned long long)esum, (unsigned long long)exor);
          st->data_bad = true;
          return GP_ERR_CUDA;
        }
    }
  }
  for (size_t i = 0; i < xs.size(); ++i) (*gbps_out)[i] = (double)bytes / (best[i] * 1e-3) / 1e9;
  return GP_OK;
}

// multi_mu and every device lock held by the caller
static int p2p_probe_locked(Ctx* ctx, int ngpus, size_t bytes, int dir, double* gbps, P2PState* st, bool keep) {
  if (bytes == 0 || bytes % 32) { set_error("bytes must be a non-zero multiple of 32"); return GP_ERR_INVALID; }
  const bool copy_engine = dir == GP_P2P_COPY_ENGINE || dir == GP_P2P_COPY_ENGINE_BIDIR;
  for (int i = 0; i < ngpus * ngpus; ++i) gbps[i] = 0.0;
  unsigned long long all_mask = 0;
  for (int i = 0; i < ngpus; ++i) all_mask |= 1ull << ctx->devs[(size_t)i]->ordinal;
  for (int i = 0; i < ngpus; ++i) {
    Device* d = ctx->devs[(size_t)i].get();
    int rc = device_prepare(ctx, d);
    if (rc) return rc;
    rc = ensure_arena(ctx, d, std::max<size_t>((copy_engine ? 2 : 1) * bytes, 1 << 20));
    if (rc) return rc;
    d->triad_operands_bytes = 0;
    // what every reader must find: the address pattern, word index from the arena base
    CU_OR_RETURN(launch_pattern_fill(reinterpret_cast<uint64_t*>(d->arena), bytes / 8, GP_PAT_ADDRESS, 0, d->sm_count, d->stream));
    if ((ctx->cfg.test_fault & GP_TEST_FAULT_P2P_CORRUPT) && i == ngpus - 1 && bytes >= 4096)  // test hook: one wrong bit
      CU_OR_RETURN(launch_bitflip(reinterpret_cast<uint64_t*>(d->arena), bytes / 16 + 3, 17, d->stream));
    if (copy_engine) CU_OR_RETURN(cudaMemsetAsync(d->arena + bytes, 0x5A, bytes, d->stream));
    CU_OR_RETURN(cudaMemsetAsync(d->d_count, 0, sizeof(unsigned long long), d->stream));
    CU_OR_RETURN(cudaStreamSynchronize(d->stream));
  }
  for (int i = 0; i < ngpus; ++i) {
    CU_OR_RETURN(cudaSetDevice(i));
    for (int j = 0; j < ngpus; ++j) {
      if (i == j) continue;
      int can = 0;
      CU_OR_RETURN(cudaDeviceCanAccessPeer(&can, i, j));
      if (!can) { set_error("device %d cannot access peer %d", i, j); return GP_ERR_UNSUPPORTED; }
      cudaError_t e = cudaDeviceEnablePeerAccess(j, 0);
      if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) return cuda_fail(e, "cudaDeviceEnablePeerAccess", nullptr);
      (void)cudaGetLastError();
    }
  }
  // a VMM arena is not covered by cudaDeviceEnablePeerAccess: grant the peers explicitly
  for (int i = 0; i < ngpus; ++i) {
    int rc = arena_grant_peers(ctx->devs[(size_t)i].get(), all_mask);
    if (rc) return rc;
  }
  auto done = [&](int rc) {
    if (!keep)
      for (int i = 0; i < ngpus; ++i) free_arena(ctx->devs[(size_t)i].get());
    return rc;
  };
  if (dir == GP_P2P_ALL_AT_ONCE) {
    // every GPU reads `bytes` from each of its n-1 peers at once; row i holds GPU i's aggregate
    // ingress divided evenly over its peers, the diagonal holds the aggregate itself
    if (ngpus < 2 || ngpus > 9) { set_error("all-at-once mode needs 2..9 GPUs"); return done(GP_ERR_INVALID); }
    const uint64_t nvec = bytes / 32;
    const int kIt = 4;
    std::vector<float> best((size_t)ngpus, 1e30f);
    for (int it = 0; it < kIt; ++it) {
      for (int i = 0; i < ngpus; ++i) {
        Device* s = ctx->devs[(size_t)i].get();
        PeerTable t{};
        for (int j = 0; j < ngpus; ++j)
          if (j != i) t.p[t.n++] = reinterpret_cast<const u64x4*>(ctx->devs[(size_t)j]->arena);
        CU_OR_RETURN(cudaSetDevice(i));
        CU_OR_RETURN(cudaEventRecord(s->events[0], s->stream));
        const int ctas_per_peer = std::max(1, s->sm_count * 4 / t.n);
        p2p_read_all_kernel<<<ctas_per_peer * t.n, kP2PThreads, 0, s->stream>>>(t, nvec, reinterpret_cast<VerifySlot*>(s->scratch));
        CU_OR_RETURN(cudaGetLastError());
        CU_OR_RETURN(cudaEventRecord(s->events[1], s->stream));
        CU_OR_RETURN(launch_verify_fold(s->scratch, s->d_results + it, s->stream));
      }
      for (int i = 0; i < ngpus; ++i) {
        Device* s = ctx->devs[(size_t)i].get();
        CU_OR_RETURN(cudaSetDevice(i));
        CU_OR_RETURN(cudaStreamSynchronize(s->stream));
        float ms = 0;
        CU_OR_RETURN(cudaEventElapsedTime(&ms, s->events[0], s->events[1]));
        if (it > 0) {
          best[(size_t)i] = std::min(best[(size_t)i], ms);
          s->stats[GP_K_P2P].launches += 1;
          s->stats[GP_K_P2P].total_ms += ms;
          s->stats[GP_K_P2P].algo_bytes += (double)bytes * (ngpus - 1);
        }
      }
    }
    uint64_t esum, exor;
    address_closed_form(bytes / 8, &esum, &exor);
    esum *= (uint64_t)(ngpus - 1);
    if (!((ngpus - 1) & 1)) exor = 0;
    for (int i = 0; i < ngpus; ++i) {
      Device* s = ctx->devs[(size_t)i].get();
      CU_OR_RETURN(cudaSetDevice(i));
      CU_OR_RETURN(cudaMemcpy(s->h_results, s->d_results, kIt * sizeof(gp_pattern_result), cudaMemcpyDeviceToHost));
      for (int it = 0; it < kIt; ++it)
        if (s->h_results[it].sum != esum || s->h_results[it].xor_ != exor) {
          set_error("device %d read the wrong bytes from its peers in the all-at-once round", i);
          st->data_bad = true;
          return done(GP_ERR_CUDA);
        }
    }
    for (int i = 0; i < ngpus; ++i) {
      const double agg = (double)bytes * (ngpus - 1) / (best[(size_t)i] * 1e-3) / 1e9;
      for (int j = 0; j < ngpus; ++j) gbps[i * ngpus + j] = i == j ? agg : agg / (ngpus - 1);
    }
    return done(GP_OK);
  }
  const bool write = dir == GP_P2P_WRITE || dir == GP_P2P_BIDIR_WRITE;
  const bool bidir = dir == GP_P2P_BIDIR_WRITE || dir == GP_P2P_BIDIR_READ || dir == GP_P2P_COPY_ENGINE_BIDIR;
  const XferMode mode = copy_engine ? kXferCopyEngine : (write ? kXferWrite : kXferRead);
  const int iters = 3;
  std::vector<double> r;
  // diagonal: the same transfer on local memory, all GPUs at once
  {
    std::vector<Xfer> xs;
    for (int i = 0; i < ngpus; ++i) xs.push_back({i, i});
    int rc = run_xfers(ctx, xs, bytes, mode, iters, &r, st);
    if (rc) return done(rc);
    for (int i = 0; i < ngpus; ++i) gbps[i * ngpus + i] = r[(size_t)i];
  }
  for (const auto& round : rounds_for(ngpus)) {
    if (bidir) {  // both directions of every pair at once
      std::vector<Xfer> xs;
      for (auto pr : round) { xs.push_back({pr.first, pr.second}); xs.push_back({pr.second, pr.first}); }
      int rc = run_xfers(ctx, xs, bytes, mode, iters, &r, st);
      if (rc) return done(rc);
      for (size_t i = 0; i < xs.size(); ++i) gbps[xs[i].src * ngpus + xs[i].dst] = r[i];
    } else {
      for (int phase = 0; phase < 2; ++phase) {
        std::vector<Xfer> xs;
        for (auto pr : round) xs.push_back(phase == 0 ? Xfer{pr.first, pr.second} : Xfer{pr.second, pr.first});
        int rc = run_xfers(ctx, xs, bytes, mode, iters, &r, st);
        if (rc) return done(rc);
        for (size_t i = 0; i < xs.size(); ++i) gbps[xs[i].src * ngpus + xs[i].dst] = r[i];
      }
    }
  }
  // a write probe must actually have landed: check what the last writer left in each arena
  if (write && ngpus > 1) {
    const auto rounds = rounds_for(ngpus);
    std::vector<int> last_writer((size_t)ngpus, -1);
    for (const auto& round : rounds)
      for (auto pr : round) { last_writer[(size_t)pr.second] = pr.first; last_writer[(size_t)pr.first] = pr.second; }
    // within a round, phase order (uni) is a->b then b->a; in both modes each arena's last writer is its partner
    for (int i = 0; i < ngpus; ++i) {
      Device* d = ctx->devs[(size_t)i].get();
      CU_OR_RETURN(cudaSetDevice(i));
      CU_OR_RETURN(cudaMemsetAsync(d->d_count, 0, sizeof(unsigned long long), d->stream));
      p2p_check_kernel<<<d->sm_count * 4, 256, 0, d->stream>>>(reinterpret_cast<const u64x4*>(d->arena), bytes / 32,
                                                              0xC0DE0000ull + (uint64_t)last_writer[(size_t)i], d->d_count);
      CU_OR_RETURN(cudaGetLastError());
      CU_OR_RETURN(cudaMemcpyAsync(d->h_count, d->d_count, sizeof(unsigned long long), cudaMemcpyDeviceToHost, d->stream));
      CU_OR_RETURN(cudaStreamSynchronize(d->stream));
      if (*d->h_count) {
        set_error("peer writes into device %d did not land intact (%llu bad vectors)", i, *d->h_count);
        st->data_bad = true;
        return done(GP_ERR_CUDA);
      }
    }
  }
  // a peer copy must have landed too: every landing half must verify as the address pattern of its source
  if (copy_engine) {
    for (int i = 0; i < ngpus; ++i) {
      Device* d = ctx->devs[(size_t)i].get();
      CU_OR_RETURN(cudaSetDevice(i));
      CU_OR_RETURN(launch_pattern_verify(reinterpret_cast<const uint64_t*>(d->arena + bytes), bytes / 8, GP_PAT_ADDRESS, 0,
                                         d->scratch, d->d_results, d->sm_count, d->stream));
      CU_OR_RETURN(cudaMemcpyAsync(d->h_results, d->d_results, sizeof(gp_pattern_result), cudaMemcpyDeviceToHost, d->stream));
      CU_OR_RETURN(cudaStreamSynchronize(d->stream));
      if (d->h_results[0].mismatches) {
        set_error("peer copies into device %d did not land intact (%llu bad words)", i, (unsigned long long)d->h_results[0].mismatches);
        st->data_bad = true;
        return done(GP_ERR_CUDA);
      }
    }
  }
  return done(GP_OK);
}

struct MultiLock {
  std::lock_guard<std::mutex> multi;
  std::vector<std::unique_lock<std::mutex>> locks;
  MultiLock(Ctx* ctx, int n) : multi(ctx->multi_mu) {
    for (int i = 0; i < n; ++i) locks.emplace_back(ctx->devs[(size_t)i]->mu);  // ordinal order: no deadlock
  }
};

int p2p_probe(Ctx* ctx, int ngpus, size_t bytes, int dir, double* gbps) {
  if (ngpus > (int)ctx->devs.size()) { set_error("ngpus %d > visible %zu", ngpus, ctx->devs.size()); return GP_ERR_NO_DEVICE; }
  MultiLock L(ctx, ngpus);
  P2PState st;
  return p2p_probe_locked(ctx, ngpus, bytes, dir, gbps, &st, ctx->cfg.keep_arena != 0);
}

// ------------------------------------------------------------------ NCCL ----
struct NcclApi {
  void* handle = nullptr;
  decltype(&ncclCommInitAll) CommInitAll = nullptr;
  decltype(&ncclCommDestroy) CommDestroy = nullptr;
  decltype(&ncclAllReduce) AllReduce = nullptr;
  decltype(&ncclGroupStart) GroupStart = nullptr;
  decltype(&ncclGroupEnd) GroupEnd = nullptr;
  decltype(&ncclGetErrorString) GetErrorString = nullptr;
  std::map<int, std::vector<ncclComm_t>> comms;  // by rank count
};

static NcclApi* nccl_api(Ctx* ctx) {
  if (ctx->nccl_handle) return reinterpret_cast<NcclApi*>(ctx->nccl_handle);
  // Use the NCCL the process already has, if any (a host that links or has loaded NCCL must not
  // end up with a second copy under the same soname); otherwise the system one, or the path in
  // GPUPROBE_NCCL_LIB.
  void* h = nullptr;
  if (const char* path = getenv("GPUPROBE_NCCL_LIB")) h = dlopen(path, RTLD_NOW | RTLD_LOCAL);  // the one environment variable of a release build
  if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);
  if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_LOCAL);
  if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_LOCAL);
  if (!h) { set_error("libnccl.so.2 not loadable: %s", dlerror()); return nullptr; }
  NcclApi* a = new NcclApi();
  a->handle = h;
  a->CommInitAll = reinterpret_cast<decltype(a->CommInitAll)>(dlsym(h, "ncclCommInitAll"));
  a->CommDestroy = reinterpret_cast<decltype(a->CommDestroy)>(dlsym(h, "ncclCommDestroy"));
  a->AllReduce = reinterpret_cast<decltype(a->AllReduce)>(dlsym(h, "ncclAllReduce"));
  a->GroupStart = reinterpret_cast<decltype(a->GroupStart)>(dlsym(h, "ncclGroupStart"));
  a->GroupEnd = reinterpret_cast<decltype(a->GroupEnd)>(dlsym(h, "ncclGroupEnd"));
  a->GetErrorString = reinterpret_cast<decltype(a->GetErrorString)>(dlsym(h, "ncclGetErrorString"));
  if (!a->CommInitAll || !a->CommDestroy || !a->AllReduce || !a->GroupStart || !a->GroupEnd) {
    set_error("libnccl is missing a required symbol");
    delete a;
    return nullptr;
  }
  ctx->nccl_handle = a;
  return a;
}

void nccl_teardown(Ctx* ctx) {
  if (!ctx->nccl_handle) return;
  NcclApi* a = reinterpret_cast<NcclApi*>(ctx->nccl_handle);
  for (auto& kv : a->comms)
    for (ncclComm_t c : kv.second) a->CommDestroy(c);
  delete a;  // the library stays mapped: unloading NCCL under live CUDA contexts is not worth the risk
  ctx->nccl_handle = nullptr;
}

#define NCCL_OR_RETURN(call)                                                                  \
  do {                                                                                        \
    ncclResult_t r__ = (call);                                                                \
    if (r__ != ncclSuccess) {                                                                 \
      set_error("%s: %s (%d)", #call, api->GetErrorString ? api->GetErrorString(r__) : "?", (int)r__); \
      return GP_ERR_NCCL;                                                                     \
    }                                                                                         \
  } while (0)

__global__ void fill_f32_kernel(float* p, uint64_t n, float v) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) p[i] = v;
}

static int allreduce_probe_locked(Ctx* ctx, int nranks, size_t bytes, uint32_t warmup, uint32_t iters, double* algbw,
                                  double* busbw, bool keep) {
  if (bytes == 0 || bytes % 4) { set_error("bytes must be a non-zero multiple of 4"); return GP_ERR_INVALID; }
  NcclApi* api = nccl_api(ctx);
  if (!api) return GP_ERR_NCCL;
  const uint64_t count = bytes / 4;
  for (int i = 0; i < nranks; ++i) {
    Device* d = ctx->devs[(size_t)i].get();
    int rc = device_prepare(ctx, d);
    if (rc) return rc;
    rc = ensure_arena(ctx, d, std::max<size_t>(bytes, 1 << 20));
    if (rc) return rc;
    d->triad_operands_bytes = 0;
    fill_f32_kernel<<<d->sm_count * 4, 256, 0, d->stream>>>(reinterpret_cast<float*>(d->arena), count, (float)(i + 1));
    CU_OR_RETURN(cudaGetLastError());
    CU_OR_RETURN(cudaStreamSynchronize(d->stream));
  }
  auto it = api->comms.find(nranks);
  if (it == api->comms.end()) {
    std::vector<ncclComm_t> comms((size_t)nranks);
    std::vector<int> devlist((size_t)nranks);
    for (int i = 0; i < nranks; ++i) devlist[(size_t)i] = i;
    NCCL_OR_RETURN(api->CommInitAll(comms.data(), nranks, devlist.data()));
    it = api->comms.emplace(nranks, comms).first;
  }
  std::vector<ncclComm_t>& comms = it->second;
  auto one = [&]() -> int {
    NCCL_OR_RETURN(api->GroupStart());
    for (int i = 0; i < nranks; ++i) {
      Device* d = ctx->devs[(size_t)i].get();
      NCCL_OR_RETURN(api->AllReduce(d->arena, d->arena, count, ncclFloat, ncclSum, comms[(size_t)i], d->stream));
    }
    NCCL_OR_RETURN(api->GroupEnd());
    return GP_OK;
  };
  // correctness: after ONE all-reduce every element is n(n+1)/2 on every rank
  int rc = one();
  if (rc) return rc;
  const float expect = (float)(nranks * (nranks + 1) / 2);
  for (int i = 0; i < nranks; ++i) {
    Device* d = ctx->devs[(size_t)i].get();
    CU_OR_RETURN(cudaSetDevice(i));
    CU_OR_RETURN(cudaStreamSynchronize(d->stream));
    float probe3[3];
    const uint64_t idx[3] = {0, count / 2, count - 1};
    for (int q = 0; q < 3; ++q)
      CU_OR_RETURN(cudaMemcpy(&probe3[q], reinterpret_cast<float*>(d->arena) + idx[q], 4, cudaMemcpyDeviceToHost));
    for (int q = 0; q < 3; ++q)
      if (probe3[q] != expect) { set_error("all-reduce result %g != %g on rank %d", probe3[q], expect, i); return GP_ERR_NCCL; }
  }
  for (uint32_t w = 0; w < warmup; ++w) { rc = one(); if (rc) return rc; }
  for (int i = 0; i < nranks; ++i) {
    Device* d = ctx->devs[(size_t)i].get();
    CU_OR_RETURN(cudaSetDevice(i));
    CU_OR_RETURN(cudaStreamSynchronize(d->stream));
  }
  for (int i = 0; i < nranks; ++i) {
    Device* d = ctx->devs[(size_t)i].get();
    CU_OR_RETURN(cudaSetDevice(i));
    CU_OR_RETURN(cudaEventRecord(d->events[0], d->stream));
  }
  for (uint32_t k = 0; k < iters; ++k) { rc = one(); if (rc) return rc; }
  double worst_ms = 0;
  for (int i = 0; i < nranks; ++i) {
    Device* d = ctx->devs[(size_t)i].get();
    CU_OR_RETURN(cudaSetDevice(i));
    CU_OR_RETURN(cudaEventRecord(d->events[1], d->stream));
    CU_OR_RETURN(cudaStreamSynchronize(d->stream));
    float ms = 0;
    CU_OR_RETURN(cudaEventElapsedTime(&ms, d->events[0], d->events[1]));
    worst_ms = std::max(worst_ms, (double)ms);  // max over ranks, on the device
  }
  const double t = worst_ms * 1e-3 / iters;
  const double alg = (double)bytes / t / 1e9;
  if (algbw) *algbw = alg;
  if (busbw) *busbw = alg * 2.0 * (nranks - 1) / nranks;
  if (!keep)
    for (int i = 0; i < nranks; ++i) free_arena(ctx->devs[(size_t)i].get());
  return GP_OK;
}

int allreduce_probe(Ctx* ctx, int nranks, size_t bytes, uint32_t warmup, uint32_t iters, double* algbw,
                    double* busbw) {
  if (nranks > (int)ctx->devs.size()) { set_error("nranks %d > visible %zu", nranks, ctx->devs.size()); return GP_ERR_NO_DEVICE; }
  MultiLock L(ctx, nranks);
  return allreduce_probe_locked(ctx, nranks, bytes, warmup, iters, algbw, busbw, ctx->cfg.keep_arena != 0);
}

// The fabric verdict: read and write matrices, their spread and worst pair, the all-reduce, graded against
// gp_config's floors.  A transfer that returns or lands the wrong bytes makes the verdict unhealthy (data_ok = 0)
// instead of failing the call: the caller gets the matrices measured so far and the reason in failed.
int fabric_probe(Ctx* ctx, int ngpus, gp_fabric_result* out) {
  const auto t0 = std::chrono::steady_clock::now();
  memset(out, 0, sizeof(*out));
  out->struct_size = (uint32_t)sizeof(*out);
  out->ngpus = (uint32_t)ngpus;
  out->worst_src = out->worst_dst = -1;
  out->data_ok = 1;
  if (ngpus > (int)ctx->devs.size()) { set_error("ngpus %d > visible %zu", ngpus, ctx->devs.size()); return GP_ERR_NO_DEVICE; }
  const size_t bytes = ctx->cfg.fabric_bytes ? (size_t)ctx->cfg.fabric_bytes : ((size_t)1 << 30);
  out->bytes = bytes;
  MultiLock L(ctx, ngpus);
  P2PState st;
  int rc = p2p_probe_locked(ctx, ngpus, bytes, GP_P2P_READ, out->read_gbps, &st, /*keep=*/true);
  if (rc == GP_OK) rc = p2p_probe_locked(ctx, ngpus, bytes, GP_P2P_WRITE, out->write_gbps, &st, /*keep=*/true);
  if (rc != GP_OK && !st.data_bad) {
    for (int i = 0; i < ngpus; ++i) if (!ctx->cfg.keep_arena) free_arena(ctx->devs[(size_t)i].get());
    return rc;
  }
  if (st.data_bad) { out->data_ok = 0; out->failed |= 8u; }
  auto stats = [&](const double* m, double* mn, double* mx, double* mean) {
    double lo = 1e300, hi = 0, sum = 0;
    int cnt = 0;
    for (int i = 0; i < ngpus; ++i)
      for (int j = 0; j < ngpus; ++j) {
        if (i == j) continue;
        const double v = m[i * ngpus + j];
        lo = std::min(lo, v); hi = std::max(hi, v); sum += v; ++cnt;
        if (out->worst_src < 0 || v < out->worst_gbps) { out->worst_gbps = v; out->worst_src = i; out->worst_dst = j; }
      }
    *mn = cnt ? lo : 0; *mx = hi; *mean = cnt ? sum / cnt : 0;
  };
  stats(out->read_gbps, &out->read_min, &out->read_max, &out->read_mean);
  stats(out->write_gbps, &out->write_min, &out->write_max, &out->write_mean);
  if (ngpus >= 2) {
    const double sr = out->read_min > 0 ? out->read_max / out->read_min : 1e30;
    const double sw = out->write_min > 0 ? out->write_max / out->write_min : 1e30;
    out->spread = out->data_ok ? std::max(sr, sw) : 0.0;
    if (out->data_ok) {
      if (ctx->cfg.min_p2p_gbps > 0 && out->worst_gbps < ctx->cfg.min_p2p_gbps) out->failed |= 1u;
      if (ctx->cfg.max_p2p_spread > 0 && out->spread > ctx->cfg.max_p2p_spread) out->failed |= 2u;
    }
    rc = allreduce_probe_locked(ctx, ngpus, bytes, 2, 5, &out->allreduce_algbw_gbps, &out->allreduce_busbw_gbps, /*keep=*/true);
    out->allreduce_error = rc;
    // NCCL not installed is "not measured", not a fabric fault; a failing or slow all-reduce is
    if (rc != GP_OK && rc != GP_ERR_NCCL) out->failed |= 4u;
    if (rc == GP_OK && ctx->cfg.min_allreduce_busbw_gbps > 0 && out->allreduce_busbw_gbps < ctx->cfg.min_allreduce_busbw_gbps)
      out->failed |= 4u;
  } else {
    out->spread = 1.0;
  }
  if (!ctx->cfg.keep_arena)
    for (int i = 0; i < ngpus; ++i) free_arena(ctx->devs[(size_t)i].get());
  out->healthy = out->failed == 0;
  out->duration_s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  return GP_OK;
}

}  // namespace gp
