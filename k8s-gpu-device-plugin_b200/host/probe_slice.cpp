// gpuprobe-slice — child process that grades ONE CUDA-visible device and writes the POD
// gp_device_result to stdout.  The parent (GpuProbeBackend::ProbeSlice) sets
// CUDA_VISIBLE_DEVICES=MIG-<uuid> before exec: a CUDA process sees a single MIG compute
// instance at a time, so per-slice probing (device/device_map.go:78-98 visits each slice,
// device/device.go:32-34 names them "i:j") needs a process per slice, not a thread.
//   argv: mask hbm_bytes pattern_bytes wo_passes gemm_mnk iters warmup [--test-hang] [--test-cuda-error N] [--sm-limit N]
// The two --test-* switches exist for the parent's own tests: a child that never answers, and a child whose
// device reports a sticky CUDA fault (injected at the API level, nothing abnormal reaches the GPU).
#include <unistd.h>

#include <cstdlib>
#include <cstring>

#include "../../include/gpuprobe.h"

int main(int argc, char** argv) {
  gp_config cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.struct_size = sizeof(cfg);
  cfg.abi_version = GP_ABI_VERSION;
  const unsigned mask = argc > 1 ? (unsigned)strtoul(argv[1], nullptr, 10) : (unsigned)GP_PROBE_ALL_SINGLE;
  if (argc > 2) cfg.hbm_bytes = strtoull(argv[2], nullptr, 10);
  if (argc > 3) cfg.pattern_bytes = strtoull(argv[3], nullptr, 10);
  if (argc > 4) cfg.wo_passes = (uint32_t)strtoul(argv[4], nullptr, 10);
  if (argc > 5) cfg.gemm_mnk = (uint32_t)strtoul(argv[5], nullptr, 10);
  if (argc > 6) cfg.iters = (uint32_t)strtoul(argv[6], nullptr, 10);
  if (argc > 7) cfg.warmup = (uint32_t)strtoul(argv[7], nullptr, 10);
  int test_cuda_error = 0;
  for (int i = 8; i < argc; ++i) {
    if (!strcmp(argv[i], "--test-hang")) { for (;;) pause(); }
    if (!strcmp(argv[i], "--test-cuda-error") && i + 1 < argc) test_cuda_error = atoi(argv[++i]);
    if (!strcmp(argv[i], "--sm-limit") && i + 1 < argc) cfg.sm_limit = (uint32_t)atoi(argv[++i]);  // size grids as a slice would
  }
  gp_device_result out;
  memset(&out, 0, sizeof(out));
  out.struct_size = sizeof(out);
  out.abi_version = GP_ABI_VERSION;
  gp_ctx* ctx = nullptr;
  int rc = gp_init(&cfg, &ctx);
  if (rc == 0) {
    if (test_cuda_error) gp_inject_cuda_error(ctx, "0", test_cuda_error);
    rc = gp_probe_device(ctx, "0", mask, &out);
    gp_shutdown(ctx);
  }
  if (rc) { out.error = rc; out.healthy = 0; out.struct_size = sizeof(out); }
  const char* p = reinterpret_cast<const char*>(&out);
  size_t off = 0;
  while (off < sizeof(out)) {
    const ssize_t n = write(STDOUT_FILENO, p + off, sizeof(out) - off);
    if (n <= 0) return 2;
    off += (size_t)n;
  }
  return rc ? 1 : 0;
}
