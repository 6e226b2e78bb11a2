/* c_abi_driver.c — the C ABI used from plain C99, as cgo will use it: proves that
 * include/gpuprobe.h is valid C (no C++ leaks), that the struct sizes the C compiler computes
 * match the library's, and drives one small probe when a GPU is present.
 *   gcc -std=c99 -pedantic -Wall -Werror -Iinclude tests/c_abi_driver.c -Lk8s-gpu-device-plugin_b200 -lgpuprobe
 * Exit code: 0 ok, 3 no GPU (still a pass for the CPU suite), anything else is a failure. */
#include <stdio.h>
#include <string.h>

#include "gpuprobe.h"

int main(void) {
  gp_config cfg;
  gp_ctx* ctx = NULL;
  gp_device_result res;
  gp_pattern_result pr;
  char uuid[64];
  int rc;

  if (gp_abi_version() != GP_ABI_VERSION) return 10;
  if (gp_sizeof(0) != sizeof(gp_config) || gp_sizeof(1) != sizeof(gp_pattern_result) ||
      gp_sizeof(2) != sizeof(gp_timing) || gp_sizeof(3) != sizeof(gp_gemm_result) ||
      gp_sizeof(4) != sizeof(gp_device_result) || gp_sizeof(5) != sizeof(gp_kernel_stat) ||
      gp_sizeof(6) != sizeof(gp_nvml_info) || gp_sizeof(7) != sizeof(gp_fabric_result))
    return 11;
  if (sizeof(gp_pattern_result) != 48) return 12;
  if (strcmp(gp_strerror(GP_OK), "ok") != 0) return 13;

  memset(&cfg, 0, sizeof(cfg));
  cfg.struct_size = sizeof(cfg);
  cfg.abi_version = GP_ABI_VERSION;
  cfg.hbm_bytes = 64u << 20;
  cfg.pattern_bytes = 16u << 20;
  cfg.wo_passes = 2;
  cfg.iters = 2;
  cfg.warmup = 1;
  rc = gp_init(&cfg, &ctx);
  if (rc == GP_ERR_CUDA || rc == GP_ERR_NO_DEVICE) {
    printf("no GPU: gp_init -> %s (%s)\n", gp_strerror(rc), gp_last_error_detail());
    return 3;
  }
  if (rc != GP_OK) return 20;
  if (gp_device_count(ctx) < 1 || gp_device_uuid(ctx, 0, uuid, sizeof(uuid)) != GP_OK) return 21;
  rc = gp_probe_device(ctx, uuid, GP_PROBE_COPY | GP_PROBE_TRIAD | GP_PROBE_PATTERN, &res);
  if (rc != GP_OK || !res.healthy || res.struct_size != sizeof(res)) return 22;
  if (res.kernels_s <= 0 || res.alloc_s <= 0 || res.free_s != 0 || res.arena_bytes == 0 || res.free_deferred != 1) return 27;
  if (gp_pattern_fill(ctx, uuid, GP_PAT_ADDRESS, 1, 4096) != GP_OK) return 23;
  if (gp_inject_bitflip(ctx, uuid, 77, 9) != GP_OK) return 24;
  if (gp_pattern_verify(ctx, uuid, GP_PAT_ADDRESS, 1, 4096, &pr) != GP_OK) return 25;
  if (pr.mismatches != 1 || pr.first_bad_index != 77 || pr.first_bad_xor != (1u << 9)) return 26;
  printf("c_abi_driver ok on %s: copy %.0f GB/s, triad %.0f GB/s, verify %.0f GB/s\n", uuid, res.copy_gbps_median,
         res.triad_gbps_median, res.verify_gbps_median);
  gp_shutdown(ctx);
  return 0;
}
