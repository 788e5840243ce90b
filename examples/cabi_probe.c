/* Plain-C consumer of the C-ABI (include/colxlip_b200.h): what a non-Python binding of the reference's loss API
 * links against.  Calls only the entry points that need no device (version, error strings, layout helpers, the
 * device check) so that it also runs on a build host without a GPU:
 *   gcc -std=c99 -Iinclude examples/cabi_probe.c -o /tmp/cabi_probe -Lcolxlip_b200 -lcolxlip_b200 \
 *       -Wl,-rpath,$PWD/colxlip_b200 && /tmp/cabi_probe
 * Compute entry points follow the same pattern with device pointers and a cudaStream_t. */
#include <stdio.h>
#include <string.h>

#include "colxlip_b200.h"

int main(void) {
  int fails = 0;
  printf("cxl_version            %d\n", cxl_version());
  fails += cxl_version() < 100;
  /* image-major argmax: pitch = Bt*T rounded up to 16 elements, 1 byte per element up to 256 patches */
  fails += cxl_argmax_ld(5, 77) != 400;
  fails += cxl_argmax_elem_bytes(196) != 1 || cxl_argmax_elem_bytes(576) != 2;
  /* bf16 image operand: row stride == P from 128 patches on, rounded up to 8 below */
  fails += cxl_img_row_stride(196) != 196 || cxl_img_row_stride(49) != 56;
  fails += strcmp(cxl_strerror(CXL_OK), "ok") != 0;
  fails += strlen(cxl_strerror(CXL_ERR_NO_DEVICE)) == 0;
  {
    const int rc = cxl_device_check();
    printf("cxl_device_check       %d (%s)\n", rc, cxl_strerror(rc));
    fails += rc != CXL_OK && rc != CXL_ERR_NO_DEVICE;
  }
  /* argument validation happens before any device work: null pointers are rejected */
  fails += cxl_maxsim_fwd(NULL, NULL, 1, 1, 1, 1, 64, CXL_DTYPE_BF16, CXL_MEAN_MASKED, NULL, NULL, NULL, NULL) !=
           CXL_ERR_INVALID_ARG;
  fails += cxl_l2norm_fwd(NULL, CXL_DTYPE_F32, 1, 8, NULL, CXL_DTYPE_BF16, NULL, NULL) != CXL_ERR_INVALID_ARG;
  /* workspace queries are pure arithmetic: the fixed-order score sums of cxl_maxsim_fwd_ws keep (T+30)/32 extra
   * [Bt,Bi] f32 slots (T = 77 -> 3), cxl_infonce_grad_ws one float per block, the statistics pass 2 floats per
   * (128-row block, column); the allocation-free statistics entry refuses what exceeds its built-in 16 MiB */
  fails += cxl_maxsim_fwd_workspace_bytes(64, 32, 77, CXL_DTYPE_BF16) != (size_t)3 * 64 * 32 * 4;
  fails += cxl_infonce_grad_workspace_bytes(100, 4096) != (size_t)2 * 100 * 4;
  fails += cxl_infonce_workspace_bytes(256, 1000) != (size_t)2 * 2 * 1000 * 4;
  fails += cxl_infonce_workspace_bytes(16384, 32768) <= CXL_INFONCE_STATIC_SCRATCH_BYTES;
  fails += strstr(cxl_strerror(CXL_ERR_STATIC_SCRATCH), "static scratch") == NULL;
  fails += cxl_maxsim_fwd_ws(NULL, NULL, 1, 1, 1, 1, 64, CXL_DTYPE_BF16, CXL_MEAN_MASKED | CXL_MEAN_ROUND16, NULL, NULL, NULL,
                             NULL, 0, NULL) != CXL_ERR_WORKSPACE;
  printf("kernel launches so far %llu\n", cxl_kernel_launches());
  printf(fails ? "CABI_PROBE FAILED (%d)\n" : "CABI_PROBE OK\n", fails);
  return fails != 0;
}
