/* The C ABI of libdtg_b200.so from plain C (C99): the entry points that need no GPU.
 *
 *   gcc -std=c99 -I ../../include host_check.c -L ../../distributed-tensorflow-guide_b200/ops \
 *       -ldtg_b200 -Wl,-rpath,$PWD/../../distributed-tensorflow-guide_b200/ops -o host_check
 *
 * A trainer written in C/C++ calls the kernels the same way (see train_step.c for the device
 * side): raw pointers, a cudaStream_t passed as void*, an int status, dtg_last_error() for text.
 */
#include <stdio.h>
#include <string.h>

#include "dtg_b200.h"

int main(void) {
  int info[6];
  int rc;
  printf("libdtg_b200 version %d\n", dtg_version());
  /* CRC-32C of "123456789" is the standard check value 0xE3069283 */
  printf("crc32c %08X\n", dtg_crc32c("123456789", 9, 0u));
  /* the tall-pair GEMM's static schedule for 8192 x 8192 on 74 CTA pairs */
  rc = dtg_tall_schedule_check(8192, 8192, 74, 0.62f, 8, info);
  if (rc != 0) {
    printf("schedule check failed: %s\n", dtg_last_error());
    return 1;
  }
  printf("tall schedule: %d big m-blocks, %d big + %d small tiles, %d extra rounds, %d pairs, load %.2f\n",
         info[0], info[1], info[2], info[3], info[4], info[5] / 1000.0);
  rc = dtg_tile_order_check(64, 32, 8);
  printf("tile order 64 x 32, group 8: %s\n", rc == 0 ? "bijection" : dtg_last_error());
  /* errors come back as a status plus a message */
  rc = dtg_tall_schedule_check(0, 1, 1, 1.0f, 1, info);
  printf("bad argument -> status %d (%s)\n", rc, dtg_last_error());
  return rc == 1 ? 0 : 1;
}
