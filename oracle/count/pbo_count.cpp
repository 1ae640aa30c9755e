/* pbo_count.cpp — CPU ORACLE instrumentation (test infrastructure): the oracle's C sources compiled with `double` replaced by
 * the counting type of count.h -> liboracle_count.so, same C ABI as libpbo.so plus pbo_count_reset / pbo_count_get.
 * Used only by tools/count_oracle_flops.py. */
#include <stdlib.h>
#include <string.h>
#include <pthread.h>
#include <stdint.h>

#include "count.h"
thread_local pbo_opcount g_ops = {0, 0, 0, 0, 0, 0};
#define double cdouble
extern "C" {
#include "../pbo_physics.c"
#include "../pbo_env.c"
#include "../pbo_api.c"
}
#undef double
extern "C" void pbo_count_reset(void) { memset(&g_ops, 0, sizeof(g_ops)); }
extern "C" void pbo_count_get(uint64_t *out6) {
  out6[0] = g_ops.add; out6[1] = g_ops.mul; out6[2] = g_ops.div; out6[3] = g_ops.sqrt_; out6[4] = g_ops.trans; out6[5] = g_ops.cmp;
}
