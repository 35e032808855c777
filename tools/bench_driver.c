/* bench_driver.c -- the calling loop of bench.py's e2e leg, in C.
 *
 * ggml calls lamm_can_mul_mat / lamm_mul_mat from C, once per mul_mat node
 * (llama.cpp-b2430/ggml.c:10858-10863).  Driving 225 such calls per token through Python ctypes adds
 * about a microsecond of interpreter overhead to each ~12 us call; this loop makes exactly the same
 * calls, on the same ggml_compute_params / ggml_tensor structs, without it.  Not part of the product:
 * bench.py loads it if it has been built (__graft_entry__.build) and falls back to the Python loop.
 */
#include <stdbool.h>
#include <time.h>

typedef bool (*can_fn)(const void *params, const void *dst);
typedef void (*mm_fn)(const void *params, void *dst);

/* `tokens` passes over the n (params, dst) pairs; returns seconds of wall clock, or -1.0 - i if the gate
 * refused call i */
double bench_drive_calls(can_fn can, mm_fn mm, void *const *params, void *const *dsts, int n, int tokens) {
  struct timespec a, b;
  clock_gettime(CLOCK_MONOTONIC, &a);
  for (int t = 0; t < tokens; t++)
    for (int i = 0; i < n; i++) {
      if (!can(params[i], dsts[i])) return -1.0 - i;
      mm(params[i], dsts[i]);
    }
  clock_gettime(CLOCK_MONOTONIC, &b);
  return (double)(b.tv_sec - a.tv_sec) + 1e-9 * (double)(b.tv_nsec - a.tv_nsec);
}
