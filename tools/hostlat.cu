// hostlat.cu -- what a synchronous host call costs on this box, independent of our kernels (diagnostics).
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o /tmp/hostlat tools/hostlat.cu && /tmp/hostlat
//
//  (a) empty kernel + cudaStreamSynchronize                      -- the launch/completion floor of host mode 0
//  (b) the same kernel writing one word to pinned host memory   -- adds the kernel-end flush over PCIe
//  (c) ping-pong with a RESIDENT kernel through pinned memory: the host writes a sequence number, one GPU
//      thread polls it over PCIe and answers in another pinned word              -- the floor of host mode 1/2
//  (d) as (c), but the GPU thread first reads a second pinned location after seeing the number (the
//      "job, then X" dependent read of host mode 1)                               -- one more PCIe round trip
// Every GPU-side poll is bounded (~2 s), so a lost message cannot hang the device.
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while (0)

__global__ void empty_kernel() {}
__global__ void flag_kernel(volatile unsigned int *f, unsigned int v) { *f = v; }

__global__ void pong_kernel(volatile unsigned int *req, volatile unsigned int *extra, volatile unsigned int *ack,
                            unsigned int n, int dependent) {
  for (unsigned int s = 1; s <= n; s++) {
    unsigned long long t0, now;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    while (*req != s) {
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
      if (now - t0 > 2000000000ull) return;
    }
    unsigned int v = s;
    if (dependent) v += *extra;                  // a second, dependent PCIe read
    *ack = v;
  }
}

static double now_us() {
  return std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

int main() {
  CK(cudaSetDevice(0));
  cudaStream_t st;
  CK(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
  unsigned int *pin;
  CK(cudaMallocHost(&pin, 4096));
  volatile unsigned int *req = pin, *extra = pin + 256, *ack = pin + 512;
  const int N = 2000;
  for (int i = 0; i < 100; i++) { empty_kernel<<<1, 32, 0, st>>>(); CK(cudaStreamSynchronize(st)); }
  double t0 = now_us();
  for (int i = 0; i < N; i++) { empty_kernel<<<1, 32, 0, st>>>(); cudaStreamSynchronize(st); }
  printf("(a) empty kernel + stream sync            : %6.2f us per call\n", (now_us() - t0) / N);
  t0 = now_us();
  for (int i = 0; i < N; i++) { flag_kernel<<<1, 32, 0, st>>>(ack, (unsigned)i); cudaStreamSynchronize(st); }
  printf("(b) + one word written to pinned memory   : %6.2f us per call\n", (now_us() - t0) / N);
  for (int dep = 0; dep < 2; dep++) {
    *req = 0; *ack = 0; *extra = 0;
    pong_kernel<<<1, 1, 0, st>>>(req, extra, ack, (unsigned)N, dep);
    t0 = now_us();
    int lost = 0;
    for (unsigned int s = 1; s <= (unsigned)N && !lost; s++) {
      *req = s;
      const double w0 = now_us();
      while (*ack != s)
        if (now_us() - w0 > 1e6) { lost = 1; break; }
    }
    const double per = (now_us() - t0) / N;
    CK(cudaStreamSynchronize(st));
    printf("(%c) resident-kernel ping-pong%s: %6.2f us per round trip%s\n", dep ? 'd' : 'c',
           dep ? ", job then dependent read " : "                           ", per, lost ? "  (LOST A MESSAGE)" : "");
  }
  return 0;
}
