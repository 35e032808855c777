// gguf_fuzz.cpp -- sanitizer fuzz of the GGUF container parser (csrc/lamm_gguf.cpp) without CUDA: the handful of
// lammcu_* functions the parser calls are stubbed here, the fixture's metadata is mutated (random bytes, bit flips,
// 0xFF runs, truncation) and every file that still opens has all its tensors and row ranges read end to end.
//   g++ -std=c++17 -g -O1 -fsanitize=address,undefined -fno-sanitize-recover=undefined -Iinclude \
//       tools/gguf_fuzz.cpp la-llama.cpp_b200/csrc/lamm_gguf.cpp -o /tmp/gguf_fuzz
//   /tmp/gguf_fuzz tests/golden/gguf_fixture_v3.gguf 100000 <seed>
// Round 1: 320 000 mutated files over 4 seeds, no ASan / UBSan report (about 12 % still parse).  Not part of the product.
// Round 2: a fourth mutation writes dimension-sized 64-bit values (2^62 + 8, 2^63 - 1, 2^32 ...) over the metadata; see
// profiles/r2s/gguf_fuzz.txt.
#include "../include/lammcu.h"
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cstdarg>
#include <vector>
#include <string>
namespace lamm { void set_error(const char *fmt, ...) { (void)fmt; } }
extern "C" {
int lammcu_blck_size(int t) { switch (t) { case 0: return 1; case 2: case 3: case 6: case 7: case 8: case 9: return 32; case 10: case 11: case 12: case 13: case 14: case 15: return 256; default: return 0; } }
size_t lammcu_type_size(int t) { switch (t) { case 0: return 4; case 2: return 18; case 3: return 20; case 6: return 22; case 7: return 24; case 8: return 34; case 9: return 36; case 10: return 84; case 11: return 110; case 12: return 144; case 13: return 176; case 14: return 210; case 15: return 292; default: return 0; } }
size_t lammcu_row_size(int t, int64_t k) { int b = lammcu_blck_size(t); return b ? (size_t)(k / b) * lammcu_type_size(t) : 0; }
int lammcu_type_supported(int t) { return lammcu_type_size(t) != 0 && t != 9 && t != 15; }
int lammcu_weight_register(const void *, int, int64_t, int64_t, int64_t, int64_t, int64_t, lammcu_weight_t *) { return -1; }
}
int main(int argc, char **argv) {
  FILE *f = fopen(argv[1], "rb"); std::vector<unsigned char> blob; int c; while ((c = fgetc(f)) != EOF) blob.push_back((unsigned char)c); fclose(f);
  unsigned seed = (unsigned)atoi(argv[3]); long ok = 0, bad = 0; unsigned long sum = 0;
  for (int it = 0; it < atoi(argv[2]); it++) {
    std::vector<unsigned char> b = blob;
    int nmut = 1 + rand_r(&seed) % 4;
    for (int m = 0; m < nmut; m++) {
      size_t pos = rand_r(&seed) % 960;            // metadata region
      int kind = rand_r(&seed) % 4;
      if (kind == 3) {                             // round 2: dimension-sized values whose products wrap 64 / 128 bits
        static const unsigned long long hostile[] = {(1ull << 62) + 8, (1ull << 63) - 1, 1ull << 32, 1ull << 61,
                                                     (1ull << 63) - 256, (1ull << 62) + 32, 1ull << 48, 0xFFFFFFFFFFFFFFFFull};
        const unsigned long long v = hostile[rand_r(&seed) % 8];
        if (pos + 8 <= b.size()) memcpy(&b[pos], &v, 8);
      } else
      if (kind == 0) b[pos] = (unsigned char)rand_r(&seed);
      else if (kind == 1) b[pos] ^= 1u << (rand_r(&seed) % 8);
      else { for (int j = 0; j < 8 && pos + j < b.size(); j++) b[pos + j] = 0xFF; }
    }
    if (rand_r(&seed) % 5 == 0) b.resize(rand_r(&seed) % b.size());
    f = fopen("/tmp/lamm_gguf_fuzz_cur.gguf", "wb"); fwrite(b.data(), 1, b.size(), f); fclose(f);
    lammcu_gguf_t g;
    if (lammcu_gguf_open("/tmp/lamm_gguf_fuzz_cur.gguf", &g) != 0) { bad++; continue; }
    ok++;
    for (int64_t i = 0; i < lammcu_gguf_tensor_count(g); i++) {
      const char *name; int t, nd; int64_t ne[4]; const void *d; uint64_t nb;
      lammcu_gguf_tensor(g, i, &name, &t, &nd, ne, &d, &nb);
      if (nb) {                                    // the recorded size is the exact product, never a wrapped one
        unsigned __int128 want = (unsigned __int128)(nb / ((uint64_t)ne[1] * ne[2] * ne[3] ? (uint64_t)ne[1] * ne[2] * ne[3] : 1));
        if ((unsigned __int128)want * (uint64_t)ne[1] * (uint64_t)ne[2] * (uint64_t)ne[3] != nb || nb > (1ull << 20)) { printf("size check wrapped\n"); abort(); }
      }
      for (uint64_t j = 0; j < nb; j += 97) sum += ((const unsigned char *)d)[j];
      if (nb) sum += ((const unsigned char *)d)[nb - 1];
      sum += strlen(name);
      const void *rd; uint64_t rn;
      if (lammcu_gguf_rows(g, i, 0, ne[1], &rd, &rn) == 0 && rn) sum += ((const unsigned char *)rd)[rn - 1];
    }
    char buf[64]; lammcu_gguf_get_str(g, "general.name", buf, 64); uint32_t u; lammcu_gguf_get_u32(g, "llama.block_count", &u);
    lammcu_gguf_close(g);
  }
  printf("parsed %ld refused %ld (sum %lu)\n", ok, bad, sum);
}
