// lamm_sibling.h -- the bookkeeping of the resident-server host path's SIBLING SPECULATION, free of CUDA and of I/O so
// that it runs on a CPU (tests/test_sibling_logic.py drives it through tests/sibling_logic_test.cpp).  The protocol
// around it -- what a job is, how jobs are posted and answered -- is mul_mat_host_server in lammcu_api.cu.
//
// ggml calls wq, wk, wv with one src1 and gate, up with another (llama.cpp-b2430/llama.cpp:5726-5800), one synchronous
// lamm_mul_mat after the other.  The tracker learns, per weight, which weight was called NEXT with a bit-identical row;
// the next time that weight is called the host posts the jobs of up to kMaxFollowers predicted followers behind its own
// job.  A following call is a HIT only if it is for the weight at the head of that queue and its row equals the row
// those jobs used; anything else empties the queue (the jobs still run, their results are ignored).
//
// W is the weight record; it must have the members
//     W *sib;  unsigned long long sib_gen;  int type;  int64_t K, row_begin, row_end;
// `gen` is a generation counter the owner bumps whenever a weight may have gone away: a link is followed only while the
// generation it was learned in is still the current one, so no dangling handle is ever dereferenced.
#pragma once
#include <cstdint>

namespace lamm {

template <class W>
struct SibTracker {
  static constexpr int kMaxFollowers = 2;
  struct Spec {
    unsigned int tag;   // tag of the job that was posted for `w`
    W *w;
  };
  Spec spec[kMaxFollowers];
  int n_spec = 0;       // speculative jobs posted behind the last real job and not yet consumed, in tag order
  W *prev = nullptr;    // weight of the previous call
  unsigned long long prev_gen = 0;

  // nothing posted is trusted any more (server stopped / relaunched, or a misprediction)
  void drop_posted() { n_spec = 0; }
  void reset() {
    n_spec = 0;
    prev = nullptr;
  }

  // Every call: the previous call's weight was followed by `w`, with the same row (row_eq) or not.
  void learn(W *w, bool row_eq, unsigned long long gen) {
    if (prev && prev_gen == gen) {
      prev->sib = row_eq ? w : nullptr;
      prev->sib_gen = gen;
    }
    prev = w;
    prev_gen = gen;
  }

  // Is this call answered by a job that is already posted?  `same`: its row equals the row the posted jobs used AND
  // the instance that was given those jobs still holds that row.  On a hit the job's tag is returned and the queue
  // advances; on anything else the queue is emptied.
  bool take_hit(const W *w, bool same, unsigned int *tag) {
    if (n_spec <= 0) return false;
    if (same && spec[0].w == w) {
      *tag = spec[0].tag;
      for (int i = 1; i < n_spec; i++) spec[i - 1] = spec[i];
      n_spec--;
      return true;
    }
    n_spec = 0;
    return false;
  }

  // The followers to post behind w's job: the chain w -> sib -> sib ... while every link is current and the follower
  // has w's type and K (it reuses the staged row and the shared-memory plan) and fits the result buffers.
  int followers(W *w, unsigned long long gen, int64_t max_rows, W **out) const {
    int n = 0;
    W *v = w;
    while (n < kMaxFollowers) {
      W *nx = (v->sib && v->sib_gen == gen) ? v->sib : nullptr;
      if (!nx || nx->type != w->type || nx->K != w->K || nx->row_end - nx->row_begin > max_rows) break;
      out[n++] = nx;
      v = nx;
    }
    return n;
  }

  // The jobs of `f[0 .. n)` were posted under the tags first_tag, first_tag + 1, ...
  void posted(unsigned int first_tag, W *const *f, int n) {
    for (int i = 0; i < n && i < kMaxFollowers; i++) {
      spec[i].tag = first_tag + (unsigned int)i;
      spec[i].w = f[i];
    }
    n_spec = n < kMaxFollowers ? n : kMaxFollowers;
  }
};

}  // namespace lamm
