// CPU test of the sibling-speculation bookkeeping (la-llama.cpp_b200/csrc/lamm_sibling.h), driven by
// tests/test_sibling_logic.py.  The weight record is a stand-in with the members the tracker needs; "posting" a job is
// just taking the next tag.  Scenarios follow ggml's call order for a decoder layer (llama.cpp:5726-5800): wq, wk, wv on
// one row, wo on another, gate and up on a third, down on a fourth.
#include <cstdint>
#include <cstdio>
#include <vector>

#include "lamm_sibling.h"

struct W {
  W *sib = nullptr;
  unsigned long long sib_gen = 0;
  int type = 3;
  int64_t K = 4096, row_begin = 0, row_end = 4096;
  const char *name = "";
};
using Tracker = lamm::SibTracker<W>;

static int failures = 0;
#define CHECK(c) do { if (!(c)) { printf("FAILED line %d: %s\n", __LINE__, #c); failures++; } } while (0)

struct Host {                       // the part of mul_mat_host_server that talks to the tracker
  Tracker t;
  unsigned long long gen = 1;
  unsigned int seq = 0;
  const void *row = nullptr;        // identity of the row the last posted job carried
  int hits = 0, posted_spec = 0;
  // one call: returns true when it was answered by a job posted ahead of it
  bool call(W *w, const void *x, int64_t max_rows = 1 << 18) {
    const bool row_eq = (x == row), same = row_eq;
    t.learn(w, row_eq, gen);
    unsigned int tag = 0;
    if (t.take_hit(w, same, &tag)) { hits++; return true; }
    row = x;
    W *fol[Tracker::kMaxFollowers];
    const int nf = t.followers(w, gen, max_rows, fol);
    const unsigned int s = ++seq;
    seq += (unsigned int)nf;
    t.posted(s + 1u, fol, nf);
    posted_spec += nf;
    return false;
  }
};

int main() {
  W q, k, v, o, gate, up, down;
  q.name = "q"; k.name = "k"; v.name = "v"; o.name = "o"; gate.name = "gate"; up.name = "up"; down.name = "down";
  gate.row_end = up.row_end = 11008;
  down.K = 11008;
  int ra, rb, rc, rd;                 // four distinct rows per token (only their addresses matter)
  W *layer[7] = {&q, &k, &v, &o, &gate, &up, &down};
  const void *rows[7] = {&ra, &ra, &ra, &rb, &rc, &rc, &rd};

  {  // 1. a regular layer: nothing on the first token, k / v / up answered ahead from the second token on
    Host h;
    for (int tok = 0; tok < 4; tok++) {
      int hits0 = h.hits;
      for (int i = 0; i < 7; i++) {
        const bool hit = h.call(layer[i], rows[i]);
        if (tok == 0) CHECK(!hit);
        else CHECK(hit == (i == 1 || i == 2 || i == 5));
      }
      CHECK(h.hits - hits0 == (tok == 0 ? 0 : 3));
      h.row = nullptr;                 // a new token brings new rows: nothing equals the previous call's row
    }
    CHECK(q.sib == &k && k.sib == &v && v.sib == nullptr && gate.sib == &up && up.sib == nullptr && o.sib == nullptr);
    CHECK(h.t.n_spec == 0);            // every posted job was consumed
  }
  {  // 2. a follower skipped: the queue is emptied, results of the posted jobs are never used, and the link is relearned
    Host h;
    for (int i = 0; i < 7; i++) h.call(layer[i], rows[i]);
    h.row = nullptr;
    CHECK(!h.call(&q, &ra));           // posts k and v behind q
    CHECK(h.t.n_spec == 2 && h.t.spec[0].w == &k && h.t.spec[1].w == &v && h.t.spec[1].tag == h.t.spec[0].tag + 1);
    CHECK(!h.call(&v, &ra));           // v although k was predicted: miss
    CHECK(h.t.n_spec == 0 && q.sib == &v);
    CHECK(!h.call(&k, &ra));           // ... and k is no hit either (the queue is gone)
    CHECK(v.sib == &k);
  }
  {  // 3. the predicted weight, but another row: miss; the link dies because the rows differed
    Host h;
    for (int i = 0; i < 3; i++) h.call(layer[i], rows[i]);
    h.row = nullptr;
    h.call(&q, &ra);
    CHECK(h.t.n_spec == 2);
    CHECK(!h.call(&k, &rb));
    CHECK(q.sib == nullptr);
    // the queue behind q is gone; the call was posted normally, with ITS follower (k -> v is still believed)
    CHECK(h.t.n_spec == 1 && h.t.spec[0].w == &v);
  }
  {  // 4. the row is equal but the instance no longer holds it (`same` false): no hit, the link survives
    Tracker t;
    W a, b;
    t.learn(&a, false, 1);
    t.learn(&b, true, 1);
    CHECK(a.sib == &b);
    W *fol[2];
    CHECK(t.followers(&a, 1, 1 << 18, fol) == 1 && fol[0] == &b);
    t.posted(10, fol, 1);
    unsigned int tag = 0;
    CHECK(!t.take_hit(&b, false, &tag) && t.n_spec == 0);
    CHECK(a.sib == &b);
  }
  {  // 5. generations: a link learned before a weight may have gone away is never followed (nor dereferenced)
    Tracker t;
    W a, b;
    t.learn(&a, false, 1);
    t.learn(&b, true, 1);
    W *fol[2];
    CHECK(t.followers(&a, 2, 1 << 18, fol) == 0);
    t.reset();
    t.learn(&a, false, 2);             // prev was reset: nothing is written through a stale pointer
    CHECK(a.sib == &b && a.sib_gen == 1);
    t.learn(&b, true, 2);
    CHECK(a.sib == &b && a.sib_gen == 2 && t.followers(&a, 2, 1 << 18, fol) == 1);
  }
  {  // 6. followers must share type and K and fit the result buffer; the chain is cut at the first that does not
    Tracker t;
    W a, b, c;
    t.learn(&a, false, 1); t.learn(&b, true, 1); t.learn(&c, true, 1);
    W *fol[2];
    CHECK(t.followers(&a, 1, 1 << 18, fol) == 2 && fol[0] == &b && fol[1] == &c);
    c.K = 2048;
    CHECK(t.followers(&a, 1, 1 << 18, fol) == 1);
    c.K = 4096; b.type = 8;
    CHECK(t.followers(&a, 1, 1 << 18, fol) == 0);
    b.type = 3; b.row_end = 1 << 20;
    CHECK(t.followers(&a, 1, 1 << 18, fol) == 0);
  }
  {  // 7. the same weight twice in a row on one row: a self link; at most kMaxFollowers jobs are posted
    Host h;
    W a;
    int r;
    CHECK(!h.call(&a, &r));
    CHECK(!h.call(&a, &r));            // learned a -> a with this call; nothing was posted for it
    CHECK(a.sib == &a && h.t.n_spec == 2);
    CHECK(h.call(&a, &r) && h.call(&a, &r));
    CHECK(h.t.n_spec == 0);
  }
  printf("sibling logic: %s\n", failures ? "FAILED" : "ok");
  return failures ? 1 : 0;
}
