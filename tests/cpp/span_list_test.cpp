// Randomised check of ballista_b200/csrc/span_list.h against a byte map (tests/test_span_list.py builds and runs it).
#include <cstdio>
#include <cstdlib>
#include <random>
#include <vector>

#include "span_list.h"

#define CHECK(c)                                                  \
  do {                                                            \
    if (!(c)) {                                                   \
      std::printf("FAILED %s at line %d\n", #c, __LINE__);        \
      return 1;                                                   \
    }                                                             \
  } while (0)

int main() {
  const size_t CAP = 1 << 16;
  b2::SpanList sl;
  sl.reset(CAP);
  std::vector<char> used(CAP, 0);
  struct Block { size_t off, len; };
  std::vector<Block> live;
  std::mt19937_64 rng(7);
  size_t o, l;
  sl.largest(&o, &l);
  CHECK(o == 0 && l == CAP);
  CHECK(!sl.take(0, CAP + 1));   // larger than the span
  CHECK(!sl.take(256, 256));     // not the start of a span
  CHECK(!sl.take(0, 0));
  for (int step = 0; step < 200000; ++step) {
    const bool alloc = live.empty() || (rng() % 100) < 55;
    if (alloc) {
      sl.largest(&o, &l);
      // the byte map agrees: [o, o + l) is free and is maximal
      size_t best = 0, run = 0;
      for (size_t i = 0; i < CAP; ++i) {
        run = used[i] ? 0 : run + 1;
        if (run > best) best = run;
      }
      CHECK(l == best);
      for (size_t i = o; i < o + l; ++i) CHECK(!used[i]);
      if (l == 0) continue;
      size_t want = 256 * (1 + rng() % 24);
      if (want > l) {
        CHECK(!sl.take(o, want));
        continue;
      }
      CHECK(sl.take(o, want));
      for (size_t i = o; i < o + want; ++i) used[i] = 1;
      live.push_back({o, want});
    } else {
      size_t k = rng() % live.size();
      Block b = live[k];
      live[k] = live.back();
      live.pop_back();
      sl.give(b.off, b.len);
      for (size_t i = b.off; i < b.off + b.len; ++i) used[i] = 0;
    }
    size_t free_map = 0;
    for (size_t i = 0; i < CAP; i += 256) free_map += used[i] ? 0 : 256;
    if (step % 64 == 0) {
      CHECK(sl.free_bytes() == free_map);
      // coalesced: as many spans as there are maximal free runs
      size_t runs = 0;
      for (size_t i = 0; i < CAP; i += 256)
        if (!used[i] && (i == 0 || used[i - 256])) ++runs;
      CHECK(sl.n_spans() == runs);
    }
  }
  for (auto& b : live) sl.give(b.off, b.len);
  sl.largest(&o, &l);
  CHECK(o == 0 && l == CAP && sl.n_spans() == 1);
  std::printf("span_list: all ok\n");
  return 0;
}
