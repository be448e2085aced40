// span_list.h — free list of the exchange's receive arena (comm.h: RecvArena): byte spans of one allocation, kept
// coalesced; a block is always taken from the START of a free span (the offer a rank publishes is the start and length
// of its largest free span) and may be given back from any thread.  Plain C++, no CUDA: tests/test_span_list.py
// compiles it with g++ and checks it against a byte map.
#pragma once
#include <cstddef>
#include <iterator>
#include <map>
#include <mutex>

namespace b2 {

struct SpanList {
  std::mutex mu;
  std::map<size_t, size_t> free_spans;  // offset -> length; no two spans touch

  void reset(size_t cap) {
    std::lock_guard<std::mutex> g(mu);
    free_spans.clear();
    if (cap) free_spans[0] = cap;
  }
  void largest(size_t* off, size_t* len) {
    std::lock_guard<std::mutex> g(mu);
    *off = 0;
    *len = 0;
    for (auto& s : free_spans)
      if (s.second > *len) {
        *off = s.first;
        *len = s.second;
      }
  }
  bool take(size_t off, size_t len) {  // [off, off + len) must lie at the start of a free span
    std::lock_guard<std::mutex> g(mu);
    auto it = free_spans.find(off);
    if (len == 0 || it == free_spans.end() || it->second < len) return false;
    size_t rest = it->second - len;
    free_spans.erase(it);
    if (rest) free_spans[off + len] = rest;
    return true;
  }
  void give(size_t off, size_t len) {
    if (len == 0) return;
    std::lock_guard<std::mutex> g(mu);
    auto nx = free_spans.lower_bound(off);
    if (nx != free_spans.end() && off + len == nx->first) {  // merge with the span behind
      len += nx->second;
      nx = free_spans.erase(nx);
    }
    if (nx != free_spans.begin()) {  // merge with the span in front
      auto pv = std::prev(nx);
      if (pv->first + pv->second == off) {
        pv->second += len;
        return;
      }
    }
    free_spans[off] = len;
  }
  size_t free_bytes() {
    std::lock_guard<std::mutex> g(mu);
    size_t n = 0;
    for (auto& s : free_spans) n += s.second;
    return n;
  }
  size_t n_spans() {
    std::lock_guard<std::mutex> g(mu);
    return free_spans.size();
  }
};

}  // namespace b2
