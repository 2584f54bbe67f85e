// TEST INFRASTRUCTURE ONLY (oracle build).  Stand-in for Vivado-HLS `hls_stream.h`
// (absent from /root/reference): an unbounded FIFO, which is
// what C-simulation gives each dataflow stage (every stage runs to completion).
#pragma once
#include <vector>
#include <string>
#include <cstdlib>
#include <iostream>
namespace hls {
template <typename T> class stream {
  std::vector<T> q;
  size_t head = 0;
 public:
  stream() {}
  stream(const char *) {}
  stream(const std::string &) {}
  stream(const stream &) = delete;
  bool empty() const { return head == q.size(); }
  size_t size() const { return q.size() - head; }
  void write(const T &x) { q.push_back(x); }
  T read() {
    if (head == q.size()) { std::cerr << "hls::stream read on empty" << std::endl; std::abort(); }
    T x = q[head++];
    if (head == q.size()) { q.clear(); head = 0; }
    return x;
  }
  void read(T &x) { x = read(); }
  stream &operator<<(const T &x) { write(x); return *this; }
  stream &operator>>(T &x) { x = read(); return *this; }
};
}  // namespace hls
