// TEST INFRASTRUCTURE ONLY (oracle build).  Stand-in for the missing
// `xilinx-tiny-cnn` submodule (@64ee7a27, .SUBMODULES.json status "missing").
// Only what the reference's sw runtime touches: input parsing/scaling and the
// planar->interleaved reorder (cnvW1A1/sw/main_python.cpp:44-65,129;
// lfcW1A1/sw/main_python.cpp:122; foldedmv-offload.h:276-281).  The published
// tiny-cnn parsers scale a byte c to  min + (max-min)*c/255  in float.
#pragma once
#include <vector>
#include <string>
#include <fstream>
#include <cstdint>
#include <cstddef>
#include <chrono>
#include <cstring>
#include <cmath>
#include <algorithm>
#include <iostream>
#include <functional>
namespace tiny_cnn {
typedef float float_t;
typedef std::vector<float_t> vec_t;
typedef size_t label_t;
struct OffloadConvParams {};
struct mse {};
struct adagrad {};
namespace activation { struct identity {}; }
typedef void (*OffloadHandler)(const vec_t &, vec_t &, unsigned int, OffloadConvParams *);
struct offloaded_layer {
  offloaded_layer(size_t, size_t, OffloadHandler, unsigned int, OffloadConvParams *) {}
};
template <typename A> struct chaninterleave_layer {
  size_t ch, px; bool deint; vec_t out;
  chaninterleave_layer(size_t channels, size_t pixelsPerChannel, bool deinterleave) : ch(channels), px(pixelsPerChannel), deint(deinterleave) {}
  const vec_t &forward_propagation(const vec_t &in, size_t) {
    out.resize(in.size());
    for (size_t c = 0; c < ch; ++c)
      for (size_t p = 0; p < px; ++p) {
        if (deint) out[c * px + p] = in[p * ch + c];
        else out[p * ch + c] = in[c * px + p];
      }
    return out;
  }
};
template <typename L, typename O> struct network {
  template <typename T> network &operator<<(const T &) { return *this; }
};
inline void parse_cifar10(const std::string &path, std::vector<vec_t> *imgs, std::vector<label_t> *labels,
                          float_t smin, float_t smax, int, int) {
  std::ifstream f(path, std::ios::binary);
  if (!f) throw "failed to open cifar file";
  std::vector<unsigned char> buf(3072);
  unsigned char lab;
  while (f.read((char *)&lab, 1)) {
    if (!f.read((char *)buf.data(), 3072)) break;
    vec_t v(3072);
    for (int i = 0; i < 3072; ++i) v[i] = smin + (smax - smin) * buf[i] / 255;
    imgs->push_back(v);
    labels->push_back(lab);
  }
}
inline uint32_t be32(std::ifstream &f) { unsigned char b[4]; f.read((char *)b, 4); return (b[0] << 24) | (b[1] << 16) | (b[2] << 8) | b[3]; }
inline void parse_mnist_images(const std::string &path, std::vector<vec_t> *imgs, float_t smin, float_t smax, int, int) {
  std::ifstream f(path, std::ios::binary);
  if (!f) throw "failed to open mnist file";
  uint32_t magic = be32(f), n = be32(f), r = be32(f), c = be32(f);
  if (magic != 0x803) throw "bad mnist magic";
  std::vector<unsigned char> buf(r * c);
  for (uint32_t i = 0; i < n; ++i) {
    f.read((char *)buf.data(), r * c);
    vec_t v(r * c);
    for (size_t k = 0; k < v.size(); ++k) v[k] = (buf[k] / 255.0f) * (smax - smin) + smin;
    imgs->push_back(v);
  }
}
}  // namespace tiny_cnn
