// One of these per network (compiled with -DBNNB_NET=\"<network>\"): lib/python_cuda-<network>.so exports the six symbols
// bnn/bnn.py binds through cffi (bnn/bnn.py:69-76) and forwards to libbnn_b200.so.  It mirrors, entry point by entry point,
// bnn/src/network/<network>/sw/main_python.cpp (file parsing included: tiny_cnn parse_cifar10 / parse_mnist_images semantics).
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <vector>
#include "../../include/bnn_b200.h"

#ifndef BNNB_NET
#error "compile with -DBNNB_NET=\"cnvW1A1\" etc."
#endif

static bnnb_ctx* g_ctx = nullptr;
static const bool g_is_cnv = BNNB_NET[0] == 'c';

static bnnb_ctx* ctx() {  // FoldedMVInit: idempotent, called at the top of every entry point (main_python.cpp:126,149,190)
  if (!g_ctx) g_ctx = bnnb_open(BNNB_NET);
  return g_ctx;
}
static bool quiet() { const char* q = std::getenv("BNN_B200_QUIET"); return q && *q && *q != '0'; }
static void report(const char* where) { std::fprintf(stderr, "[bnn_b200 %s] %s: %s\n", BNNB_NET, where, bnnb_last_error(g_ctx)); }

// CIFAR-10 binary: 3073-byte records until EOF, a trailing partial record is dropped (tiny_cnn parse_cifar10).
// MNIST idx3-ubyte: big-endian magic 2051, count, rows, cols, then count*rows*cols bytes (parse_mnist_images).
static bool read_images(const char* path, std::vector<uint8_t>* buf, int* n) {
  FILE* f = std::fopen(path, "rb");
  if (!f) { std::fprintf(stderr, "[bnn_b200 %s] could not open %s\n", BNNB_NET, path); return false; }
  std::fseek(f, 0, SEEK_END);
  long sz = std::ftell(f);
  std::fseek(f, 0, SEEK_SET);
  bool ok = true;
  if (g_is_cnv) {
    *n = (int)(sz / 3073);
    buf->resize((size_t)*n * 3073);
    ok = *n == 0 || std::fread(buf->data(), 3073, *n, f) == (size_t)*n;
  } else {
    unsigned char h[16];
    ok = std::fread(h, 1, 16, f) == 16;
    if (ok) {
      auto be = [&](int o) { return (uint32_t)h[o] << 24 | (uint32_t)h[o + 1] << 16 | (uint32_t)h[o + 2] << 8 | h[o + 3]; };
      const uint32_t magic = be(0), cnt = be(4), r = be(8), c = be(12);
      if (magic != 0x803 || r * c != 784) { std::fprintf(stderr, "[bnn_b200 %s] %s is not a 28x28 idx3-ubyte file\n", BNNB_NET, path); ok = false; }
      else {
        *n = (int)cnt;
        buf->assign((size_t)cnt * 784, 0);
        const size_t got = std::fread(buf->data(), 1, buf->size(), f);
        (void)got;  // a short file leaves zero pixels, as the reference's unchecked read does
      }
    }
  }
  std::fclose(f);
  return ok;
}

static void print_rate(float usec) {  // the two lines the reference prints (foldedmv-offload.h:411-412)
  if (quiet()) return;
  std::printf("Inference took %.2f microseconds per image\nClassification rate: %.1f images per second\n", usec, usec > 0 ? 1e6 / usec : 0.0);
}

extern "C" {

void load_parameters(const char* path) {
  if (!ctx()) { std::fprintf(stderr, "[bnn_b200 %s] cannot create context\n", BNNB_NET); return; }
  if (!quiet()) std::printf("Setting network weights and thresholds in accelerator...\n");
  if (bnnb_load_parameters(g_ctx, path) != 0) report("load_parameters");
}

int inference(const char* path, int results[64], int number_class, float* usecPerImage) {
  if (!ctx()) return -1;
  std::vector<uint8_t> buf;
  int n = 0;
  if (!read_images(path, &buf, &n) || n < 1) return -1;
  float usec = 0;
  if (g_is_cnv) {
    int16_t sc[64];
    if (bnnb_classify_cifar(g_ctx, buf.data(), 1, number_class, nullptr, sc, &usec) != 0) { report("inference"); return -1; }
    const int ncls = number_class < 0 ? 0 : (number_class > 64 ? 64 : number_class);  // BlackBoxJam writes 64 scores per image
    int best = 0;  // std::max_element over the number_class scores: first maximum (cnvW1A1/sw/main_python.cpp:138)
    for (int j = 0; j < ncls; j++) { if (results) results[j] = sc[j]; if (sc[j] > sc[best]) best = j; }
    if (usecPerImage) *usecPerImage = usec;
    print_rate(usec);
    return best;
  }
  uint64_t w = 0;
  if (bnnb_classify_mnist(g_ctx, buf.data(), 1, number_class, nullptr, &w, &usec) != 0) { report("inference"); return -1; }
  // testPrebinarized_nolabel (foldedmv-offload.cpp:140-166): mask to number_class bits, one-hot at round(log2(word)) over 64 entries
  const uint64_t mask = number_class >= 64 ? ~0ull : (number_class <= 0 ? 0 : (~0ull >> (64 - number_class)));
  w &= mask;
  const double r = w ? std::round(std::log2((double)w)) : 0;
  int best = 0; bool found = false;
  for (unsigned i = 0; i < 64; i++) {
    const int v = i == (unsigned)r;
    if (results) results[i] = v;  // the reference writes all 64 entries
    if (v && !found) { best = (int)i; found = true; }
  }
  if (usecPerImage) *usecPerImage = usec;
  print_rate(usec);
  return best;
}

// Shared by the path-based and the in-memory entry points: images already parsed (CNV: n 3073-byte records; LFC: n x 784 pixels).
static int* run_multiple(const uint8_t* data, int n, int number_class, float* usecPerImage, int enable_detail, const char* where) {
  float usec = 0;
  int* result = nullptr;
  if (g_is_cnv) {
    std::vector<int> cls(n);
    std::vector<int16_t> sc(enable_detail ? (size_t)n * 64 : 0);
    if (bnnb_classify_cifar(g_ctx, data, n, number_class, cls.data(), enable_detail ? sc.data() : nullptr, &usec) != 0) { report(where); return nullptr; }
    if (enable_detail) {  // image-major scores (cnvW1A1/sw/main_python.cpp:161-163)
      result = new (std::nothrow) int[(size_t)n * number_class + 1];
      if (result) for (int i = 0; i < n; i++) for (int j = 0; j < number_class; j++) result[(size_t)i * number_class + j] = j < 64 ? sc[(size_t)i * 64 + j] : 0;
    } else {
      result = new (std::nothrow) int[n + 1];
      if (result) std::memcpy(result, cls.data(), (size_t)n * 4);
    }
  } else {  // LFC ignores enable_detail (lfcW1A1/sw/main_python.cpp:135-156)
    result = new (std::nothrow) int[n + 1];
    if (result && bnnb_classify_mnist(g_ctx, data, n, number_class, result, nullptr, &usec) != 0) { report(where); delete[] result; return nullptr; }
  }
  if (usecPerImage) *usecPerImage = usec;
  print_rate(usec);
  return result;
}

static int* run_multiple_with_faults(const uint8_t* data, int n, int number_class, float* usecPerImage, unsigned int flip_count, int word_size,
                                     int target, int* target_layers, unsigned int num_targets, const char* where) {
  long long seed = -1;  // std::random_device like the reference, unless pinned for reproducible campaigns
  if (const char* s = std::getenv("BNN_B200_FAULT_SEED")) seed = std::atoll(s);
  std::vector<uint32_t> fl((size_t)flip_count * 5 + 5);
  const int nf = n > 0 ? bnnb_schedule_faults(g_ctx, (unsigned)n, flip_count, word_size, target, target_layers, target_layers ? num_targets : 0, seed, fl.data(), (int)flip_count) : 0;
  if (nf < 0) { report(where); return nullptr; }
  int* result = new (std::nothrow) int[n + 1];
  float usec = 0;
  if (result && bnnb_classify_with_fault_list(g_ctx, data, n, number_class, fl.data(), nf, result, &usec) != 0) { report(where); delete[] result; return nullptr; }
  if (usecPerImage) *usecPerImage = usec;  // time of the last image's segment
  return result;
}

int* inference_multiple(const char* path, int number_class, int* image_number, float* usecPerImage, int enable_detail) {
  if (!ctx()) return nullptr;
  std::vector<uint8_t> buf;
  int n = 0;
  if (!read_images(path, &buf, &n)) return nullptr;
  int* result = run_multiple(buf.data(), n, number_class, usecPerImage, enable_detail, "inference_multiple");
  if (result && image_number) *image_number = n;
  return result;
}

int* inference_multiple_with_faults(const char* path, int number_class, int* image_number, float* usecPerImage, unsigned int flip_count,
                                    int word_size, int target, int* target_layers, unsigned int num_targets) {
  if (!ctx()) return nullptr;
  std::vector<uint8_t> buf;
  int n = 0;
  if (!read_images(path, &buf, &n)) return nullptr;
  int* result = run_multiple_with_faults(buf.data(), n, number_class, usecPerImage, flip_count, word_size, target, target_layers, num_targets,
                                         "inference_multiple_with_faults");
  if (result && image_number) *image_number = n;
  return result;
}

// In-memory twins of the two calls above: the caller hands over what the file would have held after parsing, so that
// classify_images (bnn/bnn.py:280-287) needs no temp file.  Not part of the reference's cdef; bnn.py binds them when present.
// The caller states how many bytes it hands over: anything but image_count whole rows of THIS network's size (3073 for CNV, 784
// for LFC) is refused before a single byte is read.
static bool buffer_ok(const unsigned char* images, int image_count, long long image_bytes, const char* where) {
  const long long row = g_is_cnv ? 3073 : 784;
  if (image_count >= 0 && (image_count == 0 || images) && image_bytes == row * image_count) return true;
  std::fprintf(stderr, "[bnn_b200 %s] %s: expected %d images x %lld bytes = %lld bytes, got %lld\n", BNNB_NET, where, image_count, row,
               row * (long long)image_count, image_bytes);
  return false;
}

int* inference_buffer(const unsigned char* images, int image_count, long long image_bytes, int number_class, float* usecPerImage, int enable_detail) {
  if (!ctx()) return nullptr;
  if (!buffer_ok(images, image_count, image_bytes, "inference_buffer")) return nullptr;
  return run_multiple(images, image_count, number_class, usecPerImage, enable_detail, "inference_buffer");
}

int* inference_buffer_with_faults(const unsigned char* images, int image_count, long long image_bytes, int number_class, float* usecPerImage,
                                  unsigned int flip_count, int word_size, int target, int* target_layers, unsigned int num_targets) {
  if (!ctx()) return nullptr;
  if (!buffer_ok(images, image_count, image_bytes, "inference_buffer_with_faults")) return nullptr;
  return run_multiple_with_faults(images, image_count, number_class, usecPerImage, flip_count, word_size, target, target_layers, num_targets,
                                  "inference_buffer_with_faults");
}

void free_results(int* result) { delete[] result; }

void deinit(void) {
  // FoldedMVDeinit (rawhls-offload.cpp:69-74) frees the accelerator buffers; the parameters (and injected faults) survive it in the
  // reference because they live in static memories (top.cpp:51-68).  Same here: every device allocation (parameter operands,
  // activation and staging buffers, streams) is released, the host mirror of the parameters and faults is kept, and the next call
  // - possibly from another PynqBNN holding this cached library (bnn/bnn.py:79,109-115) - rebuilds the device state on demand.
  if (g_ctx) bnnb_release(g_ctx);
}

}  // extern "C"
