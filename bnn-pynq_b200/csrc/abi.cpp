// C ABI of libbnn_b200.so (include/bnn_b200.h, section 2) and the input-file parsers the six bnn.py symbols need.
#include "../../include/bnn_b200.h"
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <new>
#include <string>
#include <vector>
#include "engine.hpp"
#include "kernels.cuh"

using namespace bnnb;

struct bnnb_ctx {
  Engine* eng;
  std::string err;
};

static int fail(bnnb_ctx* c, const std::string& m) { if (c) c->err = m; return -1; }
#define GUARD(c) if (!(c) || !(c)->eng) return -1
#define TRY try {
#define CATCH(c, ret) } catch (const std::exception& e) { fail(c, e.what()); return ret; } catch (...) { fail(c, "unknown error"); return ret; }

extern "C" {

bnnb_ctx* bnnb_open(const char* network) {
  NetDesc n;
  if (!network || !make_net(network, &n)) return nullptr;
  try {
    bnnb_ctx* c = new bnnb_ctx;
    c->eng = new Engine(n);
    return c;
  } catch (...) { return nullptr; }
}
void bnnb_close(bnnb_ctx* c) { if (c) { delete c->eng; delete c; } }
const char* bnnb_last_error(bnnb_ctx* c) {
  if (!c) return "null context";
  if (!c->err.empty()) return c->err.c_str();
  return c->eng->last_error().c_str();
}
static int wrap(bnnb_ctx* c, int rc) { if (rc != 0) c->err = c->eng->last_error(); else c->err.clear(); return rc; }

int bnnb_set_devices(bnnb_ctx* c, const int* ids, int n) { GUARD(c); TRY return wrap(c, c->eng->set_devices(ids, n)); CATCH(c, -1) }
int bnnb_num_devices(bnnb_ctx* c) { GUARD(c); return c->eng->num_devices(); }
int bnnb_release(bnnb_ctx* c) { GUARD(c); TRY return wrap(c, c->eng->release()); CATCH(c, -1) }
size_t bnnb_device_bytes(bnnb_ctx* c, int slot) { return c && c->eng ? c->eng->device_bytes(slot) : 0; }
int bnnb_load_parameters(bnnb_ctx* c, const char* dir) { GUARD(c); TRY return wrap(c, c->eng->load_parameters(dir ? dir : "")); CATCH(c, -1) }

int bnnb_classify_cifar(bnnb_ctx* c, const uint8_t* rec, int n, int ncls, int* classes, int16_t* scores, float* usec) {
  GUARD(c);
  if (!c->eng->net().is_cnv) return fail(c, "not a CNV network");
  TRY return wrap(c, c->eng->infer_host(rec, n, ncls, classes, scores, nullptr, usec)); CATCH(c, -1)
}
int bnnb_classify_mnist(bnnb_ctx* c, const uint8_t* px, int n, int ncls, int* classes, uint64_t* words, float* usec) {
  GUARD(c);
  if (c->eng->net().is_cnv) return fail(c, "not an LFC network");
  TRY return wrap(c, c->eng->infer_host(px, n, ncls, classes, nullptr, words, usec)); CATCH(c, -1)
}
int bnnb_classify_packed(bnnb_ctx* c, const uint8_t* packed, int n, int ncls, int* classes, int16_t* scores, uint64_t* words, float* usec) {
  GUARD(c);
  TRY
  const bool cnv = c->eng->net().is_cnv;
  return wrap(c, c->eng->infer_host_packed(packed, n, ncls, classes, cnv ? scores : nullptr, cnv ? nullptr : words, usec));
  CATCH(c, -1)
}
int bnnb_classify_with_fault_list(bnnb_ctx* c, const uint8_t* input, int n, int ncls, const uint32_t* faults, int nf, int* classes, float* usec) {
  GUARD(c);
  TRY
  std::vector<Fault> fl(nf > 0 ? nf : 0);
  for (int i = 0; i < nf; i++) fl[i] = Fault{faults[5 * i], faults[5 * i + 1], faults[5 * i + 2], faults[5 * i + 3], faults[5 * i + 4]};
  return wrap(c, c->eng->infer_host_with_faults(input, n, ncls, fl, classes, usec));
  CATCH(c, -1)
}
int bnnb_schedule_faults(bnnb_ctx* c, unsigned num_images, unsigned flip_count, int word_size, int target, const int* layers,
                         unsigned num_targets, long long seed, uint32_t* out, int cap) {
  GUARD(c);
  TRY
  std::vector<Fault> f = schedule_random_faults(c->eng->mirror(), num_images, flip_count, word_size, target, layers, num_targets, seed, false);
  const int m = std::min<int>((int)f.size(), cap);
  for (int i = 0; i < m; i++) { out[5 * i] = f[i].time; out[5 * i + 1] = f[i].target; out[5 * i + 2] = f[i].layer; out[5 * i + 3] = f[i].bit; out[5 * i + 4] = f[i].word_size; }
  return m;
  CATCH(c, -1)
}
int bnnb_inject_fault(bnnb_ctx* c, unsigned target, unsigned layer, unsigned bit, unsigned ws) { GUARD(c); TRY return wrap(c, c->eng->inject_fault(target, layer, bit, ws)); CATCH(c, -1) }
uint64_t bnnb_mem_read(bnnb_ctx* c, unsigned tl, unsigned mem, unsigned ind, unsigned th) { if (!c) return 0; return c->eng->mem_read(tl, mem, ind, th); }
int bnnb_mem_set(bnnb_ctx* c, unsigned tl, unsigned mem, unsigned ind, unsigned th, uint64_t v) { GUARD(c); TRY return wrap(c, c->eng->mem_set(tl, mem, ind, th, v)); CATCH(c, -1) }
unsigned bnnb_weight_layer_bits(bnnb_ctx* c, int l) { if (!c || l < 0 || l >= c->eng->net().nlayers) return 0; return c->eng->mirror().weight_layer_bits(l); }
unsigned bnnb_thresh_layer_bits(bnnb_ctx* c, int l) { if (!c || l < 0 || l >= c->eng->net().nlayers) return 0; return c->eng->mirror().thresh_layer_bits(l); }

size_t bnnb_tap_acc_size(bnnb_ctx* c) { return c ? c->eng->tap_acc_size() : 0; }
size_t bnnb_tap_act_size(bnnb_ctx* c) { return c ? c->eng->tap_act_size() : 0; }
int bnnb_tap(bnnb_ctx* c, const uint8_t* input, int n, int32_t* acc, int8_t* act) { GUARD(c); TRY return wrap(c, c->eng->tap_host(input, n, acc, act)); CATCH(c, -1) }

int bnnb_infer_device(bnnb_ctx* c, int slot, const void* d_in, int n, void* d_out, int* d_cls, int ncls, void* stream) {
  GUARD(c);
  TRY return wrap(c, c->eng->infer_device(slot, d_in, n, d_out, d_cls, ncls, (cudaStream_t)stream)); CATCH(c, -1)
}
int bnnb_infer_device_static(bnnb_ctx* c, int slot, const void* d_in, int n, void* d_out, int* d_cls, int ncls, void* stream) {
  GUARD(c);
  TRY return wrap(c, c->eng->infer_device_static(slot, d_in, n, d_out, d_cls, ncls, (cudaStream_t)stream)); CATCH(c, -1)
}
int bnnb_prep_device(bnnb_ctx* c, int slot, const void* d_raw, int n, void* d_in, void* stream) {
  GUARD(c);
  TRY return wrap(c, c->eng->prep_device(slot, d_raw, n, d_in, (cudaStream_t)stream)); CATCH(c, -1)
}
int bnnb_sync(bnnb_ctx* c, int slot) { GUARD(c); return wrap(c, c->eng->sync(slot)); }
int bnnb_last_launches(bnnb_ctx* c) { GUARD(c); return c->eng->last_launches(); }
int bnnb_in_bytes(bnnb_ctx* c) { GUARD(c); return c->eng->net().in_bytes; }
int bnnb_out_bytes(bnnb_ctx* c) { GUARD(c); return c->eng->net().out_bytes; }

int bnnb_set_core(bnnb_ctx* c, int core) { GUARD(c); c->eng->set_core(core); return 0; }
int bnnb_profile_enable(bnnb_ctx* c, int on) { GUARD(c); c->eng->profile_enable(on != 0); return 0; }
int bnnb_profile_collect(bnnb_ctx* c, float* ms, int* launches) { GUARD(c); return c->eng->profile_collect(ms, launches); }

void* bnnb_alloc_host(size_t bytes) {
  void* p = nullptr;
  if (cudaHostAlloc(&p, bytes ? bytes : 16, cudaHostAllocPortable) != cudaSuccess) return nullptr;
  return p;
}
void bnnb_free_host(void* p) { if (p) cudaFreeHost(p); }

}  // extern "C"
