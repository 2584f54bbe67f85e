#include "engine.hpp"
#include "f4_common.cuh"
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>

namespace bnnb {

// CUDA call -> error string + return -1 (E: the std::string that receives the message)
#define CKE(E, call)                                                                               \
  do {                                                                                             \
    cudaError_t e_ = (call);                                                                       \
    if (e_ != cudaSuccess) {                                                                       \
      (E) = std::string(#call) + ": " + cudaGetErrorString(e_);                                    \
      return -1;                                                                                   \
    }                                                                                              \
  } while (0)
#define CK(call) CKE(err_, call)

static constexpr int kLanes = 4;  // chunk pipelines per device (see Lane)

// Chunk pipelines ("lanes") per device, each an in-order stream of its own: H2D copy -> quantise/pack -> layer kernels -> results.
// The copy of chunk i+1..i+3 overlaps the compute of chunk i, and consecutive chunks' kernels overlap their ramp-up / ramp-down
// (every kernel here is a persistent full-GPU kernel with a prologue of several microseconds: one in-order compute stream for all
// chunks cost the host-buffer path 17 %).  The parameter stream (`compute`) carries uploads and fault patches; lanes wait for it
// by event.

struct Lane {
  cudaStream_t s = nullptr;
  uint8_t* d_raw = nullptr;   // chunk x raw input bytes (CIFAR records / MNIST pixels)
  void* d_in = nullptr;       // chunk x layer-0 input
  uint32_t* act[2] = {nullptr, nullptr};  // ping-pong activations
  void* d_out = nullptr;      // chunk x out_bytes
  int* d_cls = nullptr;
  int cap = 0;                // images the staging buffers (d_raw, d_out, d_cls) hold
  int act_cap = 0;            // images d_in / act hold
  size_t act_size = 0;        // bytes of each act buffer (the size per image depends on the compute core, see act_bytes)
  bool used = false;
  // results of the chunk in flight, copied back when the lane is next needed (a device-to-pageable copy blocks the host: issuing it
  // right after the chunk would stop the feeder from running ahead)
  int pend_off = -1, pend_n = 0;
};

struct DeviceState {
  int id = 0, num_sms = 148;
  bool ready = false;         // streams, parameter buffers and the parameter upload exist
  cudaStream_t compute = nullptr;  // parameter stream: uploads, fault patches, taps, default stream of the device-resident path
  Lane lane[kLanes];
  DevLayer L[9];              // parameter set 0: what every path uses; the state that persists between calls
  DevLayer* d_layers = nullptr;
  unsigned char* arena = nullptr;  // set 0's arrays and its DevLayer table live in one allocation, so that a set is cloned by one copy
  size_t arena_bytes = 0;
  // Fault epochs: lane k >= 1 works on its own clone of the parameters (taken from set 0 at the start of the call) and applies every
  // patch to it in its own image order, so lanes never have to be ordered against each other (run_shard_faulted)
  struct ParamSet { unsigned char* arena = nullptr; DevLayer L[9]; DevLayer* d_layers = nullptr; } clone[kLanes - 1];
  DevPatch* d_patches = nullptr;
  size_t patch_cap = 0;
  int* d_status = nullptr;    // set by a kernel whose mbarrier wait timed out: mapped pinned host memory (device alias of h_status)
  int* h_status = nullptr;
  std::vector<void*> params;  // parameter-lifetime allocations
  std::vector<cudaEvent_t> tev;
  size_t bytes = 0;
  // The one-image call (bnn.py's inference()): record -> device, quantise, every layer, class -> host as ONE CUDA graph launch on
  // lane 0 instead of ten stream operations; key = everything the captured nodes hold by value (run_single).
  struct SingleGraph {
    cudaGraphExec_t exec = nullptr;
    std::vector<uint64_t> key;
    bool warm = false, disabled = false;
    uint8_t* h_in = nullptr;   // pinned staging: one record in, scores / word + class out
    uint8_t* h_out = nullptr;
  } single;

  void* alloc(size_t n, bool param) {
    void* p = nullptr;
    if (cudaMalloc(&p, n ? n : 16) != cudaSuccess) return nullptr;
    if (param) params.push_back(p);
    bytes += n ? n : 16;
    return p;
  }
  template <typename T> void release(T*& p, size_t n) {
    if (p) { cudaFree(p); bytes -= n ? n : 16; }
    p = nullptr;
  }
  void destroy() {
    if (!ready && !compute) return;
    cudaSetDevice(id);
    if (compute) cudaStreamSynchronize(compute);
    for (auto& ln : lane) {
      if (ln.s) { cudaStreamSynchronize(ln.s); cudaStreamDestroy(ln.s); }
      cudaFree(ln.d_raw); cudaFree(ln.d_in); cudaFree(ln.act[0]); cudaFree(ln.act[1]); cudaFree(ln.d_out); cudaFree(ln.d_cls);
      ln = Lane();
    }
    if (compute) cudaStreamDestroy(compute);
    compute = nullptr;
    cudaFree(d_patches);
    d_patches = nullptr; patch_cap = 0;
    for (void* p : params) cudaFree(p);
    params.clear();
    arena = nullptr; arena_bytes = 0;
    for (auto& c : clone) c = ParamSet();
    for (cudaEvent_t e : tev) cudaEventDestroy(e);
    tev.clear();
    if (single.exec) cudaGraphExecDestroy(single.exec);
    if (single.h_in) cudaFreeHost(single.h_in);
    if (single.h_out) cudaFreeHost(single.h_out);
    single = SingleGraph();
    if (h_status) cudaFreeHost(h_status);
    d_layers = nullptr; d_status = nullptr; h_status = nullptr;
    bytes = 0;
    ready = false;
  }
  ~DeviceState() { destroy(); }
};

// What run_chunk reports / needs when a parity test taps the layers.
struct TapCtx {
  int32_t* acc = nullptr;   // device, layer-major over the chunk: [layer][img][ofm^2][N]
  uint4* act0 = nullptr;    // fused CNV layers 0+1: layer 0's output as the operand rows layer 1 reads
  bool tensor = false;      // taps from the tensor-core kernels (TAP instantiations) instead of the LOP3+POPC kernels
  int ofmt[9] = {};         // activation format every layer wrote
  int raw[9] = {};          // 1: acc holds the MMA sum (needs the un-folding of tap_host), 0: the reference accumulator itself
  bool fused01 = false;
};

struct Engine::Segments {   // fault epochs of one call: patches in fault order with their times
  std::vector<Patch> patches;
  std::vector<uint32_t> ptime;
};

// Fault epochs of one shard, PER KERNEL: a fault on layer l changes nothing below layer l, so only the kernel that holds layer l is
// split at that fault's time; every other kernel runs whole chunks.  group = index of the kernel's last layer (the fused layer-0/1
// kernel is group 1).  The device patch array is ordered by group, time order kept inside a group.
struct FaultPlan {
  int group_of_layer[9] = {};
  std::vector<uint32_t> gtime[9];   // patch times of the group, ascending
  size_t goff[9] = {};              // the group's offset in the device patch array
  size_t gcur[kLanes][9] = {};      // per lane (= per parameter set): next patch of the group to apply
};

static int round_cap(int images) {  // buffer capacities grow in powers of two from 128 images
  int c = 128;
  while (c < images) c <<= 1;
  return c;
}

Engine::Engine(const NetDesc& net) : net_(net), pm_(new ParamMirror(net)) {
  if (const char* e = std::getenv("BNN_B200_CORE")) core_ = std::strcmp(e, "popc") == 0 ? 0 : (std::strcmp(e, "i8") == 0 ? 1 : 2);
  if (const char* e = std::getenv("BNN_B200_FUSE01")) fuse01_ = std::atoi(e) != 0;
  if (const char* e = std::getenv("BNN_B200_CHAIN")) chain_ = std::atoi(e) != 0;
}

Engine::~Engine() { dev_.clear(); }

int Engine::release() {
  for (auto& d : dev_) d->destroy();
  {
    std::lock_guard<std::mutex> lk(prof_mu_);
    for (auto& e : prof_) { cudaEventDestroy(e.a); cudaEventDestroy(e.b); }
    prof_.clear();
  }
  return 0;
}

size_t Engine::device_bytes(int slot) const { return slot >= 0 && slot < (int)dev_.size() ? dev_[slot]->bytes : 0; }

int Engine::check_status(DeviceState& ds, std::string& err) {
  // the flag lives in mapped pinned memory: after the stream synchronisation that precedes every call of this function the host
  // reads what a kernel wrote, without a copy
  const int st = ds.h_status ? *reinterpret_cast<volatile int*>(ds.h_status) : 0;
  if (st != 0) {
    *reinterpret_cast<volatile int*>(ds.h_status) = 0;  // reported once; the next call starts clean
    err = "a tensor-core kernel gave up waiting on an mbarrier (device " + std::to_string(ds.id) + "): results of this call are invalid";
    return -1;
  }
  return 0;
}

// Images per pass through the layer chain.  host = the pipelined host-buffer path, where a chunk is also the unit of H2D/compute
// overlap; otherwise the device-resident path, where larger chunks mean fewer launches and less tail quantisation (the path uses
// under 1% of HBM bandwidth, so nothing is gained by keeping the inter-layer tensors L2-resident there).
int Engine::chunk_images(bool host) const {
  if (const char* e = std::getenv(host ? "BNN_B200_HOST_CHUNK" : "BNN_B200_CHUNK")) { int v = std::atoi(e); if (v > 0) return v; }
  if (!net_.is_cnv) return host ? 32768 : 65536;
  if (core_ >= 1 && fuse01_) return host ? 8192 : 65536;
  const bool tern = net_.L[1].out_mode == OUT_TERN;
  // LOP3+POPC path: layer 0's output (7 200 / 14 400 B per image) exists in global memory; keep it inside the 126 MB L2
  return tern ? 2048 : 4096;
}

// Bytes of ONE activation buffer for chunks of `images`: the largest inter-layer tensor in any format a core may plan for it
// (bit planes everywhere; e2m1 operand planes, 4 bits per value and 128-image tiles, between fp4 kernels of a 1-bit network).
size_t Engine::act_bytes(int images) const {
  size_t per = 16;
  // CNV on the tensor cores: layer 0's output lives in the fused kernel's shared memory, never in these buffers
  const int first = (net_.is_cnv && core_ >= 1 && fuse01_) ? 1 : 0;
  for (int i = first; i + 1 < net_.nlayers; i++) {
    const LayerDesc& l = net_.L[i];
    const size_t px = (size_t)out_dim(l) * out_dim(l);
    per = std::max(per, px * out_words_per_pixel(l) * 4);
    if (l.out_mode == OUT_BIT || l.out_mode == OUT_TERN) per = std::max(per, px * l.N / 2);  // e2m1 codes (ACT_F4P / ACT_F4T)
  }
  // CNV: + 20 images, an ACT_F4I tensor is written in whole groups of the consumer's G <= 19 images
  return (size_t)((images + 127) / 128 * 128 + (net_.is_cnv ? 20 : 0)) * per;
}

int Engine::set_devices(const int* ids, int n) {
  if (!dev_.empty()) { err_ = "set_devices must be called before the first use"; return -1; }
  dev_ids_.assign(ids, ids + n);
  return 0;
}

// The device list.  Nothing is allocated here: a device gets its streams, parameter buffers and parameter upload the first time a
// call runs on it (upload), its activation / staging buffers sized for the chunks it is actually given (ensure_buffers / ensure_lane).
int Engine::ensure_devices() {
  if (!dev_.empty()) return 0;
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess || count == 0) {
    err_ = std::string("no CUDA device: this runtime has no CPU fallback (") + cudaGetErrorString(e) + ")";
    return -1;
  }
  if (dev_ids_.empty()) {
    if (const char* env = std::getenv("BNN_B200_DEVICES")) {
      for (const char* p = env; *p;) { dev_ids_.push_back(std::atoi(p)); while (*p && *p != ',') p++; if (*p) p++; }
    } else {
      for (int i = 0; i < count; i++) dev_ids_.push_back(i);
    }
  }
  std::vector<std::unique_ptr<DeviceState>> built;
  for (int id : dev_ids_) {
    if (id < 0 || id >= count) { err_ = "bad device id"; return -1; }
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, id));
    if (prop.major < 10) { err_ = "this runtime is built for sm_100a (Blackwell) only"; return -1; }
    std::unique_ptr<DeviceState> d(new DeviceState);
    d->id = id;
    d->num_sms = prop.multiProcessorCount;
    built.push_back(std::move(d));
  }
  dev_.swap(built);  // only a complete list is kept
  return 0;
}

// Lays one parameter set out in an arena starting at `base` (256-byte aligned pieces): every layer's folded memories and derived
// operands, then the DevLayer table the patch kernel reads.  Returns the arena size; base may be null to measure only.
static size_t carve_params(const NetDesc& net, unsigned char* base, DevLayer* L, DevLayer** table) {
  size_t off = 0;
  auto take = [&](size_t bytes) { unsigned char* p = base + off; off += (std::max<size_t>(bytes, 16) + 255) / 256 * 256; return p; };
  for (int i = 0; i < net.nlayers; i++) {
    const LayerDesc& l = net.L[i];
    DevLayer& D = L[i];
    D.d = l;
    D.kwords = l.K / 32;
    D.kpad = (l.K + 31) / 32 * 32;
    D.wf = (uint64_t*)take((size_t)l.pe * l.wmem * 8);
    D.tf = (int32_t*)take((size_t)l.pe * l.tmem * std::max(l.nthr, 1) * 4);
    const bool l0 = net.is_cnv && i == 0;
    D.wp = (uint32_t*)take(l0 ? (size_t)64 * 8 * 4 : (size_t)l.wbits * D.kwords * l.N * 4);
    D.w8 = nullptr; D.w4 = nullptr; D.w4s = nullptr; D.thr_tc = nullptr;
    if (!l0) {  // operands of the tensor-core paths
      D.w8 = (int8_t*)take((size_t)l.N * D.kpad);
      D.w4 = (uint8_t*)take((size_t)l.N * D.kpad / 2);
      D.w4s = (uint8_t*)take((size_t)l.N * D.kpad / 2);
      D.thr_tc = (int32_t*)take((size_t)l.N * 2 * 4);
    }
    D.thr = (int32_t*)take((size_t)l.N * 2 * 4);
  }
  *table = (DevLayer*)take(sizeof(DevLayer) * 9);
  return off;
}

// First use of a device (or first use after release / load_parameters): streams, parameter buffers, upload + on-device re-pack.
int Engine::upload(DeviceState& ds, std::string& err) {
  CKE(err, cudaSetDevice(ds.id));
  if (!ds.compute) {
    CKE(err, cudaStreamCreateWithFlags(&ds.compute, cudaStreamNonBlocking));
    for (auto& ln : ds.lane) {
      CKE(err, cudaStreamCreateWithFlags(&ln.s, cudaStreamNonBlocking));
    }
    DevLayer* tbl = nullptr;
    ds.arena_bytes = carve_params(net_, nullptr, ds.L, &tbl);
    ds.arena = (unsigned char*)ds.alloc(ds.arena_bytes, true);
    const bool ok = ds.arena != nullptr;
    if (ok) carve_params(net_, ds.arena, ds.L, &ds.d_layers);
    if (cudaHostAlloc((void**)&ds.h_status, 4, cudaHostAllocMapped | cudaHostAllocPortable) != cudaSuccess) ds.h_status = nullptr;
    if (ds.h_status) { *ds.h_status = 0; if (cudaHostGetDevicePointer((void**)&ds.d_status, ds.h_status, 0) != cudaSuccess) ds.d_status = nullptr; }
    if (!ok || !ds.d_layers || !ds.d_status) { err = "cudaMalloc failed (parameters, device " + std::to_string(ds.id) + ")"; ds.destroy(); return -1; }
    CKE(err, cudaMemcpy(ds.d_layers, ds.L, sizeof(DevLayer) * 9, cudaMemcpyHostToDevice));
  }
  cudaStream_t s = ds.compute;
  for (int i = 0; i < net_.nlayers; i++) {
    CKE(err, cudaMemcpyAsync(const_cast<uint64_t*>(ds.L[i].wf), pm_->weights(i).data(), pm_->weights(i).size() * 8, cudaMemcpyHostToDevice, s));
    CKE(err, cudaMemcpyAsync(const_cast<int32_t*>(ds.L[i].tf), pm_->thresholds(i).data(), pm_->thresholds(i).size() * 4, cudaMemcpyHostToDevice, s));
    launch_repack(ds.L[i], net_.is_cnv && i == 0, s);
  }
  CKE(err, cudaStreamSynchronize(s));
  CKE(err, cudaGetLastError());
  ds.ready = true;
  return 0;
}

int Engine::load_parameters(const std::string& dir) {
  std::unique_ptr<ParamMirror> pm(new ParamMirror(net_));
  std::string e;
  if (pm->load(dir, &e) != 0) { err_ = e; return -1; }
  pm_ = std::move(pm);  // a reload discards injected faults
  loaded_ = true;
  if (ensure_devices() != 0) return -1;
  for (auto& d : dev_)
    if (d->ready && upload(*d, err_) != 0) return -1;  // devices not used yet take the mirror when they are first used
  return 0;
}

// Lane k's layer-0 input and activation ping-pong for chunks of `images` (grown on demand, powers of two).
int Engine::ensure_buffers(DeviceState& ds, int k, int images, std::string& err) {
  Lane& ln = ds.lane[k];
  if (images <= ln.act_cap && act_bytes(images) <= ln.act_size) return 0;
  const int cap = std::max(round_cap(images), ln.act_cap);
  const size_t size = std::max(act_bytes(cap), ln.act_size);  // never shrinks: a later call on another core may need more per image
  CKE(err, cudaStreamSynchronize(ln.s));
  CKE(err, cudaStreamSynchronize(ds.compute));
  ds.release(ln.d_in, (size_t)ln.act_cap * net_.in_bytes);
  for (int j = 0; j < 2; j++) ds.release(ln.act[j], ln.act_size);
  ln.act_cap = 0;
  ln.act_size = 0;
  ln.d_in = ds.alloc((size_t)cap * net_.in_bytes, false);
  ln.act[0] = (uint32_t*)ds.alloc(size, false);
  ln.act[1] = (uint32_t*)ds.alloc(size, false);
  if (!ln.d_in || !ln.act[0] || !ln.act[1]) {
    cudaFree(ln.d_in); cudaFree(ln.act[0]); cudaFree(ln.act[1]);
    ln.d_in = nullptr; ln.act[0] = ln.act[1] = nullptr;
    err = "cudaMalloc failed (activation buffers for " + std::to_string(cap) + " images, device " + std::to_string(ds.id) + ")";
    return -1;
  }
  ln.act_cap = cap;
  ln.act_size = size;
  return 0;
}

// Lane k's staging buffers (raw input, results) for chunks of `images`.
int Engine::ensure_lane(DeviceState& ds, int k, int images, std::string& err) {
  Lane& ln = ds.lane[k];
  if (images <= ln.cap) return 0;
  const int cap = round_cap(images);
  const size_t raw_bytes = net_.is_cnv ? 3073 : 784;
  CKE(err, cudaStreamSynchronize(ln.s));
  ds.release(ln.d_raw, (size_t)ln.cap * raw_bytes);
  ds.release(ln.d_out, (size_t)ln.cap * net_.out_bytes);
  ds.release(ln.d_cls, (size_t)ln.cap * 4);
  ln.cap = 0;
  ln.d_raw = (uint8_t*)ds.alloc((size_t)cap * raw_bytes, false);
  ln.d_out = ds.alloc((size_t)cap * net_.out_bytes, false);
  ln.d_cls = (int*)ds.alloc((size_t)cap * 4, false);
  if (!ln.d_raw || !ln.d_out || !ln.d_cls) {
    cudaFree(ln.d_raw); cudaFree(ln.d_out); cudaFree(ln.d_cls);
    ln.d_raw = nullptr; ln.d_out = nullptr; ln.d_cls = nullptr;
    err = "cudaMalloc failed (staging buffers, device " + std::to_string(ds.id) + ")";
    return -1;
  }
  ln.cap = cap;
  return 0;
}

int Engine::ensure_patch_cap(DeviceState& ds, size_t count, std::string& err) {
  if (count <= ds.patch_cap) return 0;
  CKE(err, cudaStreamSynchronize(ds.compute));
  ds.release(ds.d_patches, ds.patch_cap * sizeof(DevPatch));
  ds.patch_cap = 0;
  const size_t cap = std::max<size_t>(64, count * 2);
  ds.d_patches = (DevPatch*)ds.alloc(cap * sizeof(DevPatch), false);
  if (!ds.d_patches) { err = "cudaMalloc failed (fault patches, device " + std::to_string(ds.id) + ")"; return -1; }
  ds.patch_cap = cap;
  return 0;
}

// Faults outside an inference call (bnnb_inject_fault, bnnb_mem_set): every device that holds parameters patches its operands in
// place; a device that is not in use yet takes the mirror (which already holds the fault) when it is first used.
int Engine::apply_patches(const std::vector<Patch>& patches) {
  if (patches.empty()) return 0;
  if (ensure_devices() != 0) return -1;
  static_assert(sizeof(Patch) == sizeof(DevPatch), "Patch and DevPatch must share a layout");
  for (auto& d : dev_) {
    if (!d->ready) continue;
    CK(cudaSetDevice(d->id));
    for (auto& ln : d->lane) CK(cudaStreamSynchronize(ln.s));
    if (ensure_patch_cap(*d, patches.size(), err_) != 0) return -1;
    CK(cudaMemcpyAsync(d->d_patches, patches.data(), patches.size() * sizeof(DevPatch), cudaMemcpyHostToDevice, d->compute));
    launch_patch_apply(d->d_layers, net_.is_cnv, d->d_patches, (int)patches.size(), d->compute);
    CK(cudaStreamSynchronize(d->compute));  // the host vector may die after return
    CK(cudaGetLastError());
  }
  return 0;
}

int Engine::inject_fault(uint32_t target, uint32_t layer, uint32_t bit, uint32_t word_size) {
  Patch p;
  if (!pm_->inject(target, layer, bit, word_size, &p)) return 0;
  return apply_patches({p});
}
uint64_t Engine::mem_read(unsigned tl, unsigned mem, unsigned ind, unsigned thresh) { return pm_->mem_read(tl, mem, ind, thresh); }
int Engine::mem_set(unsigned tl, unsigned mem, unsigned ind, unsigned thresh, uint64_t val) {
  pm_->mem_set(tl, mem, ind, thresh, val);
  const unsigned li = tl >> 1;
  if ((int)li >= net_.nlayers) return 0;
  const LayerDesc& l = net_.L[li];
  if (mem >= (unsigned)l.pe) return 0;
  Patch p{li, tl & 1, mem, ind, thresh, 0, 0};
  if (tl & 1) {
    if (l.nthr == 0 || ind >= (unsigned)l.tmem || thresh >= (unsigned)l.nthr) return 0;
    p.value = (uint64_t)(int64_t)pm_->thresholds(li)[((size_t)mem * l.tmem + ind) * l.nthr + thresh];
  } else {
    if (ind >= (unsigned)l.wmem) return 0;
    p.value = pm_->weights(li)[(size_t)mem * l.wmem + ind];
  }
  return apply_patches({p});
}

// Kernel family and activation interchange format per layer (f4_common.cuh): on the fp4 core, between two fp4 kernels of a 1-bit
// network, the producer writes e2m1 operand planes in the consumer's own shared-memory layout (staged there by TMA bulk copies).
// per_layer: the LOP3+POPC taps (every layer on the CUDA-core kernels, bit planes everywhere).
void Engine::plan_formats(const DeviceState& ds, bool per_layer, bool tensor_taps, int* fam, int* ofmt, bool* fused01, bool no_tiles) const {
  const bool popc_only = per_layer && !tensor_taps;
  for (int i = 0; i < 9; i++) { fam[i] = 0; ofmt[i] = ACT_BITS; }
  if (core_ == 2 && !popc_only)
    for (int i = 0; i < net_.nlayers; i++) fam[i] = conv_f4_supports(net_.L[i]) ? 1 : (gemm_f4_supports(ds.L[i]) ? 2 : 0);
  *fused01 = core_ >= 1 && net_.is_cnv && !popc_only && fuse01_;
  for (int i = 0; i + 1 < net_.nlayers; i++) {
    const LayerDesc& p = net_.L[i];
    const LayerDesc& c = net_.L[i + 1];
    const bool producer_f4 = fam[i] != 0 || (*fused01 && i == 1 && core_ == 2);
    if (!producer_f4 || !((p.out_mode == OUT_BIT && c.in_mode == IN_XNOR) || (p.out_mode == OUT_TERN && c.in_mode == IN_TERN))) continue;
    if (fam[i + 1] == 1 && fam[i] != 2) ofmt[i] = (no_tiles || (*fused01 && i == 1)) ? ACT_F4P : ACT_F4I;
    // (conv consumer: image planes, interleaved per group of the consumer - except after the fused kernel, whose pool helpers the
    // scattered stores slowed by 3-10 %, and in fault epochs, whose kernel runs start at arbitrary images)
    if (fam[i + 1] == 2 && !p.pool && !(*fused01 && i == 1) && !no_tiles) ofmt[i] = ACT_F4T;   // gemm consumer: 128-image tiles
  }
}

// One chunk through every layer.  Returns the number of kernels launched, or -1.
int Engine::run_chunk(DeviceState& ds, const void* d_in, int n, void* d_out, int* d_classes, int ncls, cudaStream_t s, TapCtx* tap,
                      uint32_t** bufs, std::string& err) {
  const void* cur = d_in;
  int launches = 0;
  const bool per_layer = tap != nullptr;  // when tapping, every layer writes its own buffer; else two ping-pong buffers
  const bool popc_only = per_layer && !tap->tensor;
  int first_layer = 0;
  int fam[9], ofmt[9];
  bool fused01;
  plan_formats(ds, per_layer, per_layer && tap->tensor, fam, ofmt, &fused01);
  bool classes_done = false;
  int32_t* tap_acc = tap ? tap->acc : nullptr;
  auto take_tap = [&](int i) {  // this layer's slice of the tap buffer
    int32_t* p = tap_acc;
    if (tap_acc) tap_acc += (size_t)n * net_.L[i].ofm_dim * net_.L[i].ofm_dim * net_.L[i].N;
    return p;
  };
  auto prof_begin = [&](int layer) {
    ProfEv pe{nullptr, nullptr, layer, ds.id};
    if (profile_) { cudaEventCreate(&pe.a); cudaEventCreate(&pe.b); cudaEventRecord(pe.a, s); }
    return pe;
  };
  auto prof_end = [&](ProfEv& pe) {
    if (!profile_) return;
    cudaEventRecord(pe.b, s);
    std::lock_guard<std::mutex> lk(prof_mu_);
    prof_.push_back(pe);
  };
  if (fused01) {  // layers 0+1 in one tensor-core kernel
    Tap01 tp;
    if (tap) { tp.acc0 = take_tap(0); tp.acc1 = take_tap(1); tp.act0 = tap->act0; }
    ProfEv pe = prof_begin(1);
    if (launch_cnv01_tc(ds.L[0], ds.L[1], (const int8_t*)d_in, bufs[1], n, ds.num_sms, core_ == 2, ofmt[1], ds.d_status, tp, s)) {
      prof_end(pe);
      const cudaError_t le = cudaGetLastError();
      if (le != cudaSuccess) { err = std::string("fused layer 0+1 launch: ") + cudaGetErrorString(le); return -1; }
      launches++;
      cur = bufs[1];
      first_layer = 2;
      if (tap) { tap->fused01 = true; tap->raw[0] = tap->raw[1] = 1; tap->ofmt[0] = ACT_BITS; tap->ofmt[1] = ofmt[1]; }
    } else {
      if (profile_) { cudaEventDestroy(pe.a); cudaEventDestroy(pe.b); }
      // the activation buffers of the ordinary path are sized WITHOUT layer 0's output (act_bytes): no silent fall-back there
      if (!tap) { err = "the fused layer-0/1 kernel refused the launch (" + std::to_string(n) + " images)"; return -1; }
      tap_acc = tap->acc;
      fused01 = false;
    }
  }
  for (int i = first_layer; i < net_.nlayers; i++) {
    const bool last = i == net_.nlayers - 1;
    uint32_t* out = last ? (uint32_t*)d_out : bufs[per_layer ? i : (i & 1)];
    int32_t* tp = take_tap(i);
    const int ifmt = i > 0 ? ofmt[i - 1] : ACT_BITS;
    int raw = 1;
    ProfEv pe = prof_begin(i);
    // the trailing fully-connected layers in one launch (fc_chain.cu), from the first layer whose run qualifies and fits
    if (!per_layer && chain_ && fam[i] == 2 && !last && ifmt == ACT_F4T &&
        launch_fc_chain(&ds.L[i], net_.nlayers - i, (const uint32_t*)cur, (uint32_t*)d_out, n, ds.num_sms, ifmt, ds.d_status, s, d_classes, ncls)) {
      prof_end(pe);
      const cudaError_t le = cudaGetLastError();
      if (le != cudaSuccess) { err = "layers " + std::to_string(i) + "... (chain) launch: " + cudaGetErrorString(le); return -1; }
      launches++;
      if (d_classes) classes_done = true;
      break;
    }
    if (net_.is_cnv && i == 0) { launch_conv0(ds.L[0], (const int8_t*)cur, out, n, tp, ds.num_sms, s); raw = 0; }
    // (an ACT_F4P input - the fused kernel's output - is staged in interleaved row order through a tensor map, conv_f4.cu)
    else if (fam[i] == 1 && launch_conv_f4(ds.L[i], (const uint32_t*)cur, out, n, ds.num_sms, ifmt, ofmt[i], last ? 0 : conv_f4_group(net_.L[i + 1]), ds.d_status, tp, s)) {}
    else if (fam[i] == 2 && launch_gemm_f4(ds.L[i], (const uint32_t*)cur, out, n, ds.num_sms, ifmt, ofmt[i], ds.d_status, tp, s, last ? d_classes : nullptr, ncls)) {
      if (last && d_classes && (net_.is_cnv || net_.L[i].N == 64)) classes_done = true;  // the last layer's epilogue wrote the classes
    }
    else if (core_ >= 1 && !popc_only && launch_conv_tc(ds.L[i], (const uint32_t*)cur, out, n, ds.num_sms, ds.d_status, tp, s)) {}
    else if (core_ >= 1 && !popc_only && launch_gemm_tc(ds.L[i], (const uint32_t*)cur, out, n, ds.num_sms, ds.d_status, tp, s)) {}
    else if (ifmt != ACT_BITS || ofmt[i] != ACT_BITS) { err = "layer " + std::to_string(i) + ": no kernel for the planned activation format"; return -1; }
    else { launch_layer_popc(ds.L[i], (const uint32_t*)cur, out, n, tp, ds.num_sms, s); raw = 0; }
    prof_end(pe);
    {  // a launch that the device refuses (resources, configuration) must not pass silently
      const cudaError_t le = cudaGetLastError();
      if (le != cudaSuccess) { err = "layer " + std::to_string(i) + " launch: " + cudaGetErrorString(le); return -1; }
    }
    if (tap) { tap->raw[i] = raw; tap->ofmt[i] = last ? ACT_BITS : ofmt[i]; }
    launches++;
    cur = out;
  }
  if (net_.is_cnv && d_classes && !classes_done) { launch_cnv_classes((const int16_t*)d_out, n, ncls, d_classes, s); launches++; }
  if (!net_.is_cnv && d_classes && !classes_done) { launch_lfc_classes((const uint64_t*)d_out, n, ncls, d_classes, s); launches++; }
  return launches;
}

// One chunk through every layer with fault epochs (images [img0, img0 + n) of the call), on parameter set PL / d_layers, which
// belongs to `lane`: each kernel is launched once per run of
// images between two fault times OF ITS OWN LAYER(S), the patches due at the start of a run applied just before it on the same
// stream (k_patch_apply rewrites only that layer's operands).  Activations travel as bit planes or per-image e2m1 planes (no
// 128-image tiles), so a run is a pointer offset into the chunk's buffers.
int Engine::run_chunk_faulted(DeviceState& ds, const DevLayer* PL, DevLayer* d_layers, int lane, const void* d_in, int n, long long img0,
                              void* d_out, int* d_classes, int ncls, cudaStream_t s, uint32_t** bufs, FaultPlan& fp, std::string& err) {
  int fam[9], ofmt[9];
  bool fused01;
  plan_formats(ds, false, false, fam, ofmt, &fused01, true);
  const int NL = net_.nlayers;
  auto out_stride = [&](int i) -> size_t {  // bytes per image of what layer i writes
    if (i == NL - 1) return (size_t)net_.out_bytes;
    const LayerDesc& l = net_.L[i];
    const size_t px = (size_t)out_dim(l) * out_dim(l);
    return ofmt[i] == ACT_F4P ? px * l.N / 2 : px * out_words_per_pixel(l) * 4;
  };
  const uint8_t* cur = (const uint8_t*)d_in;
  size_t cur_stride = (size_t)net_.in_bytes;
  for (int i = fused01 ? 1 : 0; i < NL; i++) {
    const bool last = i == NL - 1, fused = fused01 && i == 1;
    uint8_t* out = last ? (uint8_t*)d_out : (uint8_t*)bufs[i & 1];
    const size_t ostr = out_stride(i);
    const int grp = i, ifmt = (i > 0 && !fused) ? ofmt[i - 1] : ACT_BITS;
    const std::vector<uint32_t>& gt = fp.gtime[grp];
    size_t& gc = fp.gcur[lane][grp];
    int a = 0;
    while (a < n) {
      const size_t b = gc;
      while (gc < gt.size() && (long long)gt[gc] <= img0 + a) gc++;
      if (gc > b) launch_patch_apply(d_layers, net_.is_cnv, ds.d_patches + fp.goff[grp] + b, (int)(gc - b), s);
      int nxt = n;
      if (gc < gt.size() && (long long)gt[gc] < img0 + n) nxt = (int)((long long)gt[gc] - img0);
      const int c = nxt - a;
      const uint8_t* in_p = cur + (size_t)a * cur_stride;
      uint32_t* out_p = (uint32_t*)(out + (size_t)a * ostr);
      bool ok = true;
      if (fused) ok = launch_cnv01_tc(PL[0], PL[1], (const int8_t*)in_p, out_p, c, ds.num_sms, core_ == 2, ofmt[1], ds.d_status, Tap01(), s);
      else if (net_.is_cnv && i == 0) launch_conv0(PL[0], (const int8_t*)in_p, out_p, c, nullptr, ds.num_sms, s);
      else if (fam[i] == 1 && launch_conv_f4(PL[i], (const uint32_t*)in_p, out_p, c, ds.num_sms, ifmt, ofmt[i], 0, ds.d_status, nullptr, s)) {}
      else if (fam[i] == 2 && launch_gemm_f4(PL[i], (const uint32_t*)in_p, out_p, c, ds.num_sms, ifmt, ofmt[i], ds.d_status, nullptr, s)) {}
      else if (core_ >= 1 && launch_conv_tc(PL[i], (const uint32_t*)in_p, out_p, c, ds.num_sms, ds.d_status, nullptr, s)) {}
      else if (core_ >= 1 && launch_gemm_tc(PL[i], (const uint32_t*)in_p, out_p, c, ds.num_sms, ds.d_status, nullptr, s)) {}
      else if (ifmt != ACT_BITS || ofmt[i] != ACT_BITS) ok = false;
      else launch_layer_popc(PL[i], (const uint32_t*)in_p, out_p, c, nullptr, ds.num_sms, s);
      const cudaError_t le = cudaGetLastError();
      if (!ok || le != cudaSuccess) { err = "layer " + std::to_string(i) + " launch (fault epochs): " + (ok ? cudaGetErrorString(le) : "no kernel for the planned format"); return -1; }
      a = nxt;
    }
    cur = out;
    cur_stride = ostr;
  }
  if (net_.is_cnv && d_classes) launch_cnv_classes((const int16_t*)d_out, n, ncls, d_classes, s);
  if (!net_.is_cnv && d_classes) launch_lfc_classes((const uint64_t*)d_out, n, ncls, d_classes, s);
  return 0;
}

int Engine::infer_device(int slot, const void* d_in, int n, void* d_out, int* d_classes, int ncls, cudaStream_t stream) {
  if (!loaded_) { err_ = "load_parameters has not been called"; return -1; }
  if (ensure_devices() != 0) return -1;
  if (slot < 0 || slot >= (int)dev_.size()) { err_ = "bad device slot"; return -1; }
  DeviceState& ds = *dev_[slot];
  CK(cudaSetDevice(ds.id));
  if (!ds.ready && upload(ds, err_) != 0) return -1;
  cudaStream_t s = stream ? stream : ds.compute;
  const int chunk = chunk_images();
  if (n > 0 && ensure_buffers(ds, 0, std::min(chunk, n), err_) != 0) return -1;  // lane 0's buffers serve this path
  int launches = 0;
  for (int off = 0; off < n; off += chunk) {
    const int c = std::min(chunk, n - off);
    const int r = run_chunk(ds, (const uint8_t*)d_in + (size_t)off * net_.in_bytes, c, (uint8_t*)d_out + (size_t)off * net_.out_bytes,
                            d_classes ? d_classes + off : nullptr, ncls, s, nullptr, ds.lane[0].act, err_);
    if (r < 0) return -1;
    launches += r;
  }
  last_launches_ = launches;
  CK(cudaGetLastError());
  return 0;
}

// ---- device-resident inference that allocates nothing (no counterpart in the reference) ----
// For callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it): the
// inter-layer activation scratch - the two ping-pong buffers infer_device takes from cudaMalloc - lives in ONE file-scope device
// array of kStaticScratchBytes, which every GPU that loads this library holds until the process exits.  In host code the array's
// name is only a host-side stand-in for the symbol, not a device address: the address is taken per device with
// cudaGetSymbolAddress (device current, return code checked) and that pointer is what the kernels receive.
// Every call shares the array and finds the previous call's contents in it.  Nothing in it has to start at zero: each layer kernel
// writes every activation it later reads, and the rows of a 128-image e2m1 tile beyond the batch hold stale but finite codes that
// feed accumulator rows which are never stored - so nothing is cleared.  Calls that overlap on the GPU (two streams, two contexts
// of this library on one device) race on the array: serialise them.
static constexpr size_t kStaticScratchHalf = (size_t)16 << 20;            // one ping-pong buffer
static constexpr size_t kStaticScratchBytes = 2 * kStaticScratchHalf;     // 32 MiB per GPU
static __device__ __align__(256) unsigned char g_static_scratch[kStaticScratchBytes];
static std::mutex g_static_mu;
static unsigned char* g_static_ptr[64] = {};   // per device id: the array's device address, resolved once

int Engine::static_max_images() const {
  const size_t per = (act_bytes(256) - act_bytes(128)) / 128;  // bytes per image (act_bytes: whole 128-image tiles + slack)
  return (int)((kStaticScratchHalf / per - (net_.is_cnv ? 20 : 0)) / 128 * 128);   // act_bytes: whole 128-image tiles + its slack
}

int Engine::infer_device_static(int slot, const void* d_in, int n, void* d_out, int* d_classes, int ncls, cudaStream_t stream) {
  if (!loaded_) { err_ = "load_parameters has not been called"; return -1; }
  if (slot < 0 || slot >= (int)dev_.size()) { err_ = "bad device slot"; return -1; }
  DeviceState& ds = *dev_[slot];
  if (!ds.ready) {  // the parameter upload allocates: it belongs to an ordinary call made beforehand, never to this entry point
    err_ = "bnnb_infer_device_static allocates nothing, so the device must hold the parameters already: make one ordinary call "
           "(bnnb_infer_device / bnnb_classify_*) on this device after bnnb_load_parameters first";
    return -1;
  }
  if (profile_) { err_ = "bnnb_infer_device_static: per-layer profiling creates CUDA events; disable it (bnnb_profile_enable(c, 0))"; return -1; }
  if (ds.id < 0 || ds.id >= 64) { err_ = "bnnb_infer_device_static: device id out of range"; return -1; }
  if (n <= 0) { last_launches_ = 0; return 0; }
  const int max_n = static_max_images();
  if (n > max_n || act_bytes(n) > kStaticScratchHalf) {
    err_ = "bnnb_infer_device_static: " + std::to_string(n) + " images exceed the static scratch (" + std::to_string(kStaticScratchBytes >> 20) +
           " MiB per GPU = at most " + std::to_string(max_n) + " images per call for " + net_.name + "); split the batch or use bnnb_infer_device";
    return -1;
  }
  CK(cudaSetDevice(ds.id));
  unsigned char* base;
  {
    std::lock_guard<std::mutex> lk(g_static_mu);
    if (!g_static_ptr[ds.id]) {  // set-up, once per device, with that device current
      void* p = nullptr;
      CK(cudaGetSymbolAddress(&p, g_static_scratch));
      if (!p) { err_ = "cudaGetSymbolAddress returned a null address for the static scratch"; return -1; }
      g_static_ptr[ds.id] = (unsigned char*)p;
    }
    base = g_static_ptr[ds.id];
  }
  uint32_t* bufs[2] = {reinterpret_cast<uint32_t*>(base), reinterpret_cast<uint32_t*>(base + kStaticScratchHalf)};
  cudaStream_t s = stream ? stream : ds.compute;
  const int r = run_chunk(ds, d_in, n, d_out, d_classes, ncls, s, nullptr, bufs, err_);
  if (r < 0) return -1;
  last_launches_ = r;
  CK(cudaGetLastError());
  return 0;
}

int Engine::prep_device(int slot, const void* d_raw, int n, void* d_in, cudaStream_t stream) {
  if (ensure_devices() != 0) return -1;
  if (slot < 0 || slot >= (int)dev_.size()) { err_ = "bad device slot"; return -1; }
  DeviceState& ds = *dev_[slot];
  CK(cudaSetDevice(ds.id));
  if (!ds.ready && loaded_ && upload(ds, err_) != 0) return -1;
  cudaStream_t s = stream ? stream : ds.compute;
  if (!s) { err_ = "load_parameters has not been called"; return -1; }
  if (net_.is_cnv) launch_prep_cifar((const uint8_t*)d_raw, (int8_t*)d_in, n, s);
  else launch_prep_mnist((const uint8_t*)d_raw, (uint32_t*)d_in, n, s);
  CK(cudaGetLastError());
  return 0;
}

int Engine::sync(int slot) {
  if (slot < 0 || slot >= (int)dev_.size()) { err_ = "bad device slot"; return -1; }
  DeviceState& ds = *dev_[slot];
  if (!ds.ready) return 0;
  CK(cudaSetDevice(ds.id));
  // the device-resident path may run on the caller's stream: wait for the whole device, then read the status flag
  CK(cudaDeviceSynchronize());
  CK(cudaGetLastError());
  return check_status(ds, err_);
}

// One image through a captured CUDA graph.  Returns 0 (done), -1 (error) or 1 (this call must take the ordinary path: the first
// one-image call of a configuration runs there, which also loads every kernel and sizes the lane's buffers - nothing a capture may
// do - and a device that refuses the capture stays there for good).  BNN_B200_NO_GRAPH=1 switches the path off.
int Engine::run_single(DeviceState& ds, const uint8_t* rec, size_t raw_bytes, int ncls, int* classes, int16_t* scores, uint64_t* words,
                       float* dev_ms, std::string& err) {
  static const bool off = std::getenv("BNN_B200_NO_GRAPH") != nullptr || std::getenv("BNN_B200_TIMELINE") != nullptr;
  DeviceState::SingleGraph& sg = ds.single;
  if (off || sg.disabled) return 1;
  Lane& ln = ds.lane[0];
  if (ensure_lane(ds, 0, 1, err) != 0 || ensure_buffers(ds, 0, 1, err) != 0) return -1;
  const std::vector<uint64_t> key = {(uint64_t)(uintptr_t)ln.d_raw, (uint64_t)(uintptr_t)ln.d_in, (uint64_t)(uintptr_t)ln.act[0], (uint64_t)(uintptr_t)ln.act[1],
                                     (uint64_t)(uintptr_t)ln.d_out, (uint64_t)(uintptr_t)ln.d_cls, (uint64_t)(uintptr_t)ds.arena, (uint64_t)ncls,
                                     (uint64_t)packed_, (uint64_t)core_, (uint64_t)fuse01_, (uint64_t)chain_};
  if (key != sg.key) {  // a new configuration: its first call runs the ordinary path
    if (sg.exec) { cudaGraphExecDestroy(sg.exec); sg.exec = nullptr; }
    sg.key = key;
    sg.warm = false;
  }
  if (!sg.warm) { sg.warm = true; return 1; }
  if (!sg.h_in) {
    if (cudaHostAlloc((void**)&sg.h_in, 4096, cudaHostAllocPortable) != cudaSuccess || cudaHostAlloc((void**)&sg.h_out, 256, cudaHostAllocPortable) != cudaSuccess) {
      cudaGetLastError();
      sg.disabled = true;
      return 1;
    }
  }
  if (!sg.exec) {
    if (cudaStreamBeginCapture(ln.s, cudaStreamCaptureModeThreadLocal) != cudaSuccess) { cudaGetLastError(); sg.disabled = true; return 1; }
    cudaMemcpyAsync(packed_ ? ln.d_in : (void*)ln.d_raw, sg.h_in, raw_bytes, cudaMemcpyHostToDevice, ln.s);
    if (packed_) {}
    else if (net_.is_cnv) launch_prep_cifar(ln.d_raw, (int8_t*)ln.d_in, 1, ln.s);
    else launch_prep_mnist(ln.d_raw, (uint32_t*)ln.d_in, 1, ln.s);
    std::string cerr;
    static const bool pdl = std::getenv("BNN_B200_NO_PDL") == nullptr;
    tl_pdl = pdl;  // the layer kernels start together and wait for each other's output (kernels.cuh)
    const int r = run_chunk(ds, ln.d_in, 1, ln.d_out, ln.d_cls, ncls, ln.s, nullptr, ln.act, cerr);
    tl_pdl = false;
    cudaMemcpyAsync(sg.h_out, ln.d_out, (size_t)net_.out_bytes, cudaMemcpyDeviceToHost, ln.s);
    cudaMemcpyAsync(sg.h_out + 128, ln.d_cls, 4, cudaMemcpyDeviceToHost, ln.s);
    cudaGraph_t graph = nullptr;
    const cudaError_t ce = cudaStreamEndCapture(ln.s, &graph);
    if (r < 0 || ce != cudaSuccess || !graph || cudaGraphInstantiate(&sg.exec, graph, 0) != cudaSuccess) {
      if (graph) cudaGraphDestroy(graph);
      cudaGetLastError();
      sg.exec = nullptr;
      sg.disabled = true;
      return 1;
    }
    cudaGraphDestroy(graph);
  }
  while (ds.tev.size() < 2) {
    cudaEvent_t e;
    CKE(err, cudaEventCreate(&e));
    ds.tev.push_back(e);
  }
  std::memcpy(sg.h_in, rec, raw_bytes);
  CKE(err, cudaEventRecord(ds.tev[0], ln.s));
  CKE(err, cudaGraphLaunch(sg.exec, ln.s));
  CKE(err, cudaEventRecord(ds.tev[1], ln.s));
  CKE(err, cudaStreamSynchronize(ln.s));
  if (check_status(ds, err) != 0) return -1;
  if (net_.is_cnv) { if (scores) std::memcpy(scores, sg.h_out, 128); }
  else if (words) std::memcpy(words, sg.h_out, 8);
  if (classes) std::memcpy(classes, sg.h_out + 128, 4);
  CKE(err, cudaEventElapsedTime(dev_ms, ds.tev[0], ds.tev[1]));
  return 0;
}

// One device's shard [lo, hi): chunks go round-robin through the lanes (with fault epochs: run_shard_faulted).
int Engine::run_shard(int g, int lo, int hi, const uint8_t* input, int ncls, int* classes, int16_t* scores, uint64_t* words,
                      const Segments* seg, float* dev_ms, float* last_ms, int* last_imgs, std::string& err) {
  DeviceState& ds = *dev_[g];
  CKE(err, cudaSetDevice(ds.id));
  if (!ds.ready && upload(ds, err) != 0) return -1;
  if (seg) { *dev_ms = 0; return run_shard_faulted(g, lo, hi, input, ncls, classes, words, *seg, last_ms, last_imgs, err); }
  const size_t raw_bytes = packed_ ? (size_t)net_.in_bytes : (net_.is_cnv ? 3073 : 784);  // packed_: the caller hands over BlackBoxJam's own input
  const int chunk = chunk_images(true);
  if (hi - lo == 1 && !profile_) {
    const int rs = run_single(ds, input + (size_t)lo * raw_bytes, raw_bytes, ncls, classes ? classes + lo : nullptr, scores ? scores + (size_t)lo * 64 : nullptr,
                              words ? words + lo : nullptr, dev_ms, err);
    if (rs <= 0) { *last_ms = *dev_ms; *last_imgs = 1; return rs; }  // 1: not on the graph path (first call, capture refused)
  }
  for (auto& ln : ds.lane) { ln.used = false; ln.pend_off = -1; }
  auto flush_lane = [&](Lane& ln) -> int {  // results of the lane's chunk in flight -> host
    if (ln.pend_off < 0) return 0;
    const int off = ln.pend_off, c = ln.pend_n;
    if (net_.is_cnv) {
      if (scores) CKE(err, cudaMemcpyAsync(scores + (size_t)off * 64, ln.d_out, (size_t)c * 128, cudaMemcpyDeviceToHost, ln.s));
      if (classes) CKE(err, cudaMemcpyAsync(classes + off, ln.d_cls, (size_t)c * 4, cudaMemcpyDeviceToHost, ln.s));
    } else {  // LFC: the classes are decoded on the device (4 B per image back instead of 8, no host pass over the words)
      if (words) CKE(err, cudaMemcpyAsync(words + off, ln.d_out, (size_t)c * 8, cudaMemcpyDeviceToHost, ln.s));
      if (classes) CKE(err, cudaMemcpyAsync(classes + off, ln.d_cls, (size_t)c * 4, cudaMemcpyDeviceToHost, ln.s));
    }
    ln.pend_off = -1;
    return 0;
  };
  // timing events around the layer kernels of every chunk
  size_t tused = 0;
  auto timing_pair = [&](cudaEvent_t* a, cudaEvent_t* b) -> int {
    while (ds.tev.size() < tused + 2) {
      cudaEvent_t e;
      CKE(err, cudaEventCreate(&e));
      ds.tev.push_back(e);
    }
    *a = ds.tev[tused]; *b = ds.tev[tused + 1];
    tused += 2;
    return 0;
  };
  struct Timed { cudaEvent_t a, b; int n; };
  std::vector<Timed> timed;
  // BNN_B200_TIMELINE=1: per chunk, when its copy started and when its kernels started / ended, on stderr (a measuring aid)
  static const bool timeline = std::getenv("BNN_B200_TIMELINE") != nullptr;
  std::vector<cudaEvent_t> copy_start;
  const auto wall0 = std::chrono::steady_clock::now();
  int cur = lo, ci = 0;
  {
    while (cur < hi) {
      const int rem = hi - cur;
      // The copy engine is the bottleneck of this path (3 073 B per image over PCIe): only the compute of the LAST chunk is not
      // hidden behind a copy, so the tail of a shard is cut into halves (full chunks elsewhere keep the kernels efficient).
      int c = std::min(chunk, rem);
      if (rem <= chunk + chunk / 2 && rem > 4096) c = std::min(chunk, (rem / 2 + 1023) / 1024 * 1024);
      const int k = ci % kLanes;
      Lane& ln = ds.lane[k];
      if (flush_lane(ln) != 0) return -1;
      if (ensure_lane(ds, k, c, err) != 0 || ensure_buffers(ds, k, c, err) != 0) return -1;
      if (timeline) {
        cudaEvent_t e;
        CKE(err, cudaEventCreate(&e));
        CKE(err, cudaEventRecord(e, ln.s));
        copy_start.push_back(e);
      }
      CKE(err, cudaMemcpyAsync(packed_ ? ln.d_in : (void*)ln.d_raw, input + (size_t)cur * raw_bytes, (size_t)c * raw_bytes, cudaMemcpyHostToDevice, ln.s));
      if (packed_) {}
      else if (net_.is_cnv) launch_prep_cifar(ln.d_raw, (int8_t*)ln.d_in, c, ln.s);
      else launch_prep_mnist(ln.d_raw, (uint32_t*)ln.d_in, c, ln.s);
      const bool time_it = true;
      Timed t{nullptr, nullptr, c};
      if (time_it) { if (timing_pair(&t.a, &t.b) != 0) return -1; CKE(err, cudaEventRecord(t.a, ln.s)); }
      if (run_chunk(ds, ln.d_in, c, ln.d_out, classes ? ln.d_cls : nullptr, ncls, ln.s, nullptr, ln.act, err) < 0) return -1;
      if (time_it) { CKE(err, cudaEventRecord(t.b, ln.s)); timed.push_back(t); }
      ln.used = true;
      ln.pend_off = cur; ln.pend_n = c;
      cur += c;
      ci++;
    }
  }
  for (auto& ln : ds.lane) if (flush_lane(ln) != 0) return -1;
  for (auto& ln : ds.lane) CKE(err, cudaStreamSynchronize(ln.s));
  CKE(err, cudaStreamSynchronize(ds.compute));
  CKE(err, cudaGetLastError());
  if (check_status(ds, err) != 0) return -1;
  if (timeline) {
    const double wall = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - wall0).count();
    std::fprintf(stderr, "[timeline] device %d images %d wall %.3f ms\n", ds.id, hi - lo, wall);
    for (size_t i = 0; i < timed.size(); i++) {
      float c0 = 0, a = 0, b = 0;
      cudaEventElapsedTime(&c0, copy_start[0], copy_start[i]);
      cudaEventElapsedTime(&a, copy_start[0], timed[i].a);
      cudaEventElapsedTime(&b, copy_start[0], timed[i].b);
      std::fprintf(stderr, "[timeline]   chunk %2zu n %5d copy from %.3f kernels %.3f .. %.3f ms\n", i, timed[i].n, c0, a, b);
    }
    for (auto e : copy_start) cudaEventDestroy(e);
  }
  // Device time of the call = the time during which at least one chunk's layer kernels were in flight: the union of the chunks'
  // [start, end] intervals (lanes overlap; a plain sum would count the overlap several times, and the gaps in which every lane waits
  // for its copy are not kernel time).
  std::vector<std::pair<float, float>> iv;
  for (size_t i = 0; i < timed.size(); i++) {
    float a = 0, b = 0;
    if (i > 0) CKE(err, cudaEventElapsedTime(&a, timed[0].a, timed[i].a));
    CKE(err, cudaEventElapsedTime(&b, timed[0].a, timed[i].b));
    iv.emplace_back(a, b);
    if (i + 1 == timed.size()) { *last_ms = b - a; *last_imgs = timed[i].n; }
  }
  std::sort(iv.begin(), iv.end());
  float total = 0, end = -1e30f;
  for (auto& x : iv) {
    if (x.first > end) { total += x.second - x.first; end = x.second; }
    else if (x.second > end) { total += x.second - end; end = x.second; }
  }
  *dev_ms = total;
  return 0;
}

// A shard with fault epochs.  The lanes run exactly as without faults - each its own in-order stream, copies and kernels of
// consecutive chunks overlapping - because every lane owns a parameter set: lane 0 the device's set 0, lane k a clone taken at the
// start of the call.  A lane applies EVERY patch to its own set, in time order, just before its first kernel run that starts at or
// after the patch's time (run_chunk_faulted); it processes its chunks in increasing image order, so its set is always the
// parameter state of the images it is working on, and no lane ever has to wait for another.  Chunks are not cut at fault times:
// only the kernel of the faulted layer is split.  At the end lane 0 applies what is left, so set 0 carries every fault onwards.
int Engine::run_shard_faulted(int g, int lo, int hi, const uint8_t* input, int ncls, int* classes, uint64_t* words, const Segments& seg,
                              float* last_ms, int* last_imgs, std::string& err) {
  DeviceState& ds = *dev_[g];
  const size_t raw_bytes = packed_ ? (size_t)net_.in_bytes : (net_.is_cnv ? 3073 : 784);  // packed_: the caller hands over BlackBoxJam's own input
  const int chunk = chunk_images(true);
  for (auto& ln : ds.lane) { ln.used = false; ln.pend_off = -1; }
  for (auto& ln : ds.lane) CKE(err, cudaStreamSynchronize(ln.s));
  // patches by kernel (group), time order kept
  FaultPlan fp;
  {
    int fam[9], ofmt[9];
    bool fused01;
    plan_formats(ds, false, false, fam, ofmt, &fused01, true);
    for (int l = 0; l < 9; l++) fp.group_of_layer[l] = (fused01 && l == 0) ? 1 : l;
    std::vector<Patch> grouped;
    grouped.reserve(seg.patches.size());
    for (int grp = 0; grp < 9; grp++) {
      fp.goff[grp] = grouped.size();
      for (size_t i = 0; i < seg.patches.size(); i++)
        if (fp.group_of_layer[seg.patches[i].layer] == grp) { grouped.push_back(seg.patches[i]); fp.gtime[grp].push_back(seg.ptime[i]); }
    }
    if (!grouped.empty()) {
      if (ensure_patch_cap(ds, grouped.size(), err) != 0) return -1;
      CKE(err, cudaMemcpyAsync(ds.d_patches, grouped.data(), grouped.size() * sizeof(DevPatch), cudaMemcpyHostToDevice, ds.compute));
    }
    CKE(err, cudaStreamSynchronize(ds.compute));  // `grouped` dies with this block; set 0 is quiescent before it is cloned
  }
  // The clones of set 0 for lanes 1.. are taken HERE, before any lane has enqueued a patch: taken lazily at a lane's first chunk they
  // raced with lane 0 already patching set 0 (a torn copy - weights of a word from before a patch, its threshold offset from after -
  // makes the lane apply that patch's offset twice: a few wrong images per million, seen on the second device of a two-device run).
  const bool multi_chunk = hi - lo > 4096;  // the loop below cuts anything larger into at least two chunks (two lanes)
  if (!seg.patches.empty() && multi_chunk) {
    for (int k = 1; k < kLanes; k++) {
      DeviceState::ParamSet& c = ds.clone[k - 1];
      if (!c.arena) {
        c.arena = (unsigned char*)ds.alloc(ds.arena_bytes, true);
        if (!c.arena) { err = "cudaMalloc failed (parameter clone, device " + std::to_string(ds.id) + ")"; return -1; }
        carve_params(net_, c.arena, c.L, &c.d_layers);
      }
      CKE(err, cudaMemcpyAsync(c.arena, ds.arena, ds.arena_bytes, cudaMemcpyDeviceToDevice, ds.compute));
      CKE(err, cudaMemcpyAsync(c.d_layers, c.L, sizeof(DevLayer) * 9, cudaMemcpyHostToDevice, ds.compute));  // the table points into the clone
    }
    CKE(err, cudaStreamSynchronize(ds.compute));
  }
  auto param_set = [&](int k, const DevLayer** PL, DevLayer** tbl) -> int {  // lane k's parameter set
    if (k == 0 || seg.patches.empty() || !multi_chunk) { *PL = ds.L; *tbl = ds.d_layers; return 0; }
    DeviceState::ParamSet& c = ds.clone[k - 1];
    *PL = c.L; *tbl = c.d_layers;
    return 0;
  };
  auto flush_lane = [&](Lane& ln) -> int {  // results of the lane's chunk in flight -> host
    if (ln.pend_off < 0) return 0;
    const int off = ln.pend_off, c = ln.pend_n;
    CKE(err, cudaMemcpyAsync(classes + off, ln.d_cls, (size_t)c * 4, cudaMemcpyDeviceToHost, ln.s));
    (void)words;
    ln.pend_off = -1;
    return 0;
  };
  while (ds.tev.size() < 2) {
    cudaEvent_t e;
    CKE(err, cudaEventCreate(&e));
    ds.tev.push_back(e);
  }
  int cur = lo, ci = 0;
  bool timed = false;
  while (cur < hi) {
    const int rem = hi - cur;
    int c = std::min(chunk, rem);
    if (rem <= chunk + chunk / 2 && rem > 4096) c = std::min(chunk, (rem / 2 + 1023) / 1024 * 1024);
    const int k = ci % kLanes;
    Lane& ln = ds.lane[k];
    if (flush_lane(ln) != 0) return -1;
    if (ensure_lane(ds, k, c, err) != 0 || ensure_buffers(ds, k, c, err) != 0) return -1;
    const DevLayer* PL;
    DevLayer* tbl;
    if (param_set(k, &PL, &tbl) != 0) return -1;
    CKE(err, cudaMemcpyAsync(packed_ ? ln.d_in : (void*)ln.d_raw, input + (size_t)cur * raw_bytes, (size_t)c * raw_bytes, cudaMemcpyHostToDevice, ln.s));
    if (packed_) {}
    else if (net_.is_cnv) launch_prep_cifar(ln.d_raw, (int8_t*)ln.d_in, c, ln.s);
    else launch_prep_mnist(ln.d_raw, (uint32_t*)ln.d_in, c, ln.s);
    const bool time_it = cur + c == hi;
    if (time_it) CKE(err, cudaEventRecord(ds.tev[0], ln.s));
    if (run_chunk_faulted(ds, PL, tbl, k, ln.d_in, c, cur, ln.d_out, ln.d_cls, ncls, ln.s, ln.act, fp, err) != 0) return -1;
    if (time_it) { CKE(err, cudaEventRecord(ds.tev[1], ln.s)); timed = true; *last_imgs = c; }
    ln.used = true;
    ln.pend_off = cur; ln.pend_n = c;
    cur += c;
    ci++;
  }
  // faults scheduled at or after the last run of lane 0 (or of this shard) still persist: set 0 takes the rest
  for (int grp = 0; grp < 9; grp++)
    if (fp.gcur[0][grp] < fp.gtime[grp].size())
      launch_patch_apply(ds.d_layers, net_.is_cnv, ds.d_patches + fp.goff[grp] + fp.gcur[0][grp], (int)(fp.gtime[grp].size() - fp.gcur[0][grp]), ds.lane[0].s);
  for (auto& ln : ds.lane) if (flush_lane(ln) != 0) return -1;
  for (auto& ln : ds.lane) CKE(err, cudaStreamSynchronize(ln.s));
  CKE(err, cudaStreamSynchronize(ds.compute));
  CKE(err, cudaGetLastError());
  if (check_status(ds, err) != 0) return -1;
  if (timed) CKE(err, cudaEventElapsedTime(last_ms, ds.tev[0], ds.tev[1]));
  return 0;
}

int Engine::run_sharded(const uint8_t* input, int n, int ncls, int* classes, int16_t* scores, uint64_t* words, const Segments* seg,
                        float* dev_ms_max, float* last_ms, int* last_imgs) {
  const int G = (int)dev_.size();
  // contiguous shards [g*n/Gs, (g+1)*n/Gs): SURVEY 8(e); a small batch stays on as few devices as give each kMinShard images
  const int Gs = std::max(1, std::min(G, n / kMinShard));
  std::vector<std::string> errs(G);
  std::vector<int> rc(G, 0);
  std::vector<float> ms(G, 0.f), lms(G, 0.f);
  std::vector<int> limg(G, 0);
  auto work = [&](int g) {
    const int lo = g < Gs ? (int)((long long)g * n / Gs) : n, hi = g < Gs ? (int)((long long)(g + 1) * n / Gs) : n;
    rc[g] = run_shard(g, lo, hi, input, ncls, classes, scores, words, seg, &ms[g], &lms[g], &limg[g], errs[g]);
  };
  // devices beyond Gs hold no images; with faults they still apply the patch list if they already hold parameters (a device that
  // is not in use yet takes the mirror, faults included, when it is first used)
  std::vector<int> active;
  for (int g = 0; g < G; g++)
    if (g < Gs || (seg && !seg->patches.empty() && dev_[g]->ready)) active.push_back(g);
  if (active.size() == 1) {
    work(active[0]);
  } else {  // one feeder thread per device: no device waits for the host's attention to another
    std::vector<std::thread> th;
    for (int g : active) th.emplace_back(work, g);
    for (auto& t : th) t.join();
  }
  for (int g : active)
    if (rc[g] != 0) { err_ = errs[g]; return -1; }
  *dev_ms_max = 0;
  for (int g : active) *dev_ms_max = std::max(*dev_ms_max, ms[g]);
  *last_ms = lms[Gs - 1];
  *last_imgs = limg[Gs - 1];
  return 0;
}

int Engine::infer_host(const uint8_t* input, int n, int ncls, int* classes, int16_t* scores, uint64_t* words, float* usec) {
  if (!loaded_) { err_ = "load_parameters has not been called"; return -1; }
  if (ensure_devices() != 0) return -1;
  if (usec) *usec = 0;
  if (n <= 0) return 0;
  float worst = 0, last_ms = 0;
  int last_imgs = 0;
  if (run_sharded(input, n, ncls, classes, scores, words, nullptr, &worst, &last_ms, &last_imgs) != 0) return -1;
  if (usec) *usec = worst * 1000.0f / (float)n;
  return 0;
}

// The same call on inputs the caller has already prepared as BlackBoxJam reads them (LFC: 104 bytes = 832 bits per image, what
// binarizeAndPack leaves, foldedmv-offload.cpp:82-98; CNV: 3 072 int8 HWC bytes, quantiseAndPack<8,1>, foldedmv-offload.h:129-144):
// no quantise / pack kernel, and for LFC 7.5x fewer bytes over PCIe.
int Engine::infer_host_packed(const uint8_t* input, int n, int ncls, int* classes, int16_t* scores, uint64_t* words, float* usec) {
  packed_ = true;
  const int rc = infer_host(input, n, ncls, classes, scores, words, usec);
  packed_ = false;
  return rc;
}

int Engine::infer_host_with_faults(const uint8_t* input, int n, int ncls, const std::vector<Fault>& faults_in, int* classes,
                                   float* usec) {
  if (!loaded_) { err_ = "load_parameters has not been called"; return -1; }
  if (ensure_devices() != 0) return -1;
  if (usec) *usec = 0;
  // Every device that will run a shard must hold the PRE-fault parameters before the mirror moves on: it replays the patch list
  // in stream order between its segments.
  const int G = (int)dev_.size(), Gs = std::max(1, std::min(G, n / kMinShard));
  if (n > 0)
    for (int g = 0; g < Gs; g++)
      if (!dev_[g]->ready && upload(*dev_[g], err_) != 0) return -1;
  // host: run the whole read-modify-write chain in fault order (it does not depend on inference results)
  std::vector<Fault> faults = faults_in;
  std::stable_sort(faults.begin(), faults.end(), [](const Fault& a, const Fault& b) { return a.time < b.time; });
  Segments seg;  // one patch per effective fault, in order
  for (const Fault& f : faults) {
    Patch p;
    if (pm_->inject(f.target, f.layer, f.bit, f.word_size, &p)) { seg.patches.push_back(p); seg.ptime.push_back(f.time); }
  }
  if (n <= 0) return apply_patches(seg.patches);
  float worst = 0, last_ms = 0;
  int last_imgs = 0;
  if (run_sharded(input, n, ncls, classes, nullptr, nullptr, &seg, &worst, &last_ms, &last_imgs) != 0) return -1;
  // the reference reports the time of its LAST single-image call: device time per image of the last segment
  if (usec) *usec = last_imgs > 0 ? last_ms * 1000.0f / (float)last_imgs : 0.f;
  return 0;
}

int Engine::profile_collect(float* ms, int* launches) {
  for (int i = 0; i < 10; i++) { ms[i] = 0; launches[i] = 0; }
  std::lock_guard<std::mutex> lk(prof_mu_);
  for (auto& e : prof_) {
    cudaSetDevice(e.dev);
    cudaEventSynchronize(e.b);
    float t = 0;
    if (cudaEventElapsedTime(&t, e.a, e.b) == cudaSuccess) { ms[e.layer] += t; launches[e.layer]++; }
    cudaEventDestroy(e.a);
    cudaEventDestroy(e.b);
  }
  prof_.clear();
  return 0;
}

size_t Engine::tap_acc_size() const {
  size_t s = 0;
  for (int i = 0; i < net_.nlayers; i++) s += (size_t)net_.L[i].ofm_dim * net_.L[i].ofm_dim * net_.L[i].N;
  return s;
}
size_t Engine::tap_act_size() const {
  size_t s = 0;
  for (int i = 0; i < net_.nlayers; i++) s += (size_t)out_dim(net_.L[i]) * out_dim(net_.L[i]) * net_.L[i].N;
  return s;
}

// #(-1) weights of every logical row of a 1-bit-weight layer, from the host mirror in the reference's folded coordinates
// (row = nf*PE + pe, tile = nf*SF + sf: mvau.hpp:115,143-152): what turns an MMA sum over {0,1} activations into the reference's
// XNOR-popcount accumulator (interpret.hpp:57-69).
std::vector<int> Engine::row_neg_counts(int layer) const {
  const LayerDesc& l = net_.L[layer];
  std::vector<int> neg(l.N, 0);
  const int SF = l.K / l.simd;
  const uint64_t mask = l.simd >= 64 ? ~0ull : ((1ull << l.simd) - 1);
  for (int n = 0; n < l.N; n++) {
    const int pe = n % l.pe, nf = n / l.pe;
    for (int sf = 0; sf < SF; sf++) neg[n] += l.simd - __builtin_popcountll(pm_->weights(layer)[(size_t)pe * l.wmem + nf * SF + sf] & mask);
  }
  return neg;
}

// e2m1 code of one activation -> value.  fold: the fused kernel's layer-0 epilogue writes a set decision as 0xF (-6.0, see cnv01_tc.cu)
static inline int8_t e2m1_act(uint32_t code, bool fold) {
  if (fold) return code ? 1 : 0;
  return code == 0x2 ? 1 : (code == 0xA ? -1 : 0);
}

int Engine::tap_host(const uint8_t* input, int n, int32_t* acc, int8_t* act) {
  if (!loaded_) { err_ = "load_parameters has not been called"; return -1; }
  if (ensure_devices() != 0) return -1;
  DeviceState& ds = *dev_[0];
  CK(cudaSetDevice(ds.id));
  if (!ds.ready && upload(ds, err_) != 0) return -1;
  cudaStream_t s = ds.compute;
  const size_t raw_bytes = net_.is_cnv ? 3073 : 784;
  const size_t sa = tap_acc_size(), st = tap_act_size();
  const int step = 16, NL = net_.nlayers;
  struct Bufs {  // freed on every path out
    std::vector<void*> v;
    ~Bufs() { for (void* p : v) cudaFree(p); }
    void* get(size_t bytes) { void* p = nullptr; if (cudaMalloc(&p, bytes ? bytes : 16) != cudaSuccess) return nullptr; v.push_back(p); return p; }
  } B;
  uint8_t* d_raw = (uint8_t*)B.get(step * raw_bytes);
  void* d_in = B.get((size_t)step * net_.in_bytes);
  int32_t* d_acc = (int32_t*)B.get((size_t)step * sa * 4);
  uint4* d_act0 = (uint4*)B.get((size_t)step * 900 * 4 * 16);
  uint32_t* bufs[9];
  std::vector<size_t> bytes(NL);
  bool ok = d_raw && d_in && d_acc && d_act0;
  for (int i = 0; i < NL; i++) {
    const LayerDesc& l = net_.L[i];
    const size_t px = (size_t)out_dim(l) * out_dim(l);
    bytes[i] = 128 * std::max(px * out_words_per_pixel(l) * 4, px * l.N / 2);  // any format, one 128-image tile
    bufs[i] = (uint32_t*)B.get(bytes[i]);
    ok = ok && bufs[i];
  }
  if (!ok) { err_ = "cudaMalloc failed (tap buffers)"; return -1; }
  std::vector<std::vector<int>> neg(NL);
  for (int i = 0; i < NL; i++)
    if (net_.L[i].in_mode == IN_XNOR) neg[i] = row_neg_counts(i);
  std::vector<int32_t> hacc((size_t)step * sa);
  std::vector<uint8_t> hbuf;
  std::vector<uint4> hact0((size_t)step * 900 * 4);
  for (int off = 0; off < n; off += step) {
    const int c = std::min(step, n - off);
    CK(cudaMemcpyAsync(d_raw, input + (size_t)off * raw_bytes, c * raw_bytes, cudaMemcpyHostToDevice, s));
    if (net_.is_cnv) launch_prep_cifar(d_raw, (int8_t*)d_in, c, s);
    else launch_prep_mnist(d_raw, (uint32_t*)d_in, c, s);
    for (int i = 0; i < NL; i++) CK(cudaMemsetAsync(bufs[i], 0, bytes[i], s));
    TapCtx tap;
    tap.acc = d_acc;
    tap.act0 = d_act0;
    tap.tensor = core_ >= 1;
    if (run_chunk(ds, d_in, c, bufs[NL - 1], nullptr, 10, s, &tap, bufs, err_) < 0) return -1;
    CK(cudaMemcpyAsync(hacc.data(), d_acc, (size_t)c * sa * 4, cudaMemcpyDeviceToHost, s));
    if (tap.fused01) CK(cudaMemcpyAsync(hact0.data(), d_act0, (size_t)c * 900 * 4 * 16, cudaMemcpyDeviceToHost, s));
    CK(cudaStreamSynchronize(s));
    CK(cudaGetLastError());
    if (check_status(ds, err_) != 0) return -1;
    // accumulators: device buffer is layer-major over the chunk, the API image-major.  A tensor-core kernel reports its MMA sum S:
    // CNV layer 0: S = sum w*q, the reference accumulates 2q in ap_fixed<24,16> (top.cpp:61,214) -> 2S; 1-bit layers: S runs over
    // {0,1} activations, the reference counts XNOR matches = S + #(-1 weights of the row); ternary / +-1 inputs: S itself.
    size_t lo = 0, io = 0;
    for (int i = 0; i < NL; i++) {
      const LayerDesc& l = net_.L[i];
      const size_t rows = (size_t)l.ofm_dim * l.ofm_dim, per = rows * l.N;
      for (int k = 0; k < c; k++) {
        int32_t* dst = acc + (size_t)(off + k) * sa + io;
        const int32_t* src = hacc.data() + lo + (size_t)k * per;
        if (!tap.raw[i]) { std::memcpy(dst, src, per * 4); continue; }
        for (size_t r = 0; r < rows; r++)
          for (int ch = 0; ch < l.N; ch++) {
            const int32_t S = src[r * l.N + ch];
            dst[r * l.N + ch] = l.in_mode == IN_FIX8 ? 2 * S : (l.in_mode == IN_XNOR ? S + neg[i][ch] : S);
          }
      }
      lo += (size_t)c * per;
      io += per;
    }
    // layer outputs, decoded from the format each kernel wrote
    size_t ao = 0;
    for (int i = 0; i < NL; i++) {
      const LayerDesc& l = net_.L[i];
      const int px = out_dim(l) * out_dim(l), nw = l.N / 32, wpp = out_words_per_pixel(l);
      const bool from_act0 = tap.fused01 && i == 0;
      if (!from_act0) {
        hbuf.resize(bytes[i]);
        CK(cudaMemcpy(hbuf.data(), bufs[i], bytes[i], cudaMemcpyDeviceToHost));
      }
      const uint32_t* hw = reinterpret_cast<const uint32_t*>(hbuf.data());
      const uint4* h4 = reinterpret_cast<const uint4*>(hbuf.data());
      for (int k = 0; k < c; k++)
        for (int p = 0; p < px; p++)
          for (int ch = 0; ch < l.N; ch++) {
            int8_t v;
            if (from_act0) {
              const uint4* row = hact0.data() + ((size_t)k * 900 + p) * 4;
              if (core_ == 2) {  // e2m1 codes, 32 channels per uint4 (entries 0 and 2), in the fused kernel's nibble order; 1-bit nets: code 0xF
                const bool fold = l.out_mode == OUT_BIT;
                const uint32_t* w = reinterpret_cast<const uint32_t*>(&row[2 * (ch >> 5)]);
                const int c32 = ch & 31, wi = c32 >> 3, e = c32 & 7;
                const int nib = (e & 3) * 2 + (e >> 2);  // nibble 2k = channel k, nibble 2k+1 = channel 4+k
                v = e2m1_act((w[wi] >> (4 * nib)) & 0xF, fold);
              } else {           // int8 operands, 16 channels per uint4
                v = reinterpret_cast<const int8_t*>(&row[ch >> 4])[ch & 15];
              }
            } else if (tap.ofmt[i] == ACT_F4P) {   // uint4 [img][plane][pixel]
              const uint32_t* w = reinterpret_cast<const uint32_t*>(&h4[((size_t)k * nw + (ch >> 5)) * px + p]);
              v = e2m1_act((w[(ch & 31) >> 3] >> (4 * (ch & 7))) & 0xF, false);
            } else if (tap.ofmt[i] == ACT_F4I) {   // uint4 [img/G][plane][pixel][img%G], G = the consumer's images per group
              const int G = conv_f4_group(net_.L[i + 1]);
              const uint32_t* w = reinterpret_cast<const uint32_t*>(&h4[(((size_t)(k / G) * nw + (ch >> 5)) * px + p) * G + k % G]);
              v = e2m1_act((w[(ch & 31) >> 3] >> (4 * (ch & 7))) & 0xF, false);
            } else if (tap.ofmt[i] == ACT_F4T) {   // uint4 [img/128][chunk = pixel*C/32 + plane][img%128]
              const uint32_t* w = reinterpret_cast<const uint32_t*>(&h4[((size_t)(k >> 7) * (px * nw) + (size_t)(p * nw + (ch >> 5))) * 128 + (k & 127)]);
              v = e2m1_act((w[(ch & 31) >> 3] >> (4 * (ch & 7))) & 0xF, false);
            } else {
              const uint32_t* w = hw + ((size_t)k * px + p) * wpp;
              if (l.out_mode == OUT_SCORE) v = (int8_t)reinterpret_cast<const int16_t*>(w)[ch];
              else if (l.out_mode == OUT_BIT) v = (w[ch >> 5] >> (ch & 31)) & 1;
              else { const int nz = (w[ch >> 5] >> (ch & 31)) & 1, ps = (w[nw + (ch >> 5)] >> (ch & 31)) & 1; v = nz ? (ps ? 1 : -1) : 0; }
            }
            act[(size_t)(off + k) * st + ao + (size_t)p * l.N + ch] = v;
          }
      ao += (size_t)px * l.N;
    }
  }
  return 0;
}

}  // namespace bnnb
