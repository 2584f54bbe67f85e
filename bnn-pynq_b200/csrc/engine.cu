#include "engine.hpp"
#include "f4_common.cuh"
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>

namespace bnnb {

#define CK(call)                                                                                   \
  do {                                                                                             \
    cudaError_t e_ = (call);                                                                       \
    if (e_ != cudaSuccess) {                                                                       \
      err_ = std::string(#call) + ": " + cudaGetErrorString(e_);                                   \
      return -1;                                                                                   \
    }                                                                                              \
  } while (0)

static constexpr int kLanes = 4;  // chunk pipelines per device: H2D of chunk i+1 overlaps compute of chunk i, and the host enqueues two chunks further ahead

struct Lane {
  cudaStream_t s = nullptr;
  uint8_t* d_raw = nullptr;   // chunk x raw input bytes (CIFAR records / MNIST pixels)
  void* d_in = nullptr;       // chunk x layer-0 input
  uint32_t* buf[2] = {nullptr, nullptr};  // ping-pong activations
  void* d_out = nullptr;      // chunk x out_bytes
  int* d_cls = nullptr;
  cudaEvent_t e0 = nullptr, e1 = nullptr;
};

struct DeviceState {
  int id = 0, num_sms = 148;
  Lane lane[kLanes];
  DevLayer L[9];
  DevFolded folded;
  uint64_t* wf[9] = {};
  int32_t* tf[9] = {};
  DevPatch* d_patches = nullptr;
  size_t patch_cap = 0;
  bool uploaded = false;
  int* d_status = nullptr;  // set by a kernel whose mbarrier wait timed out
  std::vector<void*> allocs;
};

static size_t act_words_per_image(const NetDesc& n, int first = 0) {  // largest inter-layer tensor written by layers >= first
  size_t m = 0;
  for (int i = first; i + 1 < n.nlayers; i++) {
    const LayerDesc& l = n.L[i];
    const size_t w = (size_t)out_dim(l) * out_dim(l) * out_words_per_pixel(l);
    m = std::max(m, w);
  }
  return m;
}

Engine::Engine(const NetDesc& net) : net_(net), pm_(new ParamMirror(net)) {
  if (const char* e = std::getenv("BNN_B200_CORE")) core_ = std::strcmp(e, "popc") == 0 ? 0 : (std::strcmp(e, "i8") == 0 ? 1 : 2);
  if (const char* e = std::getenv("BNN_B200_FUSE01")) fuse01_ = std::atoi(e) != 0;
}

int Engine::check_status(int slot) {
  DeviceState& ds = *dev_[slot];
  int st = 0;
  CK(cudaSetDevice(ds.id));
  CK(cudaMemcpy(&st, ds.d_status, 4, cudaMemcpyDeviceToHost));
  if (st != 0) { err_ = "a tensor-core kernel timed out waiting on an mbarrier"; return -1; }
  return 0;
}

Engine::~Engine() {
  for (auto& d : dev_) {
    cudaSetDevice(d->id);
    for (auto& ln : d->lane) {
      if (ln.s) cudaStreamSynchronize(ln.s);
      if (ln.e0) cudaEventDestroy(ln.e0);
      if (ln.e1) cudaEventDestroy(ln.e1);
      if (ln.s) cudaStreamDestroy(ln.s);
    }
    for (void* p : d->allocs) cudaFree(p);
  }
}

// Images per pass through the layer chain.  host = the pipelined host-buffer path, where a chunk is also the unit of H2D/compute
// overlap; otherwise the device-resident path, where larger chunks mean fewer launches and less tail quantisation (the path uses
// under 1% of HBM bandwidth, so nothing is gained by keeping the inter-layer tensors L2-resident there).
int Engine::chunk_images(bool host) const {
  if (const char* e = std::getenv(host ? "BNN_B200_HOST_CHUNK" : "BNN_B200_CHUNK")) { int v = std::atoi(e); if (v > 0) return v; }
  if (!net_.is_cnv) return host ? 32768 : 65536;
  const bool tern = net_.L[1].out_mode == OUT_TERN;
  if (core_ >= 1 && fuse01_) return host ? 8192 : (tern ? 32768 : 65536);
  // LOP3+POPC path: layer 0's output (7 200 / 14 400 B per image) exists in global memory; keep it inside the 126 MB L2
  return tern ? 2048 : 4096;
}

int Engine::set_devices(const int* ids, int n) {
  if (!dev_.empty()) { err_ = "set_devices must be called before the first use"; return -1; }
  dev_ids_.assign(ids, ids + n);
  return 0;
}

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
  // buffers are sized for the largest chunk either core selection may use (set_core can change between calls)
  // lane 0 also serves the device-resident path (larger chunks); per lane: images and bytes of the ping-pong activation buffers
  int lane_chunk[kLanes];
  size_t lane_act[kLanes];
  {
    const int c0 = core_;
    for (int k = 0; k < kLanes; k++) {
      core_ = 0;
      const int cp = chunk_images(true);
      const size_t ap = (size_t)cp * act_words_per_image(net_) * 4;
      core_ = 2;
      const int ct = std::max(chunk_images(true), k == 0 ? chunk_images(false) : 0);
      size_t at = (size_t)ct * act_words_per_image(net_, net_.is_cnv && fuse01_ ? 1 : 0) * 4;
      // between fp4 kernels of a 1-bit network activations travel as e2m1 operand planes (4 bits per value, 128-image tiles)
      for (int i = (net_.is_cnv ? 1 : 0); i + 1 < net_.nlayers; i++)
        if (net_.L[i].out_mode == OUT_BIT)
          at = std::max(at, (size_t)((ct + 127) / 128 * 128) * out_dim(net_.L[i]) * out_dim(net_.L[i]) * net_.L[i].N / 2);
      lane_chunk[k] = std::max(cp, ct);
      lane_act[k] = std::max(ap, at);
    }
    core_ = c0;
  }
  const size_t raw_bytes = net_.is_cnv ? 3073 : 784;
  for (int id : dev_ids_) {
    if (id < 0 || id >= count) { err_ = "bad device id"; return -1; }
    CK(cudaSetDevice(id));
    std::unique_ptr<DeviceState> d(new DeviceState);
    d->id = id;
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, id));
    d->num_sms = prop.multiProcessorCount;
    if (prop.major < 10) { err_ = "this runtime is built for sm_100a (Blackwell) only"; return -1; }
    auto alloc = [&](size_t bytes) -> void* {
      void* p = nullptr;
      if (cudaMalloc(&p, bytes ? bytes : 16) != cudaSuccess) return nullptr;
      d->allocs.push_back(p);
      return p;
    };
    for (int k = 0; k < kLanes; k++) {
      Lane& ln = d->lane[k];
      const int chunk = lane_chunk[k];
      CK(cudaStreamCreateWithFlags(&ln.s, cudaStreamNonBlocking));
      CK(cudaEventCreate(&ln.e0));
      CK(cudaEventCreate(&ln.e1));
      ln.d_raw = (uint8_t*)alloc((size_t)chunk * raw_bytes);
      ln.d_in = alloc((size_t)chunk * net_.in_bytes);
      ln.buf[0] = (uint32_t*)alloc(lane_act[k]);
      ln.buf[1] = (uint32_t*)alloc(lane_act[k]);
      ln.d_out = alloc((size_t)chunk * net_.out_bytes);
      ln.d_cls = (int*)alloc((size_t)chunk * 4);
      if (!ln.d_raw || !ln.d_in || !ln.buf[0] || !ln.buf[1] || !ln.d_out || !ln.d_cls) { err_ = "cudaMalloc failed"; return -1; }
    }
    for (int i = 0; i < net_.nlayers; i++) {
      const LayerDesc& l = net_.L[i];
      DevLayer& D = d->L[i];
      D.d = l;
      D.kwords = l.K / 32;
      D.kpad = (l.K + 31) / 32 * 32;
      d->wf[i] = (uint64_t*)alloc((size_t)l.pe * l.wmem * 8);
      d->tf[i] = (int32_t*)alloc((size_t)l.pe * l.tmem * std::max(l.nthr, 1) * 4);
      const bool l0 = net_.is_cnv && i == 0;
      D.wp = (uint32_t*)alloc(l0 ? (size_t)64 * 8 * 4 : (size_t)l.wbits * D.kwords * l.N * 4);
      D.w8 = nullptr;
      D.w4 = nullptr;
      D.w4s = nullptr;
      D.thr_tc = nullptr;
      if (!l0) {  // operands of the tensor-core path
        D.w8 = (int8_t*)alloc((size_t)l.N * D.kpad);
        D.w4 = (uint8_t*)alloc((size_t)l.N * D.kpad / 2);
        D.w4s = (uint8_t*)alloc((size_t)l.N * D.kpad / 2);
        D.thr_tc = (int32_t*)alloc((size_t)l.N * 2 * 4);
        if (!D.w8 || !D.w4 || !D.w4s || !D.thr_tc) { err_ = "cudaMalloc failed"; return -1; }
      }
      D.thr = (int32_t*)alloc((size_t)l.N * 2 * 4);
      D.wf = d->wf[i];
      D.tf = d->tf[i];
      if (!d->wf[i] || !d->tf[i] || !D.wp || !D.thr) { err_ = "cudaMalloc failed"; return -1; }
      d->folded.wf[i] = d->wf[i];
      d->folded.tf[i] = d->tf[i];
      d->folded.wmem[i] = l.wmem;
      d->folded.tmem[i] = l.tmem;
      d->folded.nthr[i] = std::max(l.nthr, 1);
    }
    d->d_status = (int*)alloc(4);
    if (!d->d_status) { err_ = "cudaMalloc failed"; return -1; }
    CK(cudaMemset(d->d_status, 0, 4));
    dev_.push_back(std::move(d));
  }
  return 0;
}

int Engine::upload_all() {
  for (auto& d : dev_) {
    CK(cudaSetDevice(d->id));
    cudaStream_t s = d->lane[0].s;
    for (int i = 0; i < net_.nlayers; i++) {
      CK(cudaMemcpyAsync(d->wf[i], pm_->weights(i).data(), pm_->weights(i).size() * 8, cudaMemcpyHostToDevice, s));
      CK(cudaMemcpyAsync(d->tf[i], pm_->thresholds(i).data(), pm_->thresholds(i).size() * 4, cudaMemcpyHostToDevice, s));
      launch_repack(d->L[i], net_.is_cnv && i == 0, s);
    }
    CK(cudaStreamSynchronize(s));
    CK(cudaGetLastError());
    d->uploaded = true;
  }
  return 0;
}

int Engine::load_parameters(const std::string& dir) {
  std::unique_ptr<ParamMirror> pm(new ParamMirror(net_));
  std::string e;
  if (pm->load(dir, &e) != 0) { err_ = e; return -1; }
  pm_ = std::move(pm);  // a reload discards injected faults
  loaded_ = true;
  if (ensure_devices() != 0) return -1;
  return upload_all();
}

void Engine::enqueue_patches(DeviceState& ds, const std::vector<Patch>& p, cudaStream_t s) {
  // caller has set the device; patches are small: stage through a device array, then re-pack the touched layers
  if (p.empty()) return;
  if (ds.patch_cap < p.size()) {
    DevPatch* np = nullptr;
    cudaMalloc(&np, p.size() * sizeof(DevPatch) * 2);
    ds.allocs.push_back(np);
    ds.d_patches = np;
    ds.patch_cap = p.size() * 2;
  }
  static_assert(sizeof(Patch) == sizeof(DevPatch), "Patch and DevPatch must share a layout");
  cudaMemcpyAsync(ds.d_patches, p.data(), p.size() * sizeof(DevPatch), cudaMemcpyHostToDevice, s);
  launch_patch(ds.folded, ds.d_patches, (int)p.size(), s);
  bool touched[9] = {};
  for (const Patch& e : p) touched[e.layer] = true;
  for (int i = 0; i < net_.nlayers; i++)
    if (touched[i]) launch_repack(ds.L[i], net_.is_cnv && i == 0, s);
}

int Engine::apply_patches(const std::vector<Patch>& patches) {
  if (patches.empty()) return 0;
  if (ensure_devices() != 0) return -1;
  for (auto& d : dev_) {
    CK(cudaSetDevice(d->id));
    if (!d->uploaded) continue;
    for (auto& ln : d->lane) CK(cudaStreamSynchronize(ln.s));
    enqueue_patches(*d, patches, d->lane[0].s);
    CK(cudaStreamSynchronize(d->lane[0].s));  // the host vector may die after return
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

// One chunk through every layer.  Returns the number of kernels launched, or -1.
int Engine::run_chunk(DeviceState& ds, const void* d_in, int n, void* d_out, int* d_classes, int ncls, cudaStream_t s,
                      int32_t* tap_acc, uint32_t** bufs) {
  const void* cur = d_in;
  int launches = 0;
  const bool per_layer = tap_acc != nullptr;  // when tapping, every layer writes its own buffer; else two ping-pong buffers
  int first_layer = 0;
  // Activation format between layers (f4_common.cuh): on the fp4 core, between two fp4 kernels of a 1-bit network, the producer
  // writes e2m1 operand planes in the consumer's own shared-memory layout (the consumer stages them with TMA bulk copies).
  int fam[9] = {}, ofmt[9] = {};  // kernel family of layer i on this core: 1 = conv_f4, 2 = gemm_f4, 0 = other
  if (core_ == 2 && !per_layer)
    for (int i = 0; i < net_.nlayers; i++) fam[i] = conv_f4_supports(net_.L[i]) ? 1 : (gemm_f4_supports(ds.L[i]) ? 2 : 0);
  const bool fused01 = core_ >= 1 && net_.is_cnv && !per_layer && fuse01_;
  for (int i = 0; i + 1 < net_.nlayers; i++) {
    const LayerDesc& p = net_.L[i];
    const LayerDesc& c = net_.L[i + 1];
    const bool producer_f4 = fam[i] != 0 || (fused01 && i == 1 && core_ == 2);
    if (!producer_f4 || p.out_mode != OUT_BIT || c.in_mode != IN_XNOR) continue;
    if (fam[i + 1] == 1 && fam[i] != 2) ofmt[i] = ACT_F4P;                 // conv consumer: image planes
    if (fam[i + 1] == 2 && !p.pool && !(fused01 && i == 1)) ofmt[i] = ACT_F4T;  // gemm consumer: 128-image tiles
  }
  if (core_ >= 1 && net_.is_cnv && !per_layer && fuse01_) {  // layers 0+1 in one tensor-core kernel
    ProfEv pe{nullptr, nullptr, 1, ds.id};
    if (profile_) { cudaEventCreate(&pe.a); cudaEventCreate(&pe.b); cudaEventRecord(pe.a, s); }
    if (launch_cnv01_tc(ds.L[0], ds.L[1], (const int8_t*)d_in, bufs[1], n, ds.num_sms, core_ == 2, ofmt[1], ds.d_status, s)) {
      if (profile_) { cudaEventRecord(pe.b, s); prof_.push_back(pe); }
      {
        const cudaError_t le = cudaGetLastError();
        if (le != cudaSuccess) { err_ = std::string("fused layer 0+1 launch: ") + cudaGetErrorString(le); return -1; }
      }
      launches++;
      cur = bufs[1];
      first_layer = 2;
    } else if (profile_) { cudaEventDestroy(pe.a); cudaEventDestroy(pe.b); }
  }
  for (int i = first_layer; i < net_.nlayers; i++) {
    const bool last = i == net_.nlayers - 1;
    uint32_t* out = last ? (uint32_t*)d_out : bufs[per_layer ? i : (i & 1)];
    int32_t* tp = nullptr;
    if (tap_acc) {
      tp = tap_acc;
      tap_acc += (size_t)n * net_.L[i].ofm_dim * net_.L[i].ofm_dim * net_.L[i].N;
    }
    ProfEv pe{nullptr, nullptr, i, ds.id};
    if (profile_) { cudaEventCreate(&pe.a); cudaEventCreate(&pe.b); cudaEventRecord(pe.a, s); }
    if (net_.is_cnv && i == 0) launch_conv0(ds.L[0], (const int8_t*)cur, out, n, tp, ds.num_sms, s);
    else if (fam[i] == 1 && launch_conv_f4(ds.L[i], (const uint32_t*)cur, out, n, ds.num_sms, i > 0 ? ofmt[i - 1] : ACT_BITS, ofmt[i], ds.d_status, s)) {}
    else if (fam[i] == 2 && launch_gemm_f4(ds.L[i], (const uint32_t*)cur, out, n, ds.num_sms, i > 0 ? ofmt[i - 1] : ACT_BITS, ofmt[i], ds.d_status, s)) {}
    else if (core_ >= 1 && !tp && launch_conv_tc(ds.L[i], (const uint32_t*)cur, out, n, ds.num_sms, ds.d_status, s)) {}
    else if (core_ >= 1 && !tp && launch_gemm_tc(ds.L[i], (const uint32_t*)cur, out, n, ds.num_sms, ds.d_status, s)) {}
    else if ((i > 0 && ofmt[i - 1] != ACT_BITS) || ofmt[i] != ACT_BITS) { err_ = "layer " + std::to_string(i) + ": no kernel for the planned activation format"; return -1; }
    else launch_layer_popc(ds.L[i], (const uint32_t*)cur, out, n, tp, ds.num_sms, s);
    if (profile_) { cudaEventRecord(pe.b, s); prof_.push_back(pe); }
    {  // a launch that the device refuses (resources, configuration) must not pass silently
      const cudaError_t le = cudaGetLastError();
      if (le != cudaSuccess) { err_ = "layer " + std::to_string(i) + " launch: " + cudaGetErrorString(le); return -1; }
    }
    launches++;
    cur = out;
  }
  if (net_.is_cnv && d_classes) { launch_cnv_classes((const int16_t*)d_out, n, ncls, d_classes, s); launches++; }
  return launches;
}

int Engine::infer_device(int slot, const void* d_in, int n, void* d_out, int* d_classes, int ncls, cudaStream_t stream) {
  if (!loaded_) { err_ = "load_parameters has not been called"; return -1; }
  if (ensure_devices() != 0) return -1;
  if (slot < 0 || slot >= (int)dev_.size()) { err_ = "bad device slot"; return -1; }
  DeviceState& ds = *dev_[slot];
  CK(cudaSetDevice(ds.id));
  cudaStream_t s = stream ? stream : ds.lane[0].s;
  const int chunk = chunk_images();
  int launches = 0;
  for (int off = 0; off < n; off += chunk) {
    const int c = std::min(chunk, n - off);
    const int r = run_chunk(ds, (const uint8_t*)d_in + (size_t)off * net_.in_bytes, c, (uint8_t*)d_out + (size_t)off * net_.out_bytes,
                            d_classes ? d_classes + off : nullptr, ncls, s, nullptr, ds.lane[0].buf);
    if (r < 0) return -1;
    launches += r;
  }
  last_launches_ = launches;
  CK(cudaGetLastError());
  return 0;
}

int Engine::prep_device(int slot, const void* d_raw, int n, void* d_in, cudaStream_t stream) {
  if (ensure_devices() != 0) return -1;
  if (slot < 0 || slot >= (int)dev_.size()) { err_ = "bad device slot"; return -1; }
  DeviceState& ds = *dev_[slot];
  CK(cudaSetDevice(ds.id));
  cudaStream_t s = stream ? stream : ds.lane[0].s;
  if (net_.is_cnv) launch_prep_cifar((const uint8_t*)d_raw, (int8_t*)d_in, n, s);
  else launch_prep_mnist((const uint8_t*)d_raw, (uint32_t*)d_in, n, s);
  CK(cudaGetLastError());
  return 0;
}

int Engine::sync(int slot) {
  if (slot < 0 || slot >= (int)dev_.size()) { err_ = "bad device slot"; return -1; }
  CK(cudaSetDevice(dev_[slot]->id));
  for (auto& ln : dev_[slot]->lane) CK(cudaStreamSynchronize(ln.s));
  return 0;
}

// LFC class decode, multi-image rule (foldedmv-offload.cpp:202-220): (unsigned) log2(word & mask), 0 when no bit is set.
static int lfc_class_multi(uint64_t w, int ncls) {
  const uint64_t mask = ncls >= 64 ? ~0ull : (ncls <= 0 ? 0 : (~0ull >> (64 - ncls)));
  w &= mask;
  if (!w) return 0;
  // below 2^32 floor(log2) of the double is the index of the top bit (2^k - 1 is far enough from 2^k for k <= 32); the libm call
  // costs ~20 ns per image, which at 65 536 images per call was most of the LFC host path's time beyond its copies
  if (w < (1ull << 32)) return 31 - __builtin_clz((unsigned)w);
  return (int)(unsigned)std::log2((double)w);
}

int Engine::infer_host(const uint8_t* input, int n, int ncls, int* classes, int16_t* scores, uint64_t* words, float* usec) {
  if (!loaded_) { err_ = "load_parameters has not been called"; return -1; }
  if (ensure_devices() != 0) return -1;
  for (auto& d : dev_) if (!d->uploaded) { if (upload_all() != 0) return -1; break; }
  if (usec) *usec = 0;
  if (n <= 0) return 0;
  const int G = (int)dev_.size(), chunk = chunk_images(true);
  const size_t raw_bytes = net_.is_cnv ? 3073 : 784;
  std::vector<uint64_t> tmp_words;
  if (!net_.is_cnv && !words && classes) { tmp_words.resize(n); words = tmp_words.data(); }
  struct Timed { int dev, lane; };
  std::vector<float> dev_ms(G, 0.f);
  // contiguous shards [g*n/G, (g+1)*n/G): SURVEY 8(e)
  std::vector<int> pos(G), end(G), nchunk(G, 0);
  for (int g = 0; g < G; g++) { pos[g] = (int)((long long)g * n / G); end[g] = (int)((long long)(g + 1) * n / G); }
  bool more = true;
  while (more) {  // round-robin over devices so that all of them are fed concurrently
    more = false;
    for (int g = 0; g < G; g++) {
      if (pos[g] >= end[g]) continue;
      DeviceState& ds = *dev_[g];
      CK(cudaSetDevice(ds.id));
      Lane& ln = ds.lane[nchunk[g] % kLanes];
      if (nchunk[g] >= kLanes) {  // the lane's previous chunk must be finished before its events/buffers are reused
        CK(cudaStreamSynchronize(ln.s));
        float ms = 0;
        CK(cudaEventElapsedTime(&ms, ln.e0, ln.e1));
        dev_ms[g] += ms;
      }
      // The copy engine is the bottleneck of this path (3 073 B per image over PCIe): only the compute of the LAST chunk is not
      // hidden behind a copy, so the tail of a shard is cut into halves (full chunks elsewhere keep the kernels efficient).
      const int off = pos[g], rem = end[g] - off;
      const int c = rem > chunk + chunk / 2 ? chunk : (rem > 4096 ? std::min(chunk, (rem / 2 + 1023) / 1024 * 1024) : rem);
      CK(cudaMemcpyAsync(ln.d_raw, input + (size_t)off * raw_bytes, (size_t)c * raw_bytes, cudaMemcpyHostToDevice, ln.s));
      if (net_.is_cnv) launch_prep_cifar(ln.d_raw, (int8_t*)ln.d_in, c, ln.s);
      else launch_prep_mnist(ln.d_raw, (uint32_t*)ln.d_in, c, ln.s);
      CK(cudaEventRecord(ln.e0, ln.s));
      if (run_chunk(ds, ln.d_in, c, ln.d_out, (net_.is_cnv && classes) ? ln.d_cls : nullptr, ncls, ln.s, nullptr, ln.buf) < 0) return -1;
      CK(cudaEventRecord(ln.e1, ln.s));
      if (net_.is_cnv) {
        if (scores) CK(cudaMemcpyAsync(scores + (size_t)off * 64, ln.d_out, (size_t)c * 128, cudaMemcpyDeviceToHost, ln.s));
        if (classes) CK(cudaMemcpyAsync(classes + off, ln.d_cls, (size_t)c * 4, cudaMemcpyDeviceToHost, ln.s));
      } else if (words) {
        CK(cudaMemcpyAsync(words + off, ln.d_out, (size_t)c * 8, cudaMemcpyDeviceToHost, ln.s));
      }
      pos[g] += c;
      nchunk[g]++;
      more = true;
    }
  }
  float worst = 0;
  for (int g = 0; g < G; g++) {
    DeviceState& ds = *dev_[g];
    CK(cudaSetDevice(ds.id));
    for (int k = 0; k < std::min(nchunk[g], kLanes); k++) {
      Lane& ln = ds.lane[k];
      CK(cudaStreamSynchronize(ln.s));
      float ms = 0;
      CK(cudaEventElapsedTime(&ms, ln.e0, ln.e1));
      dev_ms[g] += ms;
    }
    CK(cudaGetLastError());
    if (check_status(g) != 0) return -1;
    worst = std::max(worst, dev_ms[g]);
  }
  if (usec) *usec = worst * 1000.0f / (float)n;
  if (!net_.is_cnv && classes)
    for (int i = 0; i < n; i++) classes[i] = lfc_class_multi(words[i], ncls);
  return 0;
}

int Engine::infer_host_with_faults(const uint8_t* input, int n, int ncls, const std::vector<Fault>& faults_in, int* classes,
                                   float* usec) {
  if (!loaded_) { err_ = "load_parameters has not been called"; return -1; }
  if (ensure_devices() != 0) return -1;
  for (auto& d : dev_) if (!d->uploaded) { if (upload_all() != 0) return -1; break; }
  if (usec) *usec = 0;
  // 1. host: run the whole read-modify-write chain in fault order (it does not depend on inference results)
  std::vector<Fault> faults = faults_in;
  std::stable_sort(faults.begin(), faults.end(), [](const Fault& a, const Fault& b) { return a.time < b.time; });
  std::vector<Patch> patches;      // one per effective fault, in order
  std::vector<uint32_t> ptime;
  for (const Fault& f : faults) {
    Patch p;
    if (pm_->inject(f.target, f.layer, f.bit, f.word_size, &p)) { patches.push_back(p); ptime.push_back(f.time); }
  }
  if (n <= 0) return apply_patches(patches);
  const int G = (int)dev_.size();
  const size_t raw_bytes = net_.is_cnv ? 3073 : 784;
  struct Shard { uint8_t* raw = nullptr; void* in = nullptr; void* out = nullptr; int* cls = nullptr; DevPatch* pt = nullptr; int lo = 0, hi = 0; };
  std::vector<Shard> sh(G);
  std::vector<uint64_t> words;
  if (!net_.is_cnv) words.resize(n);
  int rc = 0;
  float last_ms = 0;
  for (int g = 0; g < G && rc == 0; g++) {
    DeviceState& ds = *dev_[g];
    Shard& S = sh[g];
    S.lo = (int)((long long)g * n / G); S.hi = (int)((long long)(g + 1) * n / G);
    const int cnt = S.hi - S.lo;
    CK(cudaSetDevice(ds.id));
    cudaStream_t s = ds.lane[0].s;
    for (auto& ln : ds.lane) CK(cudaStreamSynchronize(ln.s));
    CK(cudaMalloc(&S.raw, std::max<size_t>(16, (size_t)cnt * raw_bytes)));
    CK(cudaMalloc(&S.in, std::max<size_t>(16, (size_t)cnt * net_.in_bytes)));
    CK(cudaMalloc(&S.out, std::max<size_t>(16, (size_t)cnt * net_.out_bytes)));
    CK(cudaMalloc(&S.cls, std::max<size_t>(16, (size_t)cnt * 4)));
    CK(cudaMalloc(&S.pt, std::max<size_t>(16, patches.size() * sizeof(DevPatch))));
    if (!patches.empty()) CK(cudaMemcpyAsync(S.pt, patches.data(), patches.size() * sizeof(DevPatch), cudaMemcpyHostToDevice, s));
    if (cnt > 0) {
      CK(cudaMemcpyAsync(S.raw, input + (size_t)S.lo * raw_bytes, (size_t)cnt * raw_bytes, cudaMemcpyHostToDevice, s));
      if (net_.is_cnv) launch_prep_cifar(S.raw, (int8_t*)S.in, cnt, s);
      else launch_prep_mnist(S.raw, (uint32_t*)S.in, cnt, s);
    }
    // 2. epochs: apply every fault with time <= image index before that image (faults.h:138-146)
    size_t pi = 0;
    auto apply_upto = [&](uint32_t t_incl) {  // patches with time <= t_incl
      const size_t b = pi;
      bool touched[9] = {};
      while (pi < patches.size() && ptime[pi] <= t_incl) { touched[patches[pi].layer] = true; pi++; }
      if (pi > b) {
        launch_patch(ds.folded, S.pt + b, (int)(pi - b), s);
        for (int i = 0; i < net_.nlayers; i++)
          if (touched[i]) launch_repack(ds.L[i], net_.is_cnv && i == 0, s);
      }
    };
    int cur = S.lo;
    const int chunk = chunk_images();
    while (cur < S.hi) {
      apply_upto((uint32_t)cur);
      int nxt = S.hi;
      if (pi < patches.size() && (long long)ptime[pi] < (long long)S.hi) nxt = (int)ptime[pi];
      for (int off = cur; off < nxt; off += chunk) {  // segment [cur, nxt) with a fixed parameter set
        const int c = std::min(chunk, nxt - off);
        const bool timed = (g == G - 1) && (off + c == S.hi);  // the reference reports the LAST image's time
        if (timed) CK(cudaEventRecord(ds.lane[0].e0, s));
        if (run_chunk(ds, (const uint8_t*)S.in + (size_t)(off - S.lo) * net_.in_bytes, c, (uint8_t*)S.out + (size_t)(off - S.lo) * net_.out_bytes,
                      net_.is_cnv ? S.cls + (off - S.lo) : nullptr, ncls, s, nullptr, ds.lane[0].buf) < 0) return -1;
        if (timed) CK(cudaEventRecord(ds.lane[0].e1, s));
      }
      cur = nxt;
    }
    apply_upto(0xffffffffu);  // faults scheduled at or after this shard's end still persist on this device
    if (cnt > 0) {
      if (net_.is_cnv) CK(cudaMemcpyAsync(classes + S.lo, S.cls, (size_t)cnt * 4, cudaMemcpyDeviceToHost, s));
      else CK(cudaMemcpyAsync(words.data() + S.lo, S.out, (size_t)cnt * 8, cudaMemcpyDeviceToHost, s));
    }
  }
  for (int g = 0; g < G; g++) {
    DeviceState& ds = *dev_[g];
    cudaSetDevice(ds.id);
    cudaError_t e = cudaStreamSynchronize(ds.lane[0].s);
    if (e == cudaSuccess) e = cudaGetLastError();
    if (e != cudaSuccess) { err_ = cudaGetErrorString(e); rc = -1; }
    if (g == G - 1 && sh[g].hi > sh[g].lo && rc == 0) cudaEventElapsedTime(&last_ms, ds.lane[0].e0, ds.lane[0].e1);
    cudaFree(sh[g].raw); cudaFree(sh[g].in); cudaFree(sh[g].out); cudaFree(sh[g].cls); cudaFree(sh[g].pt);
  }
  if (rc != 0) return rc;
  if (!net_.is_cnv)
    for (int i = 0; i < n; i++) classes[i] = lfc_class_multi(words[i], ncls);
  if (usec) *usec = last_ms * 1000.0f;
  return 0;
}

int Engine::profile_collect(float* ms, int* launches) {
  for (int i = 0; i < 10; i++) { ms[i] = 0; launches[i] = 0; }
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

int Engine::tap_host(const uint8_t* input, int n, int32_t* acc, int8_t* act) {
  if (!loaded_) { err_ = "load_parameters has not been called"; return -1; }
  if (ensure_devices() != 0) return -1;
  DeviceState& ds = *dev_[0];
  CK(cudaSetDevice(ds.id));
  if (!ds.uploaded && upload_all() != 0) return -1;
  cudaStream_t s = ds.lane[0].s;
  const size_t raw_bytes = net_.is_cnv ? 3073 : 784;
  const size_t sa = tap_acc_size(), st = tap_act_size();
  const int step = 16;
  uint8_t* d_raw; void* d_in; int32_t* d_acc; uint32_t* bufs[9];
  CK(cudaMalloc(&d_raw, step * raw_bytes));
  CK(cudaMalloc(&d_in, (size_t)step * net_.in_bytes));
  CK(cudaMalloc(&d_acc, (size_t)step * sa * 4));
  std::vector<size_t> words(net_.nlayers);
  for (int i = 0; i < net_.nlayers; i++) {
    const LayerDesc& l = net_.L[i];
    words[i] = (size_t)out_dim(l) * out_dim(l) * out_words_per_pixel(l);
    CK(cudaMalloc(&bufs[i], step * words[i] * 4));
  }
  std::vector<int32_t> hacc((size_t)step * sa);
  std::vector<uint32_t> hbuf;
  for (int off = 0; off < n; off += step) {
    const int c = std::min(step, n - off);
    CK(cudaMemcpyAsync(d_raw, input + (size_t)off * raw_bytes, c * raw_bytes, cudaMemcpyHostToDevice, s));
    if (net_.is_cnv) launch_prep_cifar(d_raw, (int8_t*)d_in, c, s);
    else launch_prep_mnist(d_raw, (uint32_t*)d_in, c, s);
    if (run_chunk(ds, d_in, c, bufs[net_.nlayers - 1], nullptr, 10, s, d_acc, bufs) < 0) return -1;
    CK(cudaMemcpyAsync(hacc.data(), d_acc, (size_t)c * sa * 4, cudaMemcpyDeviceToHost, s));
    CK(cudaStreamSynchronize(s));
    CK(cudaGetLastError());
    // device accumulators are layer-major over the chunk; the API is image-major
    size_t lo = 0, io = 0;
    for (int i = 0; i < net_.nlayers; i++) {
      const LayerDesc& l = net_.L[i];
      const size_t per = (size_t)l.ofm_dim * l.ofm_dim * l.N;
      for (int k = 0; k < c; k++) std::memcpy(acc + (size_t)(off + k) * sa + io, hacc.data() + lo + (size_t)k * per, per * 4);
      lo += (size_t)c * per;
      io += per;
    }
    size_t ao = 0;
    for (int i = 0; i < net_.nlayers; i++) {
      const LayerDesc& l = net_.L[i];
      const int px = out_dim(l) * out_dim(l), nw = l.N / 32;
      hbuf.resize((size_t)c * words[i]);
      CK(cudaMemcpy(hbuf.data(), bufs[i], (size_t)c * words[i] * 4, cudaMemcpyDeviceToHost));
      for (int k = 0; k < c; k++)
        for (int p = 0; p < px; p++)
          for (int ch = 0; ch < l.N; ch++) {
            int8_t v;
            const uint32_t* w = hbuf.data() + (size_t)k * words[i] + (size_t)p * out_words_per_pixel(l);
            if (l.out_mode == OUT_SCORE) v = (int8_t)reinterpret_cast<const int16_t*>(w)[ch];
            else if (l.out_mode == OUT_BIT) v = (w[ch >> 5] >> (ch & 31)) & 1;
            else { const int nz = (w[ch >> 5] >> (ch & 31)) & 1, ps = (w[nw + (ch >> 5)] >> (ch & 31)) & 1; v = nz ? (ps ? 1 : -1) : 0; }
            act[(size_t)(off + k) * st + ao + (size_t)p * l.N + ch] = v;
          }
      ao += (size_t)px * l.N;
    }
  }
  cudaFree(d_raw); cudaFree(d_in); cudaFree(d_acc);
  for (int i = 0; i < net_.nlayers; i++) cudaFree(bufs[i]);
  return 0;
}

void* Engine::alloc_pinned(size_t bytes) {
  void* p = nullptr;
  if (cudaHostAlloc(&p, bytes ? bytes : 16, cudaHostAllocPortable) != cudaSuccess) return nullptr;
  return p;
}
void Engine::free_pinned(void* p) { if (p) cudaFreeHost(p); }

}  // namespace bnnb
