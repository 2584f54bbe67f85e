// The CUDA runtime behind the C ABI: per-device replicated parameters, a few chunk pipelines (lanes) per device, buffers grown on
// demand, fault epochs as in-place operand patches on a per-lane parameter set, and contiguous sharding
// of an image batch over the visible GPUs (no collective; one feeder thread per device).
// Replaces the host library's accelerator glue: FoldedMVInit/Deinit, FoldedMVOffloadBinarized (rawhls-offload.cpp:54-94),
// testPrebuiltCIFAR10_multiple_images / _from_image (foldedmv-offload.h:316-416),
// testPrebinarized_nolabel(_multiple_images) (foldedmv-offload.cpp:114-224) and BlackBoxJam's dispatch (top.cpp:238-291).
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <memory>
#include <mutex>
#include <string>
#include <vector>
#include "kernels.cuh"
#include "netdesc.hpp"
#include "params.hpp"

namespace bnnb {

struct DeviceState;
struct TapCtx;
struct FaultPlan;

class Engine {
 public:
  explicit Engine(const NetDesc& net);
  ~Engine();
  const NetDesc& net() const { return net_; }
  const std::string& last_error() const { return err_; }

  // Devices used for sharding (default: every visible device, or BNN_B200_DEVICES="0,1,..").
  int set_devices(const int* ids, int n);
  int num_devices() const { return (int)dev_.size(); }

  int load_parameters(const std::string& dir);
  bool loaded() const { return loaded_; }
  ParamMirror& mirror() { return *pm_; }
  // Frees every device allocation (parameters, buffers, streams).  The host mirror - parameters and injected faults - is kept, so
  // the next call re-creates the device state on demand: what deinit() needs (FoldedMVDeinit frees buffers while the parameters
  // survive in static memories, rawhls-offload.cpp:69-74, top.cpp:51-68; every entry point re-initialises, main_python.cpp:126).
  int release();
  // Bytes currently allocated on device slot `slot` (tests: footprint of small calls).
  size_t device_bytes(int slot) const;

  // Host-buffer inference.  `input`: CNV n x 3073-byte CIFAR records, LFC n x 784 pixel bytes.
  // classes (n, multi-image rule) / scores (CNV: n x 64 int16) / words (LFC: n x uint64) may be null.
  // usec_per_image: device time of the inference kernels only (the reference's timed region, foldedmv-offload.h:389-392): per
  // device the union of its chunks' CUDA-event intervals (lanes overlap), max over devices, / n.
  // Batches are sharded in contiguous ranges over the devices, but never thinner than kMinShard images per device.
  int infer_host(const uint8_t* input, int n, int ncls, int* classes, int16_t* scores, uint64_t* words, float* usec_per_image);

  // infer_host on inputs already in BlackBoxJam's format (LFC 104 B / CNV 3 072 B per image): no preparation kernel.
  int infer_host_packed(const uint8_t* input, int n, int ncls, int* classes, int16_t* scores, uint64_t* words, float* usec_per_image);

  // Same, with faults applied between image segments ("epochs"): every fault with time i is applied before image i,
  // cumulatively, and they persist afterwards (faults.h:115-148).  All devices apply every fault; each runs only its shard.
  // usec_per_image: device time per image of the last segment (the reference reports its last single-image call).
  int infer_host_with_faults(const uint8_t* input, int n, int ncls, const std::vector<Fault>& faults, int* classes,
                             float* usec_per_image);

  // Device-resident inference on device slot `slot` (index into the device list): `d_in` is the layer-0 input exactly as
  // BlackBoxJam reads it (CNV int8 HWC 3072 B/img; LFC 104 B/img), `d_out` what it writes (CNV 64 x int16; LFC uint64),
  // `d_classes` optional int32 (CNV only).  Enqueued on `stream` (0 = the engine's own stream); returns without syncing.
  int infer_device(int slot, const void* d_in, int n, void* d_out, int* d_classes, int ncls, cudaStream_t stream);
  // infer_device without any allocation: the activation scratch is a static 32 MiB device array (engine.cu), so n is limited to
  // static_max_images() per call, the device must already hold the parameters (one ordinary call made before), and calls that overlap
  // on one GPU race on the shared array.  No counterpart in the reference.
  int infer_device_static(int slot, const void* d_in, int n, void* d_out, int* d_classes, int ncls, cudaStream_t stream);
  int static_max_images() const;
  // raw records / pixels (device) -> layer-0 input (device)
  int prep_device(int slot, const void* d_raw, int n, void* d_in, cudaStream_t stream);
  // Waits for the slot's streams and reports a kernel that gave up waiting on an mbarrier (device status flag) as an error.
  int sync(int slot);
  // Number of kernels the last infer_device call enqueued (per call; for bench.py's gpu_launches).
  int last_launches() const { return last_launches_; }

  // Debug taps for parity tests, ON THE SELECTED CORE: per image, concatenated over layers, int32 accumulators [ofm^2][N] (the
  // reference's accumulator values) and the layer outputs after pooling as int8 values [out_dim^2][N] (last CNV layer: low 8 bits
  // of the score).  Core 0: the LOP3+POPC kernels.  Cores 1 / 2: the TAP instantiations of the tensor-core kernels that ship
  // (same source, extra stores), run with the production activation formats; outputs are decoded from those formats.
  int tap_host(const uint8_t* input, int n, int32_t* acc, int8_t* act);
  size_t tap_acc_size() const;
  size_t tap_act_size() const;

  // Deterministic fault door (inject_fault's arguments) and accelerator-memory access (FoldedMVMemRead/MemSet).
  int inject_fault(uint32_t target, uint32_t layer, uint32_t bit, uint32_t word_size);
  uint64_t mem_read(unsigned tl, unsigned mem, unsigned ind, unsigned thresh);
  int mem_set(unsigned tl, unsigned mem, unsigned ind, unsigned thresh, uint64_t val);

  // Per-layer device timing: while enabled, run_chunk brackets every layer kernel with CUDA events on the launching stream.
  // profile_collect synchronises, returns per-layer total milliseconds and launch counts since the last collect (arrays of 10:
  // layers 0..8 then the class-decode kernel) and clears the pool.
  void profile_enable(bool on) { profile_ = on; }
  int profile_collect(float* ms, int* launches);

  // Compute core: 0 = LOP3+POPC CUDA-core kernels everywhere, 1 = tcgen05 kind::i8 where instantiated,
  // 2 (default) = tcgen05 kind::mxf4 (exact fp4) where instantiated, else kind::i8, else LOP3+POPC.
  void set_core(int c) { core_ = c; }
  int core() const { return core_; }

  static constexpr int kMinShard = 2048;  // images: below this a device is not worth its launch latency

 private:
  struct Segments;  // fault epochs of one call
  int ensure_devices();
  int upload(DeviceState& ds, std::string& err);
  int check_status(DeviceState& ds, std::string& err);
  int apply_patches(const std::vector<Patch>& patches);                 // all uploaded devices, stream-ordered
  int ensure_buffers(DeviceState& ds, int lane, int images, std::string& err);  // lane's d_in + activation ping-pong
  int ensure_lane(DeviceState& ds, int lane, int images, std::string& err);     // lane's staging buffers
  int ensure_patch_cap(DeviceState& ds, size_t count, std::string& err);
  void plan_formats(const DeviceState& ds, bool per_layer, bool tensor_taps, int* fam, int* ofmt, bool* fused01, bool no_tiles = false) const;
  int run_chunk_faulted(DeviceState& ds, const DevLayer* PL, DevLayer* d_layers, int lane, const void* d_in, int n, long long img0, void* d_out,
                        int* d_classes, int ncls, cudaStream_t s, uint32_t** bufs, FaultPlan& fp, std::string& err);
  int run_chunk(DeviceState& ds, const void* d_in, int n, void* d_out, int* d_classes, int ncls, cudaStream_t s, TapCtx* tap,
                uint32_t** bufs, std::string& err);
  int run_single(DeviceState& ds, const uint8_t* rec, size_t raw_bytes, int ncls, int* classes, int16_t* scores, uint64_t* words, float* dev_ms,
                 std::string& err);
  int run_shard(int g, int lo, int hi, const uint8_t* input, int ncls, int* classes, int16_t* scores, uint64_t* words,
                const Segments* seg, float* dev_ms, float* last_ms, int* last_imgs, std::string& err);
  int run_shard_faulted(int g, int lo, int hi, const uint8_t* input, int ncls, int* classes, uint64_t* words, const Segments& seg,
                        float* last_ms, int* last_imgs, std::string& err);
  int run_sharded(const uint8_t* input, int n, int ncls, int* classes, int16_t* scores, uint64_t* words, const Segments* seg,
                  float* dev_ms_max, float* last_ms, int* last_imgs);
  int chunk_images(bool host = false) const;
  size_t act_bytes(int images) const;
  std::vector<int> row_neg_counts(int layer) const;  // #(-1) weights per logical row, from the host mirror (taps)

  NetDesc net_;
  std::unique_ptr<ParamMirror> pm_;
  std::vector<std::unique_ptr<DeviceState>> dev_;
  std::vector<int> dev_ids_;
  bool loaded_ = false;
  int last_launches_ = 0;
  bool profile_ = false;
  int core_ = 2;
  bool fuse01_ = true;
  bool chain_ = true;     // trailing fully-connected layers in one launch (fc_chain.cu); BNN_B200_CHAIN=0 launches them one by one
  bool packed_ = false;  // the host-buffer call in progress takes prepared inputs (infer_host_packed)
  struct ProfEv { cudaEvent_t a, b; int layer, dev; };
  std::vector<ProfEv> prof_;
  std::mutex prof_mu_;
  std::string err_;
};

}  // namespace bnnb
