// Host-side parameter mirror in the REFERENCE's folded coordinates [layer][pe][ind] / [layer][pe][ind][thresh],
// the .bin loader, the accelerator-memory read/write semantics and the fault address map.
// Replaces: FoldedMVLoadLayerMem/LoadWeightPE/LoadThreshPE (bnn/src/library/host/foldedmv-offload.cpp:281-336),
// FoldedMVMemSet/MemRead (rawhls-offload.cpp:76-88), DoMemInit/DoMemRead (network/cnvW1A1/hw/top.cpp:78-178),
// inject_fault/inject_fault_impl (foldedmv-offload.h:146-214), FINNTopology (topology.h:29-178),
// inject_random_fault / make_faulty_classification_func scheduling (faults.h:60-148).
#pragma once
#include <cstdint>
#include <random>
#include <string>
#include <vector>
#include "netdesc.hpp"

namespace bnnb {

// One parameter word to (re)write on the device: folded coordinates + the value now stored there.
struct Patch {
  uint32_t layer, is_thr, pe, ind, thresh, pad;
  uint64_t value;  // weights: the SIMD*WPI-bit word; thresholds: int32 raw threshold (sign-extended into 64 bits)
};

struct Fault {      // one scheduled fault, in the reference injector's terms (inject_fault arguments)
  uint32_t time;    // image index before which it is applied (faults.h:138-146)
  uint32_t target;  // 1 = weights, 2 = thresholds ("Activations")
  uint32_t layer, bit, word_size;
};

class ParamMirror {
 public:
  explicit ParamMirror(const NetDesc& n);
  const NetDesc& net() const { return net_; }
  // Returns 0, or -1 and fills err (the reference throws "Could not open file", foldedmv-offload.cpp:321-323).
  int load(const std::string& dir, std::string* err);
  void mem_set(unsigned target_layer, unsigned mem, unsigned ind, unsigned thresh, uint64_t val);
  uint64_t mem_read(unsigned target_layer, unsigned mem, unsigned ind, unsigned thresh) const;
  // Applies one fault to the mirror (read-modify-write) and reports the rewritten word.  Returns false if nothing changed.
  bool inject(uint32_t target, uint32_t layer, uint32_t bit, uint32_t word_size, Patch* out);
  uint32_t weight_layer_bits(int l) const;   // topology.h:55-62
  uint32_t thresh_layer_bits(int l) const;   // topology.h:63-70
  const std::vector<uint64_t>& weights(int l) const { return w_[l]; }
  const std::vector<int32_t>& thresholds(int l) const { return t_[l]; }

 private:
  NetDesc net_;
  std::vector<std::vector<uint64_t>> w_;  // [layer][pe*wmem + ind]
  std::vector<std::vector<int32_t>> t_;   // [layer][(pe*tmem + ind)*nthr + thresh]
};

// Draws `flip_count` random faults over `num_images` exactly as the reference schedules them
// (fault times uniform in [0,N), target class proportional to bit counts, layer proportional to its bits among the targets,
// bit uniform).  `seed` < 0: std::random_device like the reference (not reproducible).
// `word_size` goes through the reference's bool collapse (faults.h:42 vs :55) unless `exact_word_size`.
std::vector<Fault> schedule_random_faults(const ParamMirror& pm, uint32_t num_images, uint32_t flip_count, int word_size,
                                          int target, const int* target_layers, uint32_t num_targets, long long seed,
                                          bool exact_word_size);

}  // namespace bnnb
