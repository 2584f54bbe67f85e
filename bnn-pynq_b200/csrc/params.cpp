#include "params.hpp"
#include <algorithm>
#include <cstdio>
#include <cstring>

namespace bnnb {

static inline int64_t sext(uint64_t v, int bits) {
  const uint64_t m = 1ull << (bits - 1);
  v &= (m << 1) - 1;
  return (int64_t)((v ^ m) - m);
}

ParamMirror::ParamMirror(const NetDesc& n) : net_(n) {
  w_.resize(n.nlayers);
  t_.resize(n.nlayers);
  for (int i = 0; i < n.nlayers; i++) {
    const LayerDesc& l = n.L[i];
    w_[i].assign((size_t)l.pe * l.wmem, 0);
    t_[i].assign((size_t)l.pe * l.tmem * std::max(l.nthr, 1), 0);
  }
}

void ParamMirror::mem_set(unsigned tl, unsigned mem, unsigned ind, unsigned thresh, uint64_t val) {
  const unsigned li = tl >> 1;
  if ((int)li >= net_.nlayers) return;
  const LayerDesc& l = net_.L[li];
  if (mem >= (unsigned)l.pe) return;
  if (!(tl & 1)) {
    if (ind >= (unsigned)l.wmem) return;
    const int bits = l.simd * l.wbits;  // ap_uint<SIMD*WPI> = ap_uint<64> truncates (top.cpp:81)
    w_[li][(size_t)mem * l.wmem + ind] = bits >= 64 ? val : (val & ((1ull << bits) - 1));
  } else {
    if (l.nthr == 0 || ind >= (unsigned)l.tmem || thresh >= (unsigned)l.nthr) return;  // top.cpp:131-133
    int32_t& p = t_[li][((size_t)mem * l.tmem + ind) * l.nthr + thresh];
    // CNV layer 0: the word is bit-cast to ap_fixed<64,56> and assigned to ap_fixed<24,16> => low 24 bits, raw (top.cpp:84)
    p = (net_.is_cnv && li == 0) ? (int32_t)sext(val, 24) : (int32_t)sext(val, 16);
  }
}

uint64_t ParamMirror::mem_read(unsigned tl, unsigned mem, unsigned ind, unsigned thresh) const {
  const unsigned li = tl >> 1;
  if ((int)li >= net_.nlayers) return 0;
  const LayerDesc& l = net_.L[li];
  if (mem >= (unsigned)l.pe) return 0;
  if (!(tl & 1)) return ind < (unsigned)l.wmem ? w_[li][(size_t)mem * l.wmem + ind] : 0;
  if (l.nthr == 0 || ind >= (unsigned)l.tmem || thresh >= (unsigned)l.nthr) return 0;
  const int32_t r = t_[li][((size_t)mem * l.tmem + ind) * l.nthr + thresh];
  if (net_.is_cnv && li == 0) {
    // static_cast<ap_uint<64>>(ap_fixed<24,16>) is a VALUE cast (top.cpp:142): the 8 fraction bits are dropped,
    // toward zero (C-native conversion of the HLS fixed-point type); a later MemSet bit-casts, so an RMW divides by 256.
    int64_t q = r >> 8;
    if (r < 0 && (r & 0xff)) ++q;
    return (uint64_t)q;
  }
  return (uint64_t)(int64_t)r;  // ap_int<16> -> ap_uint<64> sign-extends
}

int ParamMirror::load(const std::string& dir, std::string* err) {
  for (int li = 0; li < net_.nlayers; li++) {
    const LayerDesc& l = net_.L[li];
    for (int pe = 0; pe < l.pe; pe++) {
      std::string path = dir + "/" + std::to_string(li) + "-" + std::to_string(pe) + "-weights.bin";
      FILE* f = std::fopen(path.c_str(), "rb");
      if (!f) { if (err) *err = "Could not open file " + path; return -1; }
      for (int line = 0; line < l.wmem; line++) {
        uint64_t e = 0;  // a short file leaves the remaining words 0 (foldedmv-offload.cpp:283-284)
        if (std::fread(&e, 8, 1, f) != 1) e = 0;
        mem_set(2 * li, pe, line, 0, e);
      }
      std::fclose(f);
      if (l.nthr > 0) {
        path = dir + "/" + std::to_string(li) + "-" + std::to_string(pe) + "-thres.bin";
        f = std::fopen(path.c_str(), "rb");
        if (!f) { if (err) *err = "Could not open file " + path; return -1; }
        for (int line = 0; line < l.tmem; line++)
          for (int i = 0; i < l.nthr; i++) {
            uint64_t e = 0;
            if (std::fread(&e, 8, 1, f) != 1) e = 0;
            mem_set(2 * li + 1, pe, line, i, e);
          }
        std::fclose(f);
      }
    }
  }
  return 0;
}

uint32_t ParamMirror::weight_layer_bits(int l) const {
  const LayerDesc& L = net_.L[l];
  return (uint32_t)L.wbits * L.simd * L.pe * L.wmem;
}
uint32_t ParamMirror::thresh_layer_bits(int l) const {
  const LayerDesc& L = net_.L[l];
  return (uint32_t)L.tmem * L.pe * L.api * L.thr_bits;
}

bool ParamMirror::inject(uint32_t target, uint32_t layer, uint32_t bit, uint32_t word_size, Patch* out) {
  if ((int)layer >= net_.nlayers) return false;
  const LayerDesc& l = net_.L[layer];
  uint32_t mem, ind, thresh, tl;
  if (target == 1) {  // foldedmv-offload.h:192-200
    const uint32_t es = l.simd * l.wbits, el = bit / es;
    thresh = 0; ind = el % l.wmem; mem = (el / l.wmem) % l.pe; bit %= es; tl = layer * 2;
  } else {            // foldedmv-offload.h:202-211
    const uint32_t es = l.thr_bits;
    if (!es) return false;  // a layer without threshold bits is never drawn (weight 0 in the layer distribution)
    const uint32_t el = bit / es;
    thresh = el % l.api; ind = (el / l.api) % l.tmem; mem = ((el / l.api) / l.tmem) % l.pe; bit %= es; tl = layer * 2 + 1;
  }
  if (word_size == 0) return false;  // reference: bit_pos / word_size divides by zero (foldedmv-offload.h:159); we flip nothing
  uint64_t val = mem_read(tl, mem, ind, thresh), mask = 0;
  for (uint32_t i = 0; i < word_size && i < 64; i++) mask |= 1ull << i;
  const uint32_t sh = (bit / word_size) * word_size;
  val ^= sh < 64 ? (mask << sh) : 0;
  mem_set(tl, mem, ind, thresh, val);
  if (out) {
    out->layer = layer; out->is_thr = tl & 1; out->pe = mem; out->ind = ind; out->thresh = thresh; out->pad = 0;
    if (tl & 1) {
      if (l.nthr == 0 || thresh >= (uint32_t)l.nthr) return false;
      out->value = (uint64_t)(int64_t)t_[layer][((size_t)mem * l.tmem + ind) * l.nthr + thresh];
    } else {
      out->value = w_[layer][(size_t)mem * l.wmem + ind];
    }
  }
  return true;
}

std::vector<Fault> schedule_random_faults(const ParamMirror& pm, uint32_t num_images, uint32_t flip_count, int word_size,
                                          int target, const int* target_layers, uint32_t num_targets, long long seed,
                                          bool exact_word_size) {
  std::vector<Fault> out;
  if (num_images == 0) return out;
  const NetDesc& net = pm.net();
  std::random_device rd;
  auto mk = [&](uint32_t salt) { return seed < 0 ? std::mt19937{rd()} : std::mt19937{(uint32_t)(seed * 2654435761u + salt)}; };
  std::mt19937 time_gen = mk(0), any_gen = mk(1), topo_gen = mk(2);
  // RandomFaultArgs is built through a std::function whose word_size parameter is `bool` (faults.h:42) although the
  // injector takes uint8_t (:55): every non-zero word_size reaches inject_fault as 1.
  const uint32_t ws = exact_word_size ? (uint32_t)(uint8_t)word_size : (uint32_t)((uint8_t)word_size != 0);
  std::vector<uint32_t> layers;
  for (uint32_t i = 0; target_layers && i < num_targets; i++)
    if (target_layers[i] >= 0 && target_layers[i] < net.nlayers) layers.push_back((uint32_t)target_layers[i]);
  std::vector<double> wbits(net.nlayers), tbits(net.nlayers);
  for (int l = 0; l < net.nlayers; l++) { wbits[l] = pm.weight_layer_bits(l); tbits[l] = pm.thresh_layer_bits(l); }
  auto space = [&](const std::vector<double>& b) {  // topology.h:80-106 (sum over the target list, or all layers)
    double s = 0;
    if (layers.empty()) for (double v : b) s += v;
    else for (size_t i = 0; i < std::min(layers.size(), b.size()); i++) s += b[layers[i]];
    return s;
  };
  auto pick = [&](const std::vector<double>& b, uint32_t* layer, uint32_t* bit) {  // topology.h:108-160
    uint32_t l;
    if (layers.empty()) {
      std::discrete_distribution<uint32_t> d(b.begin(), b.end());
      l = d(topo_gen);
    } else {
      std::vector<double> f;
      for (uint32_t i : layers) f.push_back(b[i]);
      std::discrete_distribution<uint32_t> d(f.begin(), f.end());
      l = layers[d(topo_gen)];
    }
    if (b[l] < 1) return false;
    std::uniform_int_distribution<uint32_t> bd{0, (uint32_t)b[l] - 1};
    *layer = l; *bit = bd(topo_gen);
    return true;
  };
  std::uniform_int_distribution<size_t> loop_dist{0, (size_t)num_images - 1};  // faults.h:127
  std::vector<uint32_t> times(flip_count);
  for (auto& t : times) t = (uint32_t)loop_dist(time_gen);
  std::stable_sort(times.begin(), times.end());
  for (uint32_t t : times) {
    uint32_t tgt;
    if (target < 0) {  // Any: weights vs thresholds in proportion to their bit counts (faults.h:63-75)
      std::discrete_distribution<uint32_t> wa{space(wbits), space(tbits)};
      tgt = wa(any_gen) == 0 ? 1 : 2;
    } else {
      tgt = target == 0 ? 1 : 2;  // main_python.cpp:105-108
    }
    Fault f{t, tgt, 0, 0, ws};
    if (!pick(tgt == 1 ? wbits : tbits, &f.layer, &f.bit)) continue;
    out.push_back(f);
  }
  return out;
}

}  // namespace bnnb
