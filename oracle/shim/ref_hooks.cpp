// TEST INFRASTRUCTURE ONLY (oracle build).  Deterministic doors into the UNMODIFIED reference
// runtime, compiled into oracle/_ref/python_sw-<net>.so next to the reference sources:
//  * ref_inject_fault: drives the reference's own injector (inject_fault<Topology>, foldedmv-offload.h:169-214)
//    through the RandomFaultArgs object that make_random_fault_args builds (<net>/sw/main_python.cpp:84-118), so the
//    word_size bool quirk (faults.h:42 vs :55) is reproduced by the reference code itself, with a caller-chosen
//    (layer, bit) instead of the random_device draw (faults.h:60-94) that cannot be replayed.
//  * ref_mem_read / ref_mem_set: FoldedMVMemRead / FoldedMVMemSet (rawhls-offload.cpp:76-88).
//  * ref_blackboxjam: BlackBoxJam compute mode on caller-packed buffers (foldedmv-offload.h:391).
#include "tiny_cnn/tiny_cnn.h"
#include "foldedmv-offload.h"
#include "faults.h"
RandomFaultArgs make_random_fault_args(int word_size, int target, int *target_layers, unsigned int num_layers);
extern "C" void ref_inject_fault(int target, int word_size, unsigned layer, unsigned bit) {
  RandomFaultArgs a = make_random_fault_args(word_size, target, nullptr, 0);
  a.injection_func(a.topology, target == 0 ? TargetType::Weights : TargetType::Activations, a.word_size, layer, bit);
}
extern "C" unsigned long long ref_mem_read(unsigned tl, unsigned mem, unsigned ind, unsigned th) {
  FoldedMVInit("hook");
  return FoldedMVMemRead(tl, mem, ind, th);
}
extern "C" void ref_mem_set(unsigned tl, unsigned mem, unsigned ind, unsigned th, unsigned long long v) {
  FoldedMVInit("hook");
  FoldedMVMemSet(tl, mem, ind, th, v);
}
extern "C" void ref_blackboxjam(unsigned long long *in, unsigned long long *out, unsigned count) {
  BlackBoxJam((ap_uint<64> *)in, (ap_uint<64> *)out, 0, 0, 0, 0, 0, 0, count, -1);
}
