#include "netdesc.hpp"
#include <cstring>
#include <cstdio>

namespace bnnb {

static LayerDesc conv(int ic, int id, int oc, int od, int pe, int simd, int wb, int nthr, int api, int tb, int im, int om, int pool) {
  LayerDesc l{};
  l.kind = KIND_CONV; l.ifm_ch = ic; l.ifm_dim = id; l.ofm_ch = oc; l.ofm_dim = od;
  l.K = 9 * ic; l.N = oc; l.pe = pe; l.simd = simd; l.wbits = wb; l.nthr = nthr; l.api = api; l.thr_bits = tb;
  l.in_mode = im; l.out_mode = om; l.pool = pool;
  return l;
}
static LayerDesc fc(int K, int N, int pe, int simd, int wb, int nthr, int api, int tb, int im, int om) {
  LayerDesc l{};
  l.kind = KIND_FC; l.ifm_ch = K; l.ifm_dim = 1; l.ofm_ch = N; l.ofm_dim = 1;
  l.K = K; l.N = N; l.pe = pe; l.simd = simd; l.wbits = wb; l.nthr = nthr; l.api = api; l.thr_bits = tb;
  l.in_mode = im; l.out_mode = om; l.pool = 0;
  return l;
}

bool make_net(const char* name, NetDesc* n) {
  std::memset(n, 0, sizeof *n);
  std::snprintf(n->name, sizeof n->name, "%s", name);
  const bool w1a1 = !std::strcmp(name, "cnvW1A1"), w1a2 = !std::strcmp(name, "cnvW1A2"), w2a2 = !std::strcmp(name, "cnvW2A2");
  if (w1a1 || w1a2 || w2a2) {
    n->is_cnv = 1; n->nlayers = 9; n->in_bytes = 3072; n->out_bytes = 128;
    const int wb = w2a2 ? 2 : 1, nt = w1a1 ? 1 : 2;
    const int om = w1a1 ? OUT_BIT : OUT_TERN, im = w1a1 ? IN_XNOR : IN_TERN;
    // cnvW1A1/hw/config.h:18-202 (same folding for cnvW1A2); cnvW2A2/hw/config.h
    static const int pe1[9] = {16, 32, 16, 16, 4, 1, 1, 1, 4}, si1[9] = {3, 32, 32, 32, 32, 32, 4, 8, 1};
    static const int pe2[9] = {8, 16, 8, 8, 4, 1, 1, 2, 4}, si2[9] = {3, 16, 16, 16, 8, 8, 2, 2, 1};
    const int* pe = w2a2 ? pe2 : pe1; const int* si = w2a2 ? si2 : si1;
    n->L[0] = conv(3, 32, 64, 30, pe[0], si[0], wb, nt, nt, 24, IN_FIX8, om, 0);
    n->L[1] = conv(64, 30, 64, 28, pe[1], si[1], wb, nt, nt, 16, im, om, 1);
    n->L[2] = conv(64, 14, 128, 12, pe[2], si[2], wb, nt, nt, 16, im, om, 0);
    n->L[3] = conv(128, 12, 128, 10, pe[3], si[3], wb, nt, nt, 16, im, om, 1);
    n->L[4] = conv(128, 5, 256, 3, pe[4], si[4], wb, nt, nt, 16, im, om, 0);
    n->L[5] = conv(256, 3, 256, 1, pe[5], si[5], wb, nt, nt, 16, im, om, 0);
    n->L[6] = fc(256, 512, pe[6], si[6], wb, nt, nt, 16, im, om);
    n->L[7] = fc(512, 512, pe[7], si[7], wb, nt, nt, 16, im, om);
    // last layer: loader reads no thresholds (cnvW1A1/sw/main_python.cpp:81); topology API = L8_API, element bits 0 (:92-102)
    n->L[8] = fc(512, 64, pe[8], si[8], wb, 0, w1a1 ? 16 : 1, 0, im, OUT_SCORE);
  } else if (!std::strcmp(name, "lfcW1A1") || !std::strcmp(name, "lfcW1A2")) {
    const bool a2 = !std::strcmp(name, "lfcW1A2");
    n->is_cnv = 0; n->nlayers = 4; n->in_bytes = 104; n->out_bytes = 8;
    const int nt = a2 ? 2 : 1, om = a2 ? OUT_TERN : OUT_BIT, im = a2 ? IN_TERN : IN_XNOR;
    // lfcW1A1/hw/config.h:17-83; datatypes lfcW1A1/hw/top.cpp:156-163, lfcW1A2/hw/top.cpp:156-163
    n->L[0] = fc(832, 1024, 32, 64, 1, nt, nt, 16, a2 ? IN_BIN : IN_XNOR, om);
    n->L[1] = fc(1024, 1024, 64, 32, 1, nt, nt, 16, im, om);
    n->L[2] = fc(1024, 1024, 32, 64, 1, nt, nt, 16, im, om);
    n->L[3] = fc(1024, 64, 16, 8, 1, 1, 1, 16, im, OUT_BIT);
  } else {
    return false;
  }
  for (int i = 0; i < n->nlayers; i++) {
    LayerDesc& l = n->L[i];
    l.wmem = (l.K / l.simd) * (l.N / l.pe);
    l.tmem = l.N / l.pe;
  }
  return true;
}

}  // namespace bnnb
