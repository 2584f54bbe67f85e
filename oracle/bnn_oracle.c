/* TEST INFRASTRUCTURE ONLY -- never linked, imported or executed by the product path.
 *
 * CPU restatement of the BNN-PYNQ software runtime's hot path (BlackBoxJam -> DoCompute)
 * for cnvW1A1 / cnvW1A2 / cnvW2A2 / lfcW1A1 / lfcW1A2, written from the reference's
 * algorithm, in "logical" (unfolded) int8 arithmetic so that it shares no formula with
 * the CUDA kernels (which use bit-plane popcounts / int8 MMA on re-packed parameters).
 *
 * Parity status: PINNED.  tests/test_oracle_golden.py checks this file against
 * (a) the reference's own known-answer tests (deer.bin -> 4, 3.image-idx3-ubyte -> 3;
 * make-hw.sh:123-129, tests/test_qnn.py), (b) outputs of the UNMODIFIED reference sources
 * compiled here by oracle/build_ref.sh (oracle/_ref/python_sw-<net>.so), recorded as fixtures
 * under tests/golden/ref_<net>.npz by tests/golden/make_golden.py, and (c) EVERY LAYER's
 * accumulators and outputs as the reference's own templates produce them when DoCompute's call
 * list is replayed layer by layer (oracle/tap.cpp -> oracle/_ref/tap-<net>.so, fixtures
 * tests/golden/taps_<net>.npz by tests/golden/make_golden_taps.py).
 *
 * Every function cites the reference file:line it follows (paths relative to
 * /root/reference/bnn/src/).
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <stdio.h>
#include <math.h>

enum { K_CONV = 0, K_FC = 1 };
enum { IN_FIX8 = 0,   /* Slice<ap_fixed<8,1>>  (CNV layer 0)                         */
       IN_XNOR = 1,   /* Recast<XnorMul>: acc = #(w_bit == x_bit)                    */
       IN_BIN = 2,    /* Recast<Binary> input: x in {-1,+1}  (lfcW1A2 layer 0)       */
       IN_TERN = 3 }; /* Slice<ap_int<2>>: x in {-1,0,+1}                            */
enum { OUT_BIT = 0,   /* ThresholdsActivation<..,1,..,ap_uint<1>,0>                  */
       OUT_TERN = 1,  /* ThresholdsActivation<..,2,..,ap_int<2>,-1>                  */
       OUT_SCORE = 2 };/* PassThroughActivation<ap_uint<16>>                         */

typedef struct {
  int kind, ifm_ch, ifm_dim, ofm_ch, ofm_dim; /* conv geometry (FC: dims = 1)         */
  int mw, mh;                                 /* MatrixW (K), MatrixH (N)             */
  int pe, simd, wbits;                        /* folding + weight precision (WPI)     */
  int nthr;                                   /* thresholds per row given to the loader */
  int api;                                    /* L*_API of config.h (fault topology)  */
  int thr_bits;                               /* 24 (CNV L0), 16, or 0 (CNV L8) for the fault topology */
  int in_mode, out_mode, pool;
  /* derived */
  int wmem, tmem;
  uint64_t *w;  /* [pe][wmem]    folded weight memory, as m_weights (weights.hpp:68-100,159-199) */
  int32_t *t;   /* [pe][tmem][nthr] raw thresholds (int16 or 24-bit fixed raw), activations.hpp:126-148 */
  int8_t *wv;   /* [mh][mw] logical weight values (unfolded lazily)                   */
} layer_t;

typedef struct bo_net {
  char name[16];
  int is_cnv, nlayers, dirty;
  layer_t L[9];
} bo_net;

static void mk_conv(layer_t *l, int ic, int id, int oc, int od, int pe, int simd, int wb, int nthr, int api,
                    int tb, int im, int om, int pool) {
  memset(l, 0, sizeof *l);
  l->kind = K_CONV; l->ifm_ch = ic; l->ifm_dim = id; l->ofm_ch = oc; l->ofm_dim = od;
  l->mw = 9 * ic; l->mh = oc; /* convlayer.h:85-121: MatrixW = K*K*IFM_CH, MatrixH = OFM_CH */
  l->pe = pe; l->simd = simd; l->wbits = wb; l->nthr = nthr; l->api = api; l->thr_bits = tb;
  l->in_mode = im; l->out_mode = om; l->pool = pool;
}
static void mk_fc(layer_t *l, int mw, int mh, int pe, int simd, int wb, int nthr, int api, int tb, int im, int om) {
  memset(l, 0, sizeof *l);
  l->kind = K_FC; l->ifm_ch = mw; l->ifm_dim = 1; l->ofm_ch = mh; l->ofm_dim = 1;
  l->mw = mw; l->mh = mh; l->pe = pe; l->simd = simd; l->wbits = wb; l->nthr = nthr; l->api = api; l->thr_bits = tb;
  l->in_mode = im; l->out_mode = om; l->pool = 0;
}

/* Network tables: network/<net>/hw/config.h (folding) + hw/top.cpp (datatypes per layer). */
bo_net *bo_open(const char *name) {
  bo_net *n = (bo_net *)calloc(1, sizeof *n);
  snprintf(n->name, sizeof n->name, "%s", name);
  if (!strncmp(name, "cnv", 3)) {
    int w2 = !strcmp(name, "cnvW2A2"), a2 = strcmp(name, "cnvW1A1") != 0;
    if (strcmp(name, "cnvW1A1") && strcmp(name, "cnvW1A2") && !w2) { free(n); return NULL; }
    n->is_cnv = 1; n->nlayers = 9;
    int wb = w2 ? 2 : 1, nt = a2 ? 2 : 1, om = a2 ? OUT_TERN : OUT_BIT, im = a2 ? IN_TERN : IN_XNOR;
    /* cnvW1A1/hw/config.h:18-202, cnvW2A2/hw/config.h (PE/SIMD) */
    static const int pe1[9] = {16, 32, 16, 16, 4, 1, 1, 1, 4}, si1[9] = {3, 32, 32, 32, 32, 32, 4, 8, 1};
    static const int pe2[9] = {8, 16, 8, 8, 4, 1, 1, 2, 4}, si2[9] = {3, 16, 16, 16, 8, 8, 2, 2, 1};
    const int *pe = w2 ? pe2 : pe1, *si = w2 ? si2 : si1;
    mk_conv(&n->L[0], 3, 32, 64, 30, pe[0], si[0], wb, nt, nt, 24, IN_FIX8, om, 0);
    mk_conv(&n->L[1], 64, 30, 64, 28, pe[1], si[1], wb, nt, nt, 16, im, om, 1);
    mk_conv(&n->L[2], 64, 14, 128, 12, pe[2], si[2], wb, nt, nt, 16, im, om, 0);
    mk_conv(&n->L[3], 128, 12, 128, 10, pe[3], si[3], wb, nt, nt, 16, im, om, 1);
    mk_conv(&n->L[4], 128, 5, 256, 3, pe[4], si[4], wb, nt, nt, 16, im, om, 0);
    mk_conv(&n->L[5], 256, 3, 256, 1, pe[5], si[5], wb, nt, nt, 16, im, om, 0);
    mk_fc(&n->L[6], 256, 512, pe[6], si[6], wb, nt, nt, 16, im, om);
    mk_fc(&n->L[7], 512, 512, pe[7], si[7], wb, nt, nt, 16, im, om);
    /* L8: loader is called with 0 thresholds (cnvW1A1/sw/main_python.cpp:81); topology API = L8_API
       (16 for W1A1, 1 for W1A2/W2A2) with activation_element_bits 0 (main_python.cpp:92-102). */
    mk_fc(&n->L[8], 512, 64, pe[8], si[8], wb, 0, a2 ? 1 : 16, 0, im, OUT_SCORE);
  } else if (!strcmp(name, "lfcW1A1") || !strcmp(name, "lfcW1A2")) {
    int a2 = !strcmp(name, "lfcW1A2");
    n->is_cnv = 0; n->nlayers = 4;
    int nt = a2 ? 2 : 1, om = a2 ? OUT_TERN : OUT_BIT, im = a2 ? IN_TERN : IN_XNOR;
    /* lfcW1A1/hw/config.h:17-83; lfcW1A2/hw/top.cpp:156-163 (datatypes) */
    mk_fc(&n->L[0], 832, 1024, 32, 64, 1, nt, nt, 16, a2 ? IN_BIN : IN_XNOR, om);
    mk_fc(&n->L[1], 1024, 1024, 64, 32, 1, nt, nt, 16, im, om);
    mk_fc(&n->L[2], 1024, 1024, 32, 64, 1, nt, nt, 16, im, om);
    mk_fc(&n->L[3], 1024, 64, 16, 8, 1, 1, 1, 16, im, OUT_BIT);
  } else { free(n); return NULL; }
  for (int i = 0; i < n->nlayers; i++) {
    layer_t *l = &n->L[i];
    l->wmem = (l->mw / l->simd) * (l->mh / l->pe); /* mvau.hpp:100,104: SF*NF */
    l->tmem = l->mh / l->pe;
    l->w = (uint64_t *)calloc((size_t)l->pe * l->wmem, 8);
    l->t = (int32_t *)calloc((size_t)l->pe * l->tmem * (l->nthr ? l->nthr : 1), 4);
    l->wv = (int8_t *)calloc((size_t)l->mh * l->mw, 1);
  }
  n->dirty = 1;
  return n;
}
void bo_close(bo_net *n) {
  if (!n) return;
  for (int i = 0; i < n->nlayers; i++) { free(n->L[i].w); free(n->L[i].t); free(n->L[i].wv); }
  free(n);
}
int bo_num_layers(const bo_net *n) { return n->nlayers; }
/* geometry query for the tests: out[0..7] = mw, mh, ofm_dim, pool, out_mode, in_mode, pe, simd */
void bo_layer_info(const bo_net *n, int i, int *out) {
  const layer_t *l = &n->L[i];
  out[0] = l->mw; out[1] = l->mh; out[2] = l->ofm_dim; out[3] = l->pool; out[4] = l->out_mode; out[5] = l->in_mode;
  out[6] = l->pe; out[7] = l->simd; out[8] = l->wmem; out[9] = l->tmem; out[10] = l->nthr; out[11] = l->wbits;
}

static int64_t sext(uint64_t v, int bits) { uint64_t m = 1ull << (bits - 1); v &= (m << 1) - 1; return (int64_t)((v ^ m) - m); }

/* DoMemInit: cnvW1A1/hw/top.cpp:78-135, lfcW1A1/hw/top.cpp (same shape).
 * targetLayer = 2*L (weights) or 2*L+1 (thresholds). */
void bo_mem_set(bo_net *n, unsigned tl, unsigned mem, unsigned ind, unsigned thresh, uint64_t val) {
  unsigned li = tl >> 1;
  if ((int)li >= n->nlayers) return;
  layer_t *l = &n->L[li];
  if (!(tl & 1)) {
    int bits = l->simd * l->wbits; /* ap_uint<SIMD*wb> = ap_uint<64>: truncation (top.cpp:81) */
    l->w[(size_t)mem * l->wmem + ind] = bits >= 64 ? val : (val & ((1ull << bits) - 1));
    n->dirty = 1;
  } else {
    if (l->nthr == 0) return; /* top.cpp:131-133: no threshold memory behind PassThroughActivation */
    int32_t *p = &l->t[((size_t)mem * l->tmem + ind) * l->nthr + thresh];
    if (n->is_cnv && li == 0) *p = (int32_t)sext(val, 24); /* bit-cast to ap_fixed<64,56> -> ap_fixed<24,16> (top.cpp:84) */
    else *p = (int32_t)sext(val, 16);                      /* ap_int<16> = ap_uint<64> (top.cpp:90) */
  }
}
/* DoMemRead: cnvW1A1/hw/top.cpp:137-178 */
uint64_t bo_mem_read(const bo_net *n, unsigned tl, unsigned mem, unsigned ind, unsigned thresh) {
  unsigned li = tl >> 1;
  if ((int)li >= n->nlayers) return 0;
  const layer_t *l = &n->L[li];
  if (!(tl & 1)) return l->w[(size_t)mem * l->wmem + ind];
  if (l->nthr == 0) return 0;
  int32_t r = l->t[((size_t)mem * l->tmem + ind) * l->nthr + thresh];
  if (n->is_cnv && li == 0) { /* static_cast<ap_uint<64>>(ap_fixed<24,16>): VALUE cast (top.cpp:142), C-native = toward zero */
    int64_t q = r >> 8;
    if (r < 0 && (r & 0xff)) ++q;
    return (uint64_t)q;
  }
  return (uint64_t)(int64_t)r; /* ap_int<16> -> ap_uint<64>: sign extension */
}

/* FoldedMVLoadLayerMem / LoadWeightPE / LoadThreshPE: library/host/foldedmv-offload.cpp:281-336.
 * Short files load zeros ("ExtMemWord e = 0; wf.read(...)" fails silently, :283-284). */
int bo_load(bo_net *n, const char *dir) {
  char path[1024];
  for (int li = 0; li < n->nlayers; li++) {
    layer_t *l = &n->L[li];
    for (int pe = 0; pe < l->pe; pe++) {
      snprintf(path, sizeof path, "%s/%d-%d-weights.bin", dir, li, pe);
      FILE *f = fopen(path, "rb");
      if (!f) return -1; /* reference: throw "Could not open file" (:321-323) */
      for (int line = 0; line < l->wmem; line++) {
        uint64_t e = 0;
        if (fread(&e, 8, 1, f) != 1) e = 0;
        bo_mem_set(n, 2 * li, pe, line, 0, e);
      }
      fclose(f);
      if (l->nthr > 0) {
        snprintf(path, sizeof path, "%s/%d-%d-thres.bin", dir, li, pe);
        f = fopen(path, "rb");
        if (!f) return -1;
        for (int line = 0; line < l->tmem; line++)
          for (int i = 0; i < l->nthr; i++) {
            uint64_t e = 0;
            if (fread(&e, 8, 1, f) != 1) e = 0;
            bo_mem_set(n, 2 * li + 1, pe, line, i, e);
          }
        fclose(f);
      }
    }
  }
  return 0;
}

/* inject_fault + inject_fault_impl: library/host/foldedmv-offload.h:146-214, for topologies with one
 * module per layer (cnvW1A1/sw/main_python.cpp:92-102).  target: 1 = weights, 2 = thresholds ("Activations").
 * `bit` is the n-th bit of the layer's weight (or threshold) space.  word_size is what the injector finally
 * receives (the reference's bool quirk, faults.h:42 vs :55, is applied by the caller). */
void bo_inject(bo_net *n, int target, unsigned layer, unsigned bit, unsigned word_size) {
  const layer_t *l = &n->L[layer];
  unsigned mem, ind, thresh, tl;
  if (target == 1) {
    unsigned es = l->simd * l->wbits, el = bit / es;
    thresh = 0; ind = el % l->wmem; mem = (el / l->wmem) % l->pe; bit %= es; tl = layer * 2;
  } else {
    unsigned es = l->thr_bits;
    if (!es) return; /* the reference would divide by zero; such a layer has 0 threshold bits and is never drawn */
    unsigned el = bit / es;
    thresh = el % l->api; ind = (el / l->api) % l->tmem; mem = ((el / l->api) / l->tmem) % l->pe; bit %= es; tl = layer * 2 + 1;
  }
  if (word_size == 0) return; /* reference: bit_pos/word_size divides by zero; treated as "flip nothing" */
  uint64_t val = bo_mem_read(n, tl, mem, ind, thresh), mask = 0;
  for (unsigned i = 0; i < word_size; i++) mask |= 1ull << i;
  val ^= mask << ((bit / word_size) * word_size);
  bo_mem_set(n, tl, mem, ind, thresh, val);
}

/* Unfold the folded memories to logical W[row][col] (mvau.hpp:115,143-152,166-169;
 * weights.hpp:89-98,180-191; interpret.hpp:84-90): row = nf*PE + pe, col = sf*SIMD + s, tile = nf*SF + sf. */
static void unfold(bo_net *n) {
  for (int li = 0; li < n->nlayers; li++) {
    layer_t *l = &n->L[li];
    int SF = l->mw / l->simd;
    for (int r = 0; r < l->mh; r++) {
      int pe = r % l->pe, nf = r / l->pe;
      for (int c = 0; c < l->mw; c++) {
        int sf = c / l->simd, s = c % l->simd;
        uint64_t word = l->w[(size_t)pe * l->wmem + nf * SF + sf];
        int v;
        if (l->wbits == 1) v = ((word >> s) & 1) ? 1 : -1;
        else v = (int)sext(word >> (2 * s), 2);
        l->wv[(size_t)r * l->mw + c] = (int8_t)v;
      }
    }
  }
  n->dirty = 0;
}

static inline int dot8(const int8_t *a, const int8_t *b, int k) {
  int s = 0;
  for (int i = 0; i < k; i++) s += (int)a[i] * (int)b[i];
  return s;
}

/* One MVAU layer on one image.  in: [ifm_dim^2][ifm_ch] int8 VALUES (NHWC); for IN_XNOR the values are bits {0,1}.
 * acc_out (optional): [ofm_dim^2][mh] int32; act (pre-pool): [ofm_dim^2][mh] int8 values (or int16 scores in sc). */
static void run_layer(const layer_t *l, const int8_t *in, int32_t *acc_out, int8_t *act, int16_t *sc) {
  int od = l->ofm_dim, id = l->ifm_dim, C = l->ifm_ch, K = l->mw, N = l->mh;
  int8_t *x = (int8_t *)malloc(K);
  for (int oy = 0; oy < od; oy++)
    for (int ox = 0; ox < od; ox++) {
      /* ConvolutionInputGenerator: slidingwindow.h:133-155 -> column (ky*3+kx)*C + c, stride 1, no padding */
      if (l->kind == K_CONV)
        for (int ky = 0; ky < 3; ky++) memcpy(x + ky * 3 * C, in + ((size_t)(oy + ky) * id + ox) * C, 3 * C);
      else memcpy(x, in, K);
      if (l->in_mode == IN_XNOR) for (int i = 0; i < K; i++) x[i] = x[i] ? 1 : -1;
      for (int r = 0; r < N; r++) {
        int d = dot8(l->wv + (size_t)r * K, x, K), acc;
        switch (l->in_mode) {
          case IN_FIX8: acc = 2 * d; break;       /* ap_fixed<8,1> raw q -> ap_fixed<24,16> raw 2q (top.cpp:61,214) */
          case IN_XNOR: acc = (K + d) / 2; break; /* XnorMul: #equal bits (interpret.hpp:57-69) */
          default: acc = d; break;                /* Binary*x = +-x (interpret.hpp:86-103); Identity: w*x */
        }
        size_t o = ((size_t)oy * od + ox) * N + r;
        if (acc_out) acc_out[o] = acc;
        if (l->out_mode == OUT_SCORE) { sc[o] = (int16_t)(uint16_t)(acc & 0xffff); continue; } /* activations.hpp:85-92 + foldedmv-offload.h:121-127 */
        int pe = r % l->pe, nf = r / l->pe;
        const int32_t *t = &l->t[((size_t)pe * l->tmem + nf) * l->nthr];
        /* ThresholdsActivation::activate: result = ActVal + sum_i (thr[i] < accu), strict (activations.hpp:139-147) */
        int a16 = l->thr_bits == 24 ? acc : (int16_t)acc;
        if (l->out_mode == OUT_BIT) act[o] = (int8_t)(t[0] < a16);
        else act[o] = (int8_t)(-1 + (t[0] < a16) + (t[1] < a16));
      }
    }
  free(x);
}

/* StreamingMaxPool_Batch (maxpool.h:62-112: OR) / StreamingMaxPool_Precision_Batch<..,ap_int<2>,-1> (:129-208: signed max, init -1) */
static void pool2(const int8_t *in, int dim, int C, int8_t *out) {
  int od = dim / 2;
  for (int y = 0; y < od; y++)
    for (int x = 0; x < od; x++)
      for (int c = 0; c < C; c++) {
        int m = -1;
        for (int dy = 0; dy < 2; dy++)
          for (int dx = 0; dx < 2; dx++) {
            int v = in[((size_t)(2 * y + dy) * dim + 2 * x + dx) * C + c];
            if (v > m) m = v;
          }
        out[((size_t)y * od + x) * C + c] = (int8_t)m;
      }
}

/* Offsets of each layer's outputs inside the tap buffers (per image). */
static size_t tap_rows(const layer_t *l) { return (size_t)l->ofm_dim * l->ofm_dim; }
size_t bo_tap_acc_size(const bo_net *n) { size_t s = 0; for (int i = 0; i < n->nlayers; i++) s += tap_rows(&n->L[i]) * n->L[i].mh; return s; }
size_t bo_tap_act_size(const bo_net *n) {
  size_t s = 0;
  for (int i = 0; i < n->nlayers; i++) { const layer_t *l = &n->L[i]; int d = l->pool ? l->ofm_dim / 2 : l->ofm_dim; s += (size_t)d * d * l->mh; }
  return s;
}

/* DoCompute for ONE image given the layer-0 input values.  scores: 64 x int16 (CNV) ; word: 64-bit output (LFC).
 * tap_acc: concatenation over layers of [ofm^2][mh] int32 accumulators (pre-threshold);
 * tap_act: concatenation over layers of the layer's output activations AFTER its max-pool (int8 values; scores
 *          for the last CNV layer are stored wrapped to int8 pairs? no: last CNV layer contributes its 64 scores as
 *          int8-truncated values and should be read from `scores` instead). */
static void compute_one(const bo_net *n, const int8_t *x0, int16_t *scores, uint64_t *word, int32_t *tap_acc, int8_t *tap_act) {
  int8_t *cur = NULL, *act, *pooled;
  const int8_t *in = x0;
  for (int li = 0; li < n->nlayers; li++) {
    const layer_t *l = &n->L[li];
    size_t rows = tap_rows(l), sz = rows * l->mh;
    act = (int8_t *)calloc(sz, 1);
    int16_t sc[64];
    run_layer(l, in, tap_acc, act, sc);
    if (tap_acc) tap_acc += sz;
    if (l->out_mode == OUT_SCORE) { memcpy(scores, sc, 128); for (int i = 0; i < 64; i++) act[i] = (int8_t)sc[i]; }
    if (l->pool) {
      int d = l->ofm_dim / 2;
      pooled = (int8_t *)malloc((size_t)d * d * l->mh);
      pool2(act, l->ofm_dim, l->mh, pooled);
      free(act); act = pooled; sz = (size_t)d * d * l->mh;
    }
    if (tap_act) { memcpy(tap_act, act, sz); tap_act += sz; }
    free(cur); cur = act; in = cur;
  }
  if (!n->is_cnv) { /* Stream2Mem: bit r of the word = output r of the last layer (LSB first, streamtools.h:255-315) */
    uint64_t w = 0;
    for (int r = 0; r < 64; r++) if (cur[r]) w |= 1ull << r;
    *word = w;
  }
  free(cur);
}

/* CNV input quantisation: tiny_cnn parse_cifar10 scale [-1,1] (cnvW1A1/sw/main_python.cpp:129) then
 * ap_fixed<8,1,AP_RND,AP_SAT> (foldedmv-offload.h:139), planar RGB -> HWC (chaninterleave, :381-385). */
void bo_cifar_lut(int8_t lut[256]) {
  for (int c = 0; c < 256; c++) {
    float f = -1.0f + (1.0f - -1.0f) * (float)c / 255; /* min + (max-min)*c/255 in float */
    double r = floor(ldexp((double)f, 7) + 0.5);
    if (r > 127) r = 127;
    if (r < -128) r = -128;
    lut[c] = (int8_t)r;
  }
}
/* records: n x 3073 bytes (label + 1024 R + 1024 G + 1024 B).  q: n x 3072 int8, HWC. */
void bo_cifar_quantise(const uint8_t *records, int n, int8_t *q) {
  int8_t lut[256];
  bo_cifar_lut(lut);
  for (int i = 0; i < n; i++) {
    const uint8_t *p = records + (size_t)i * 3073 + 1;
    for (int px = 0; px < 1024; px++)
      for (int c = 0; c < 3; c++) q[(size_t)i * 3072 + px * 3 + c] = lut[p[c * 1024 + px]];
  }
}

/* BlackBoxJam(numReps=n) for CNV on pre-quantised inputs.  scores: n x 64 int16.  Taps optional (n x tap sizes). */
int bo_cnv_compute(bo_net *n, const int8_t *q, int count, int16_t *scores, int32_t *tap_acc, int8_t *tap_act) {
  if (!n->is_cnv) return -1;
  if (n->dirty) unfold(n);
  size_t sa = bo_tap_acc_size(n), st = bo_tap_act_size(n);
#pragma omp parallel for schedule(dynamic, 1)
  for (int i = 0; i < count; i++)
    compute_one(n, q + (size_t)i * 3072, scores + (size_t)i * 64, NULL, tap_acc ? tap_acc + (size_t)i * sa : NULL,
                tap_act ? tap_act + (size_t)i * st : NULL);
  return 0;
}

/* LFC: pixels n x 784 bytes.  binarizeAndPack (foldedmv-offload.cpp:82-98): bit = (scaled >= 0) <=> byte >= 128
 * with tiny_cnn's (c/255)*2-1 float scaling; 784 -> 832 columns, pad bits 0.  words: n x uint64. */
int bo_lfc_compute(bo_net *n, const uint8_t *pixels, int count, uint64_t *words, int32_t *tap_acc, int8_t *tap_act) {
  if (n->is_cnv) return -1;
  if (n->dirty) unfold(n);
  size_t sa = bo_tap_acc_size(n), st = bo_tap_act_size(n);
  int l0bin = n->L[0].in_mode == IN_BIN;
#pragma omp parallel for schedule(dynamic, 16)
  for (int i = 0; i < count; i++) {
    int8_t x[832];
    for (int k = 0; k < 832; k++) {
      int bit = 0;
      if (k < 784) { float f = ((float)pixels[(size_t)i * 784 + k] / 255.0f) * (1.0f - -1.0f) + -1.0f; bit = f >= 0; }
      x[k] = l0bin ? (bit ? 1 : -1) : bit; /* Recast<Binary>: bit 0 -> -1 (interpret.hpp:84) */
    }
    compute_one(n, x, NULL, &words[i], tap_acc ? tap_acc + (size_t)i * sa : NULL, tap_act ? tap_act + (size_t)i * st : NULL);
  }
  return 0;
}

/* Class decode.  CNV multi-image rule (foldedmv-offload.h:398-406): first score strictly above a running max that starts at 0. */
int bo_cnv_class_multi(const int16_t *sc, int ncls) {
  int mi = 0; int16_t mv = 0;
  for (int j = 0; j < ncls; j++) if (sc[j] > mv) { mv = sc[j]; mi = j; }
  return mi;
}
/* CNV single-image rule: std::max_element, first maximum (cnvW1A1/sw/main_python.cpp:138). */
int bo_cnv_class_single(const int16_t *sc, int ncls) {
  int mi = 0;
  for (int j = 1; j < ncls; j++) if (sc[j] > sc[mi]) mi = j;
  return mi;
}
/* LFC multi-image rule (foldedmv-offload.cpp:202-220): (unsigned) log2(word & mask), 0 if none. */
int bo_lfc_class_multi(uint64_t w, int ncls) {
  uint64_t m = 0xFFFFFFFFFFFFFFFFull >> (64 - ncls);
  w &= m;
  if (!w) return 0;
  return (int)(unsigned)log2((double)w);
}
/* LFC single-image rule (foldedmv-offload.cpp:155-166 + lfcW1A1/sw/main_python.cpp:132): one-hot at round(log2(word)). */
int bo_lfc_class_single(uint64_t w, int ncls, int onehot[64]) {
  uint64_t m = 0xFFFFFFFFFFFFFFFFull >> (64 - ncls);
  w &= m;
  double r = w ? round(log2((double)w)) : 0;
  int best = -1;
  for (unsigned i = 0; i < 64; i++) { int v = (i == (unsigned)r); if (onehot) onehot[i] = v; if (v && best < 0) best = i; }
  return best < 0 ? 0 : best;
}

/* Fault topology: library/host/topology.h:55-70 (bit counts per layer). */
unsigned bo_weight_layer_bits(const bo_net *n, int l) { const layer_t *L = &n->L[l]; return (unsigned)L->wbits * L->simd * L->pe * L->wmem; }
unsigned bo_thresh_layer_bits(const bo_net *n, int l) { const layer_t *L = &n->L[l]; return (unsigned)L->tmem * L->pe * L->api * L->thr_bits; }

/* make_faulty_classification_func with an explicit fault list (faults.h:115-148): every fault with time i is applied
 * before image i; faults are cumulative.  Images are classified one at a time (numReps = 1) with the multi-image rule.
 * faults: nf x 5 uint32 {time, target(1|2), layer, bit, word_size}; applied in list order within equal times. */
int bo_classify_with_faults(bo_net *n, const void *inputs, int count, int ncls, const uint32_t *faults, int nf, int *classes) {
  for (int i = 0; i < count; i++) {
    for (int f = 0; f < nf; f++)
      if (faults[5 * f] == (uint32_t)i) bo_inject(n, faults[5 * f + 1], faults[5 * f + 2], faults[5 * f + 3], faults[5 * f + 4]);
    if (n->is_cnv) {
      int16_t sc[64];
      bo_cnv_compute(n, (const int8_t *)inputs + (size_t)i * 3072, 1, sc, NULL, NULL);
      classes[i] = bo_cnv_class_multi(sc, ncls);
    } else {
      uint64_t w;
      bo_lfc_compute(n, (const uint8_t *)inputs + (size_t)i * 784, 1, &w, NULL, NULL);
      classes[i] = bo_lfc_class_multi(w, ncls);
    }
  }
  return 0;
}
