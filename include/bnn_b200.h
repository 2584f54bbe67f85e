/* bnn_b200 -- C ABI of the B200 (sm_100a) runtime for BNN-PYNQ's CNV / LFC networks.
 *
 * Two layers of entry points, both plain C (pointers and sizes only):
 *
 * (1) The six per-network symbols that bnn/bnn.py binds through cffi (cdef at bnn/bnn.py:69-76).  They are exported,
 *     with exactly these names, by one shared object per network, lib/python_cuda-<network>.so, next to the
 *     reference's python_hw-<network>-<platform>.so and python_sw-<network>-<platform>.so.  They replace the definitions in
 *     bnn/src/network/<network>/sw/main_python.cpp (cnvW1A1: :67,120,141,171,225,229; lfcW1A1: :65,112,135,158,210,214).
 *
 * (2) A handle-based superset (bnnb_*), exported by lib/libbnn_b200.so, which the per-network objects are thin shims over:
 *     in-memory inputs (no temp files), 16-bit scores / raw LFC words, deterministic fault lists, accelerator-memory access
 *     (FoldedMVMemRead/MemSet, rawhls-offload.cpp:76-88), per-layer taps, and device-resident inference for benchmarks.
 *
 * Error behaviour: the reference throws string literals through the extern "C" frame and terminates the process
 * (foldedmv-offload.cpp:321-323, foldedmv-offload.h:373-376).  This library never throws across the boundary: functions
 * return -1 / NULL and bnnb_last_error() holds the message; the six bnn.py symbols also print it to stderr.
 * There is no CPU fallback: without a CUDA device every compute call fails.
 */
#ifndef BNN_B200_H
#define BNN_B200_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* ------------------------------------------------------------------------------------------------------------------
 * (1) bnn.py's cffi surface -- exported by lib/python_cuda-<network>.so
 * ------------------------------------------------------------------------------------------------------------------ */
/* <net>/sw/main_python.cpp:67 -- reads <path>/<L>-<PE>-weights.bin and <L>-<PE>-thres.bin (FoldedMVLoadLayerMem,
 * foldedmv-offload.cpp:315-336), re-packs them on every visible GPU; discards previously injected faults. */
void load_parameters(const char* path);
/* :120 -- first CIFAR record (CNV) / first idx3 image (LFC) of the file.  results may be NULL; CNV: number_class scores,
 * class = first maximum (std::max_element, :138); LFC: 64-entry one-hot at round(log2(word)) (foldedmv-offload.cpp:155-166). */
int inference(const char* path, int results[64], int number_class, float* usecPerImage);
/* :141 -- every record of the file.  Returns new int[N] class ids, or (CNV, enable_detail != 0) new int[N*number_class] scores.
 * CNV class = first score strictly above a running maximum that starts at 0 (foldedmv-offload.h:398-406);
 * LFC class = index of the highest set bit of (word & mask(number_class)), 0 if none (foldedmv-offload.cpp:202-220). */
int* inference_multiple(const char* path, int number_class, int* image_number, float* usecPerImage, int enable_detail);
/* :171 -- as inference_multiple (class ids only) with flip_count random parameter faults scheduled over the batch
 * (faults.h:60-148).  target: <0 any, 0 weights, >0 thresholds; target_layers may be NULL (= all layers). */
int* inference_multiple_with_faults(const char* path, int number_class, int* image_number, float* usecPerImage,
                                    unsigned int flip_count, int word_size, int target, int* target_layers,
                                    unsigned int num_targets);
/* In-memory twins of the two calls above (extension; bnn.py binds them when the library exports them).  They take what the
 * file would hold after parsing -- CNV: image_count 3073-byte CIFAR records (label byte ignored); LFC: image_count x 784
 * pixels, no idx3 header -- so that classify_images / image_to_cifar (bnn/bnn.py:226-242,280-287) need no temp file.
 * image_bytes = the size of the buffer handed over; it must equal image_count rows of THIS network's size or the call is refused
 * (NULL) before anything is read.  Same results, same result ownership (free_results), same fault scheduling as the path-based calls. */
int* inference_buffer(const unsigned char* images, int image_count, long long image_bytes, int number_class, float* usecPerImage,
                      int enable_detail);
int* inference_buffer_with_faults(const unsigned char* images, int image_count, long long image_bytes, int number_class,
                                  float* usecPerImage, unsigned int flip_count, int word_size, int target, int* target_layers,
                                  unsigned int num_targets);
/* :225 */
void free_results(int* result);
/* :229 -- FoldedMVDeinit: releases every device allocation (bnnb_release).  The parameters and injected faults are kept in the host
 * mirror, as the reference keeps them in static memories, and every entry point re-initialises on demand (FoldedMVInit). */
void deinit(void);

/* ------------------------------------------------------------------------------------------------------------------
 * (2) handle-based API -- exported by lib/libbnn_b200.so
 * ------------------------------------------------------------------------------------------------------------------ */
typedef struct bnnb_ctx bnnb_ctx;

/* network: "cnvW1A1" | "cnvW1A2" | "cnvW2A2" | "lfcW1A1" | "lfcW1A2".  NULL for an unknown name. */
bnnb_ctx* bnnb_open(const char* network);
void bnnb_close(bnnb_ctx* c);
const char* bnnb_last_error(bnnb_ctx* c);
/* GPUs the image batch is sharded over (contiguous ranges, parameters replicated, no collective).  Default: all visible
 * devices or the comma list in $BNN_B200_DEVICES.  Must precede the first load/inference. */
int bnnb_set_devices(bnnb_ctx* c, const int* device_ids, int n);
int bnnb_num_devices(bnnb_ctx* c);
int bnnb_load_parameters(bnnb_ctx* c, const char* dir);
/* Frees every device allocation of the context (parameter operands, buffers, streams); parameters and faults stay in the host mirror and
 * the next call rebuilds the device state on demand.  Device memory is otherwise taken lazily: a device is set up the first time a call
 * runs on it, and its buffers are sized for the chunks it is given (a one-image call takes a few MB on one device). */
int bnnb_release(bnnb_ctx* c);
size_t bnnb_device_bytes(bnnb_ctx* c, int slot);   /* bytes currently allocated on device slot `slot` */

/* Host-buffer inference.  records: n x 3073 bytes (label + planar RGB, the CIFAR-10 binary record bnn.py writes, bnn/bnn.py:226-242);
 * pixels: n x 784 bytes (idx3 payload).  Output pointers may be NULL.  scores: n x 64 int16 (what BlackBoxJam writes per image);
 * words: n x uint64.  classes use the multi-image rules above.  usec_per_image: device time of the inference kernels only. */
int bnnb_classify_cifar(bnnb_ctx* c, const uint8_t* records, int n, int number_class, int* classes, int16_t* scores, float* usec_per_image);
int bnnb_classify_mnist(bnnb_ctx* c, const uint8_t* pixels, int n, int number_class, int* classes, uint64_t* words, float* usec_per_image);

/* The same call on inputs the caller prepared itself, exactly as BlackBoxJam reads them (bnnb_in_bytes per image): LFC 104 bytes =
 * 832 bits, bit k of the image = (pixel k >= 128), LSB first, what binarizeAndPack leaves (foldedmv-offload.cpp:82-98) - the reference
 * packs on the host too; CNV 3 072 int8 bytes HWC, what quantiseAndPack<8,1> leaves (foldedmv-offload.h:129-144).  No preparation kernel;
 * for LFC 7.5x fewer bytes cross PCIe.  scores (CNV) / words (LFC): whichever applies, the other is ignored; both may be NULL. */
int bnnb_classify_packed(bnnb_ctx* c, const uint8_t* packed, int n, int number_class, int* classes, int16_t* scores, uint64_t* words,
                         float* usec_per_image);

/* Deterministic fault list: nf x 5 uint32 {time, target (1 weights | 2 thresholds), layer, bit, word_size}, the arguments of the
 * reference's inject_fault (foldedmv-offload.h:169-214); a fault with time i is applied before image i, cumulatively (faults.h:138-146).
 * word_size is used as given (the reference's bool collapse, faults.h:42, is applied by bnnb_schedule_faults only). */
int bnnb_classify_with_fault_list(bnnb_ctx* c, const uint8_t* input, int n, int number_class, const uint32_t* faults, int nf,
                                  int* classes, float* usec_per_image);
/* Draws the reference's random schedule (faults.h:60-94,123-133; topology.h:108-160) into out (cap x 5 uint32); returns the count.
 * seed < 0: std::random_device, as the reference. */
int bnnb_schedule_faults(bnnb_ctx* c, unsigned num_images, unsigned flip_count, int word_size, int target, const int* target_layers,
                         unsigned num_targets, long long seed, uint32_t* out, int cap);
/* One fault now (persists until the next load). */
int bnnb_inject_fault(bnnb_ctx* c, unsigned target, unsigned layer, unsigned bit, unsigned word_size);
/* FoldedMVMemRead / FoldedMVMemSet (rawhls-offload.cpp:76-88 -> DoMemRead/DoMemInit, cnvW1A1/hw/top.cpp:78-178):
 * target_layer = 2*L (weights) or 2*L+1 (thresholds). */
uint64_t bnnb_mem_read(bnnb_ctx* c, unsigned target_layer, unsigned mem, unsigned ind, unsigned thresh);
int bnnb_mem_set(bnnb_ctx* c, unsigned target_layer, unsigned mem, unsigned ind, unsigned thresh, uint64_t val);
unsigned bnnb_weight_layer_bits(bnnb_ctx* c, int layer);   /* topology.h:55-62 */
unsigned bnnb_thresh_layer_bits(bnnb_ctx* c, int layer);   /* topology.h:63-70 */

/* Per-layer taps (parity tests) ON THE SELECTED CORE (bnnb_set_core): per image, concatenated over layers: int32 accumulators
 * [ofm_dim^2][N] before the threshold (the reference's accumulator values, mvau.hpp:110,166-169), and int8 layer outputs
 * [out_dim^2][N] after the layer's max-pool (values 0/1, -1/0/+1, or the low byte of the score).  Cores 1 / 2 run the TAP
 * instantiations of the tensor-core kernels that ship (same source + extra stores) with the production activation formats. */
size_t bnnb_tap_acc_size(bnnb_ctx* c);
size_t bnnb_tap_act_size(bnnb_ctx* c);
int bnnb_tap(bnnb_ctx* c, const uint8_t* input, int n, int32_t* acc, int8_t* act);

/* Device-resident inference (the reference's timed region: BlackBoxJam on inputs already in accelerator memory,
 * foldedmv-offload.h:389-392).  d_in: CNV int8 HWC 3072 B/img, LFC 104 B/img; d_out: CNV 64 x int16, LFC uint64 per image;
 * d_classes: optional int32 (CNV).  All device pointers on device slot `slot`; stream: a cudaStream_t (NULL = the library's).
 * Asynchronous: call bnnb_sync or synchronise the stream.  bnnb_last_launches: kernels enqueued by the last call. */
int bnnb_infer_device(bnnb_ctx* c, int slot, const void* d_in, int n, void* d_out, int* d_classes, int number_class, void* stream);
int bnnb_prep_device(bnnb_ctx* c, int slot, const void* d_raw, int n, void* d_in, void* stream);
/* bnnb_infer_device that allocates NOTHING.  No counterpart in the reference: added for callers
 * that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same arguments, same
 * results, same asynchrony as bnnb_infer_device; the inter-layer activation scratch is a static device array of 32 MiB that every GPU
 * which loads the library holds until the process exits, so:
 *   - n is limited per call by that array (two buffers of 16 MiB, each holding the widest inter-layer tensor the compute core in use
 *     writes to global memory): 1 792 images for cnvW1A1 / cnvW1A2 / cnvW2A2, 32 768 for lfcW1A1 / lfcW1A2 on the default core; more is refused
 *     (-1) with the limit in bnnb_last_error;
 *   - the device must already hold the parameters: make one ordinary call (bnnb_infer_device / bnnb_classify_*) on that device
 *     after bnnb_load_parameters, outside any capture - the upload allocates, this call never does (no cudaMalloc / cudaFree, no
 *     allocating library routine; per-layer profiling must be off).  Its first call on a device also looks the array's address up
 *     (cudaGetSymbolAddress: no allocation): make that first call outside graph capture too;
 *   - every call shares the array and finds the previous call's contents; nothing in it needs clearing (each kernel writes every
 *     activation it reads).  Calls that overlap on one GPU - two streams, two contexts - race on it: serialise them. */
int bnnb_infer_device_static(bnnb_ctx* c, int slot, const void* d_in, int n, void* d_out, int* d_classes, int number_class, void* stream);  /* records/pixels -> layer-0 input */
/* Waits for the device and fails if a kernel gave up waiting on an mbarrier since the last check (results invalid). */
int bnnb_sync(bnnb_ctx* c, int slot);
int bnnb_last_launches(bnnb_ctx* c);
int bnnb_in_bytes(bnnb_ctx* c);    /* per image: 3072 | 104 */
int bnnb_out_bytes(bnnb_ctx* c);   /* per image: 128 | 8 */

/* Compute core per layer: 0 = LOP3+POPC CUDA-core kernels everywhere; 1 = tcgen05 kind::i8 tensor-core kernels for the layers that
 * have an instantiation; 2 (default) = tcgen05 kind::mxf4 (e2m1 operands, unit scales: an exact integer MMA at twice the int8
 * rate) where instantiated, else kind::i8, else LOP3+POPC.  $BNN_B200_CORE=popc|i8|f4 sets the default. */
int bnnb_set_core(bnnb_ctx* c, int core);

/* Per-layer device timing (CUDA events on the launching stream around every layer kernel while enabled).
 * collect: ms[10] / launches[10] = totals since the last collect for layers 0..8 and (index 9) the class-decode kernel. */
int bnnb_profile_enable(bnnb_ctx* c, int on);
int bnnb_profile_collect(bnnb_ctx* c, float* ms, int* launches);

/* Integer-pipe / tensor-pipe micro-benchmarks on device slot 0 (the roofline denominators of DESIGN.md):
 * which: 0 = POPC (returns 32-bit popcounts per second), 1 = mma.sync.m16n8k32.s8 (returns int8 MACs per second),
 * 2 = LOP3 (returns 32-bit logic ops per second); 3 shf.l.wrap, 4 prmt, 5 mad.hi, 6 mad.lo, 7 shr, 8 min.f32 (thread-instructions
 * per second: what one epilogue instruction per accumulator column costs).  clock_mhz (optional): SM clock implied by clock64(). */
double bnnb_microbench(int which, double* clock_mhz);

/* Pinned host memory for full-rate asynchronous copies. */
void* bnnb_alloc_host(size_t bytes);
void bnnb_free_host(void* p);

#ifdef __cplusplus
}
#endif
#endif
