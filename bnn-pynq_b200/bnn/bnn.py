"""Host-side mirror of BNN-PYNQ's bnn/bnn.py for the CUDA (sm_100a) runtime.

Same classes, methods, argument meaning and printed lines as the reference module (PynqBNN bnn/bnn.py:94-207,
CnvClassifier :210-345, LfcClassifier :348-388, available_params :82-91), with a third runtime string, RUNTIME_CUDA,
next to RUNTIME_HW / RUNTIME_SW.  The library is found as "{runtime}-{network}.so" under lib/ (the reference appends the
board name, bnn/bnn.py:108); pynq and the BOARD variable are not needed for this runtime (bnn/bnn.py:30,55-60 made that
impossible).  cffi dlopen + the same cdef as bnn/bnn.py:69-76.
"""
import io
import os
import tempfile

import cffi
import numpy as np

RUNTIME_HW = "python_hw"
RUNTIME_SW = "python_sw"
RUNTIME_CUDA = "python_cuda"

NETWORK_CNVW1A1 = "cnvW1A1"
NETWORK_CNVW1A2 = "cnvW1A2"
NETWORK_CNVW2A2 = "cnvW2A2"
NETWORK_LFCW1A1 = "lfcW1A1"
NETWORK_LFCW1A2 = "lfcW1A2"
# The fork's hardened variants (bnn/bnn.py:41-53): the names exist so that campaign scripts import cleanly, but RUNTIME_CUDA does not
# build them.  They replicate or interleave the FPGA's parameter BRAMs (TMR majority vote, interleaved weight words) and compute the
# same function as their base network only while no fault is injected, so a faults.py campaign on them is NOT reproduced by the base
# network: asking for one is an explicit error (DESIGN.md "out of scope").
NETWORK_CNVW1A1_INTERLEAVED = "cnvW1A1-interleaved"
NETWORK_CNVW1A1_RESILIENT_INTERLEAVED = "cnvW1A1-resilient-interleaved"
NETWORK_CNVW1A1_TMR = "cnvW1A1-TMR"
NETWORK_CNVW1A2_INTERLEAVED = "cnvW1A2-interleaved"
NETWORK_CNVW1A2_RESILIENT_INTERLEAVED = "cnvW1A2-resilient-interleaved"
NETWORK_CNVW2A2_INTERLEAVED = "cnvW2A2-interleaved"
NETWORK_CNVW2A2_RESILIENT_INTERLEAVED = "cnvW2A2-resilient-interleaved"
NETWORK_CNVW2A2_TMR = "cnvW2A2-TMR"
NETWORK_LFCW1A2_INTERLEAVED = "lfcW1A2-interleaved"
HARDENED_SUFFIXES = ("-TMR", "-interleaved", "-resilient-interleaved")

BNN_ROOT_DIR = os.path.dirname(os.path.realpath(__file__))
BNN_LIB_DIR = os.environ.get("BNN_B200_LIB_DIR", os.path.join(os.path.dirname(BNN_ROOT_DIR), "lib"))
# parameter sets: $BNN_PARAM_DIR, else a params/ directory next to this package (the reference ships bnn/params)
BNN_PARAM_DIR = os.environ.get("BNN_PARAM_DIR", os.path.join(BNN_ROOT_DIR, "params"))

_ffi = cffi.FFI()
_ffi.cdef("""
void load_parameters(const char* path);
int inference(const char* path, int results[64], int number_class, float *usecPerImage);
int* inference_multiple(const char* path, int number_class, int *image_number, float *usecPerImage, int enable_detail);
int* inference_multiple_with_faults(const char* path, int number_class, int *image_number, float *usecPerImage, unsigned int flip_count, int word_size, int target, int* target_layers, unsigned int num_targets);
void free_results(int * result);
void deinit();
""")
# in-memory twins exported by this runtime's libraries only (include/bnn_b200.h); bound lazily so the cdef above stays the reference's
_ffi.cdef("""
int* inference_buffer(const unsigned char* images, int image_count, long long image_bytes, int number_class, float *usecPerImage, int enable_detail);
int* inference_buffer_with_faults(const unsigned char* images, int image_count, long long image_bytes, int number_class, float *usecPerImage, unsigned int flip_count, int word_size, int target, int* target_layers, unsigned int num_targets);
""")

_libraries = {}


def set_param_dir(path):
    """Point the package at a directory laid out like the reference's bnn/params (<dataset>/<network>/...)."""
    global BNN_PARAM_DIR
    BNN_PARAM_DIR = path


def available_params(network):
    """Datasets that ship parameters for `network` (bnn/bnn.py:82-91)."""
    ret = []
    if not os.path.isdir(BNN_PARAM_DIR):
        return ret
    for d in sorted(os.listdir(BNN_PARAM_DIR)):
        if os.path.isdir(os.path.join(BNN_PARAM_DIR, d)) and network in os.listdir(os.path.join(BNN_PARAM_DIR, d)):
            ret.append(d)
    return ret


def _quiet():
    return os.environ.get("BNN_B200_QUIET", "0") not in ("", "0")


def _say(msg):
    if not _quiet():
        print(msg)


class PynqBNN:
    """cffi door to one per-network shared object (bnn/bnn.py:94-207)."""

    def __init__(self, runtime, network, load_overlay=True):
        if runtime in (RUNTIME_HW, RUNTIME_SW):
            raise RuntimeError("runtime %r needs a PYNQ board; this package provides %r" % (runtime, RUNTIME_CUDA))
        if network.endswith(HARDENED_SUFFIXES):
            raise NotImplementedError("hardened variant %r is not supported by %s: it models FPGA parameter-memory redundancy "
                                      "(TMR vote / interleaved words); use the base network %r for fault-free inference"
                                      % (network, RUNTIME_CUDA, network.split("-")[0]))
        dllname = "{0}-{1}.so".format(runtime, network)
        if dllname not in _libraries:
            path = os.path.join(BNN_LIB_DIR, dllname)
            if not os.path.exists(path):
                raise RuntimeError("%s not built: run `make -C bnn-pynq_b200` (no CPU fallback exists)" % path)
            _libraries[dllname] = _ffi.dlopen(path)
        self.interface = _libraries[dllname]
        self.network = network
        self.is_cnv = network.startswith("cnv")  # decides the row size every in-memory entry point accepts
        self.num_classes = 0
        self.classes = []
        self.usecPerImage = 0.0

    def __del__(self):
        try:
            self.interface.deinit()
        except Exception:
            pass

    def load_parameters(self, params):
        if not os.path.isabs(params):
            params = os.path.join(BNN_PARAM_DIR, params)
        if os.path.isdir(params):
            self.interface.load_parameters(params.encode())
            with open(os.path.join(params, "classes.txt")) as f:
                self.classes = [c.strip() for c in f.readlines()]
        else:
            print("\nERROR: No such parameter directory \"" + params + "\"")

    def _rate(self, usec, count=None):
        if count is None:
            _say("Inference took %.2f microseconds" % usec)
        else:
            _say("Inference took %.2f microseconds, %.2f usec per image" % (usec * count, usec))
        _say("Classification rate: %.2f images per second" % (1000000.0 / usec if usec > 0 else 0.0))
        self.usecPerImage = usec

    def inference(self, path):
        usecperimage = _ffi.new("float *")
        result = self.interface.inference(path.encode(), _ffi.NULL, len(self.classes), usecperimage)
        self._rate(usecperimage[0])
        return result

    def detailed_inference(self, path):
        # the reference allocates len(classes) ints although the LFC library writes 64: allocate 64, return len(classes)
        details_ptr = _ffi.new("int[]", max(64, len(self.classes)))
        usecperimage = _ffi.new("float *")
        self.interface.inference(path.encode(), details_ptr, len(self.classes), usecperimage)
        self._rate(usecperimage[0])
        return np.array([details_ptr[i] for i in range(len(self.classes))], dtype=np.int32)

    def _take(self, result_ptr, count):
        if result_ptr == _ffi.NULL:
            raise RuntimeError("bnn_b200: inference failed (see stderr)")
        out = np.copy(np.frombuffer(_ffi.buffer(result_ptr, count * 4), dtype=np.int32))
        self.interface.free_results(result_ptr)
        return out

    def inference_multiple(self, path):
        size_ptr = _ffi.new("int *")
        usecperimage = _ffi.new("float *")
        result_ptr = self.interface.inference_multiple(path.encode(), len(self.classes), size_ptr, usecperimage, 0)
        out = self._take(result_ptr, size_ptr[0])
        self._rate(usecperimage[0], size_ptr[0])
        return out

    def inference_multiple_with_faults(self, path, num_faults, word_size, target_type, target_layers=[]):
        # the reference passes np.int (int64) through an int*, reading [0,1,2] as [0,0,1]: int32 here
        targets = self._targets_ptr(target_layers)
        size_ptr = _ffi.new("int *")
        usecperimage = _ffi.new("float *")
        result_ptr = self.interface.inference_multiple_with_faults(
            path.encode(), len(self.classes), size_ptr, usecperimage, num_faults, word_size, target_type, targets, len(target_layers))
        out = self._take(result_ptr, size_ptr[0])
        self._rate(usecperimage[0], size_ptr[0])
        return out

    def _targets_ptr(self, target_layers):
        if len(target_layers) == 0:
            return _ffi.cast("int *", 0)
        self._targets = np.ascontiguousarray(target_layers, dtype=np.int32)
        return _ffi.cast("int *", _ffi.from_buffer(self._targets))

    def _as_images(self, images):
        """uint8 array -> (contiguous [n, row] array, n) for THIS library's network kind: CNV rows are 3073-byte CIFAR records or
        3072 planar RGB bytes (a zero label byte is prepended), LFC rows 784 pixels.  A flat array is split by that row size.
        Anything else is a ValueError here; the library checks the byte count once more before it reads."""
        a = np.ascontiguousarray(images, dtype=np.uint8)
        rows = (3073, 3072) if self.is_cnv else (784,)
        if a.ndim == 1:
            fit = [r for r in rows if a.size % r == 0]
            if a.size and not fit:
                raise ValueError("%s: a flat input must hold whole rows of %s bytes, got %d bytes" % (self.network, " / ".join(map(str, rows)), a.size))
            a = a.reshape(-1, fit[0] if a.size else rows[0])
        if a.shape[0] == 0:
            return a.reshape(0, rows[0]), 0
        a = a.reshape(a.shape[0], -1)
        if a.shape[1] not in rows:
            raise ValueError("%s expects rows of %s bytes, got %d" % (self.network, " / ".join(map(str, rows)), a.shape[1]))
        if a.shape[1] == 3072:
            a = np.concatenate([np.zeros((a.shape[0], 1), np.uint8), a], axis=1)
        return np.ascontiguousarray(a), a.shape[0]

    def inference_buffer(self, images, detail=False):
        """inference_multiple / inference_multiple_detail on images already in memory (no file)."""
        a, n = self._as_images(images)
        usecperimage = _ffi.new("float *")
        ptr = _ffi.cast("const unsigned char *", _ffi.from_buffer(a)) if n else _ffi.NULL
        result_ptr = self.interface.inference_buffer(ptr, n, a.size, len(self.classes), usecperimage, 1 if detail else 0)
        out = self._take(result_ptr, n * len(self.classes) if detail else n)
        self._rate(usecperimage[0], n)
        return out

    def inference_buffer_with_faults(self, images, num_faults, word_size, target_type, target_layers=[]):
        a, n = self._as_images(images)
        usecperimage = _ffi.new("float *")
        ptr = _ffi.cast("const unsigned char *", _ffi.from_buffer(a)) if n else _ffi.NULL
        result_ptr = self.interface.inference_buffer_with_faults(
            ptr, n, a.size, len(self.classes), usecperimage, num_faults, word_size, target_type, self._targets_ptr(target_layers), len(target_layers))
        out = self._take(result_ptr, n)
        self._rate(usecperimage[0], n)
        return out

    def inference_multiple_detail(self, path):
        size_ptr = _ffi.new("int *")
        usecperimage = _ffi.new("float *")
        result_ptr = self.interface.inference_multiple(path.encode(), len(self.classes), size_ptr, usecperimage, 1)
        out = self._take(result_ptr, len(self.classes) * size_ptr[0])
        self._rate(usecperimage[0], size_ptr[0])
        return out

    def class_name(self, index):
        return self.classes[index]


class _Classifier:
    def __init__(self, network, params, runtime=RUNTIME_CUDA):
        if params in available_params(network):
            self.net = network
            self.params = params
            self.runtime = runtime
            self.usecPerImage = 0.0
            self.bnn = PynqBNN(runtime, network)
            self.bnn.load_parameters(os.path.join(params, network))
            self.classes = self.bnn.classes
        else:
            print("ERROR: parameters are not availlable for {0}".format(network))

    def _done(self, result):
        self.usecPerImage = self.bnn.usecPerImage
        return result

    def class_name(self, index):
        return self.bnn.classes[index]


class CnvClassifier(_Classifier):
    """bnn/bnn.py:210-345."""

    def image_to_cifar(self, img, fp):
        """PIL image -> one CIFAR-10 binary record (label byte + planar RGB), bnn/bnn.py:226-242."""
        from PIL import Image
        img.thumbnail((32, 32), getattr(Image, "LANCZOS", getattr(Image, "ANTIALIAS", 1)))
        background = Image.new('RGBA', (32, 32), (255, 255, 255, 0))
        background.paste(img, (int((32 - img.size[0]) / 2), int((32 - img.size[1]) / 2)))
        img = np.array(background)
        fp.write(np.identity(1, dtype=np.uint8).tobytes())
        for c in range(3):
            fp.write(img[:, :, c].flatten().tobytes())

    def _with_tmp(self, imgs, fn, *args):
        with tempfile.NamedTemporaryFile() as tmp:
            for img in imgs:
                self.image_to_cifar(img, tmp)
            tmp.flush()
            return self._done(fn(tmp.name, *args))

    def images_to_records(self, imgs):
        """PIL images -> uint8 [n][3073] CIFAR records through image_to_cifar, in memory (the reference writes a temp file,
        bnn/bnn.py:280-287)."""
        fp = io.BytesIO()
        for img in imgs:
            self.image_to_cifar(img, fp)
        return np.frombuffer(fp.getvalue(), np.uint8).reshape(-1, 3073)

    def classify_records(self, records, details=False):
        """uint8 [n][3073] records (or [n][3072] planar RGB) already in memory -> class ids, or [n*classes] scores."""
        return self._done(self.bnn.inference_buffer(records, details))

    def classify_records_with_faults(self, records, num_faults, word_size, target_type, target_layers=[]):
        return self._done(self.bnn.inference_buffer_with_faults(records, num_faults, word_size, target_type, target_layers))

    def classify_image(self, img):
        return self._with_tmp([img], self.bnn.inference)

    def classify_cifar(self, path):
        return self._done(self.bnn.inference(path))

    def classify_image_details(self, img):
        return self._with_tmp([img], self.bnn.detailed_inference)

    def classify_cifar_details(self, path):
        return self._done(self.bnn.detailed_inference(path))

    def classify_path(self, path):
        from PIL import Image
        return self.classify_image(Image.open(path))

    def classify_images(self, imgs):
        return self.classify_records(self.images_to_records(imgs))

    def classify_images_with_faults(self, imgs, num_faults, word_size, target_type, target_layers=[]):
        return self.classify_records_with_faults(self.images_to_records(imgs), num_faults, word_size, target_type, target_layers)

    def classify_cifars(self, path):
        return self._done(self.bnn.inference_multiple(path))

    def classify_cifars_with_faults(self, path, num_faults, word_size, target_type, target_layers=[]):
        return self._done(self.bnn.inference_multiple_with_faults(path, num_faults, word_size, target_type, target_layers))

    def classify_images_details(self, imgs):
        return self.classify_records(self.images_to_records(imgs), details=True)

    def classify_cifars_details(self, path):
        return self._done(self.bnn.inference_multiple_detail(path))

    def classify_paths(self, paths):
        from PIL import Image
        return self.classify_images([Image.open(p) for p in paths])


class LfcClassifier(_Classifier):
    """bnn/bnn.py:348-388."""

    def classify_mnist(self, mnist_format_file):
        return self._done(self.bnn.inference(mnist_format_file))

    def classify_mnists(self, mnist_format_file):
        return self._done(self.bnn.inference_multiple(mnist_format_file))

    def classify_mnists_with_faults(self, mnist_format_file, num_faults, flip_word, target_type, target_layers=[]):
        return self._done(self.bnn.inference_multiple_with_faults(mnist_format_file, num_faults, flip_word, target_type, target_layers))

    def classify_pixels(self, pixels):
        """uint8 [n][784] (or [n][28][28]) images already in memory -> class ids (no idx3 file)."""
        return self._done(self.bnn.inference_buffer(pixels))

    def classify_pixels_with_faults(self, pixels, num_faults, flip_word, target_type, target_layers=[]):
        return self._done(self.bnn.inference_buffer_with_faults(pixels, num_faults, flip_word, target_type, target_layers))
