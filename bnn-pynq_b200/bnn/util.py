"""Helpers of the fault campaigns on the CUDA runtime: the counterpart of BNN-PYNQ's bnn/util/util.py.

Same function names and return values (calculate_accuracy :11-16, dict_of_dicts_merge :19-28, dict_to_str / write_dict_to_file
:30-39, load_cifar10_testset :43-68, load_gtsrb_testset :72-92, load_svhn_testset :96-117, load_mnist_testset :121-134).  The
reference formats its JSON through yapf (bnn/util/util.py:7,30-33); here it is plain json with an indent, so the module has no
dependency beyond numpy (PIL / scipy are imported only by the two loaders that need them).
"""
import copy
import csv
import json
import os

import numpy as np


def calculate_accuracy(results, labels):
    """Percentage of positions where results[i] == labels[i], over len(labels)."""
    n = len(labels)
    if n == 0:
        return 0.0
    r = np.asarray(results[:n])
    return float(np.count_nonzero(r == np.asarray(labels))) * 100 / n


def dict_of_dicts_merge(*dicts):
    """Recursive union: sub-dictionaries under the same key are merged, any other clash keeps the first value."""
    out = {}
    for d in dicts:
        for key, value in d.items():
            if key not in out:
                out[key] = copy.deepcopy(value)
            elif isinstance(out[key], dict) and isinstance(value, dict):
                out[key] = dict_of_dicts_merge(out[key], value)
    return out


def dict_to_str(d):
    return json.dumps(d, indent=2, sort_keys=False) + "\n"


def write_dict_to_file(file_name, output_dict):
    folder = os.path.dirname(file_name)
    if folder:
        os.makedirs(folder, exist_ok=True)
    with open(file_name, "w") as f:
        f.write(dict_to_str(output_dict))


def load_cifar10_testset(folder, num_images=10000):
    """(path of a CIFAR-10 binary batch holding the first num_images test records, their labels).
    A truncated copy test_batch_<n>.bin is written next to test_batch.bin the first time it is asked for."""
    num_images = min(num_images, 10000)
    full = os.path.join(folder, "test_batch.bin")
    path = full if num_images == 10000 else os.path.join(folder, "test_batch_%d.bin" % num_images)
    if not os.path.exists(path):
        rec = np.fromfile(full, np.uint8, num_images * 3073)
        rec.tofile(path)
    rec = np.fromfile(path, np.uint8, num_images * 3073).reshape(-1, 3073)
    return path, [int(x) for x in rec[:, 0]]


def load_mnist_testset(folder, num_images=10000):
    """(path of the idx3 image file, the first num_images labels of the idx1 label file)."""
    num_images = min(num_images, 10000)
    lbl = np.fromfile(os.path.join(folder, "t10k-labels-idx1-ubyte"), np.uint8)
    return os.path.join(folder, "t10k-images-idx3-ubyte"), [int(x) for x in lbl[8:8 + num_images]]


def load_gtsrb_testset(gt_file, images_folder, num_images=12630):
    """(PIL images, labels) of the GTSRB test set: ';'-separated ground-truth file, class id in column 7."""
    from PIL import Image
    num_images = min(num_images, 12630)
    files, labels = [], []
    with open(gt_file) as f:
        rows = csv.reader(f, delimiter=";")
        next(rows)
        for row in rows:
            if len(files) == num_images:
                break
            files.append(os.path.join(images_folder, row[0]))
            labels.append(int(row[7]))
    images = []
    for p in files:
        im = Image.open(p)
        im.load()
        images.append(im)
    return images, labels


def load_svhn_testset(file, num_images=26032):
    """(PIL images, labels) of the SVHN test .mat file; labels are shifted by -1 to match the classifier's classes."""
    import scipy.io
    from PIL import Image
    num_images = min(num_images, 26032)
    data = scipy.io.loadmat(file)
    labels = [int(x) for x in (data["y"].reshape(-1) - 1)[:num_images]]
    images = [Image.fromarray(data["X"][:, :, :, i]) for i in range(num_images)]
    return images, labels
