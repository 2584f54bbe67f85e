"""Fault-injection campaigns on the CUDA runtime: the counterpart of BNN-PYNQ's bnn/faults/faults.py.

Same classes and call signatures — FaultTest (+ TargetType), CNVFaultTest, LFCFaultTest (bnn/faults/faults.py:26-108) and
NetworkTest (+ TestType, test_network, comprehensive_test; :126-261) — and the same result dictionaries / file names, so existing
campaign scripts only change their import.  Differences, all consequences of the runtime:
  * the reference hard-wires RUNTIME_HW (faults.py:53); here the runtime is a constructor argument defaulting to RUNTIME_CUDA;
  * the reference builds a new classifier (bitstream download + parameter load) for every run to reset the injected faults; here
    one classifier is kept per FaultTest and `load_parameters` is called between runs — in this runtime that is exactly what
    discards injected faults, and it is a device-side re-pack, not a reconfiguration;
  * a 100-run x 10 000-image campaign is seconds, so `quiet=True` suppresses the per-run lines if wanted.
"""
import os

from . import bnn, util


class FaultTest:
    class TargetType:
        @staticmethod
        def any():
            return -1

        @staticmethod
        def weights():
            return 0

        @staticmethod
        def thresholds():
            return 1

    def __init__(self, classifier_cls, network, dataset, input_file, labels, runtime=bnn.RUNTIME_CUDA, quiet=False):
        self.classifier_cls = classifier_cls
        self.network = network
        self.dataset = dataset
        self.input_file = input_file
        self.labels = labels
        self.runtime = runtime
        self.quiet = quiet
        self._classifier = None

    def _fresh_classifier(self):
        """A classifier whose parameters carry no injected faults."""
        if self._classifier is None:
            self._classifier = self.classifier_cls(self.network, self.dataset, self.runtime)
        else:
            self._classifier.bnn.load_parameters(os.path.join(self.dataset, self.network))   # reload = faults discarded
        return self._classifier

    def _describe(self, i, num_runs, num_flips, word_size, target_type, target_layers):
        what = " weight" if target_type == 0 else " threshold" if target_type == 1 else ""
        where = "any layer" if len(target_layers) == 0 else "layer(s) {}".format(target_layers)
        return "{}-{} run {} of {} (flipping {}{} {}(s) in {})".format(
            self.network, self.dataset, i + 1, num_runs, num_flips, what, "word" if word_size > 1 else "bit", where)

    def run_test(self, num_runs, num_flips, word_size=1, target_type=-1, target_layers=[]):
        """num_runs independent runs, each injecting num_flips faults of word_size adjacent bits at random times over the image
        set.  Returns (per-run class lists, per-run usec per image, per-run accuracies in percent)."""
        results, times, accuracies = [], [], []
        for i in range(num_runs):
            c = self._fresh_classifier()
            if not self.quiet:
                print(self._describe(i, num_runs, num_flips, word_size, target_type, target_layers))
            if self.dataset == "cifar10":
                r = c.classify_cifars_with_faults(self.input_file, num_flips, word_size, target_type, target_layers)
            elif self.dataset == "mnist":
                r = c.classify_mnists_with_faults(self.input_file, num_flips, word_size, target_type, target_layers)
            else:
                r = c.classify_images_with_faults(self.input_file, num_flips, word_size, target_type, target_layers)
            r = [int(x) for x in r]
            results.append(r)
            times.append(c.usecPerImage)
            accuracies.append(util.calculate_accuracy(r, self.labels))
            if not self.quiet:
                print("Accuracy:", accuracies[-1])
                print()
        return results, times, accuracies


class CNVFaultTest(FaultTest):
    def __init__(self, network, dataset, input_file, labels, runtime=bnn.RUNTIME_CUDA, quiet=False):
        super().__init__(bnn.CnvClassifier, network, dataset, input_file, labels, runtime, quiet)

    @classmethod
    def CIFARTest(cls, network, input_file, labels, **kw):
        return cls(network, "cifar10", input_file, labels, **kw)

    @classmethod
    def SVHNTest(cls, network, input_file, labels, **kw):
        return cls(network, "streetview", input_file, labels, **kw)

    @classmethod
    def GTSRBTest(cls, network, input_file, labels, **kw):
        return cls(network, "road-signs", input_file, labels, **kw)


class LFCFaultTest(FaultTest):
    def __init__(self, network, dataset, input_file, labels, runtime=bnn.RUNTIME_CUDA, quiet=False):
        super().__init__(bnn.LfcClassifier, network, dataset, input_file, labels, runtime, quiet)

    @classmethod
    def MNISTTest(cls, network, input_file, labels, **kw):
        return cls(network, "mnist", input_file, labels, **kw)


class NetworkTest:
    """Runs a FaultTest over combinations of target (any / weights / thresholds) and upset size (bit / word) and writes the raw
    results and their statistics as JSON under <output_folder>/<network>/<dataset>/<n>flips/."""

    class TestType:
        def __init__(self, target_type, word_size):
            self.target_type = target_type
            self.word_size = word_size
            target = {FaultTest.TargetType.any(): "any", FaultTest.TargetType.weights(): "weight"}.get(target_type, "threshold")
            self.name = target + " " + ("word" if word_size > 1 else "bit")

        @classmethod
        def any_bit(cls):
            return cls(FaultTest.TargetType.any(), 1)

        @classmethod
        def any_word(cls, word_size=8):
            return cls(FaultTest.TargetType.any(), word_size)

        @classmethod
        def weight_bit(cls):
            return cls(FaultTest.TargetType.weights(), 1)

        @classmethod
        def weight_word(cls, word_size=8):
            return cls(FaultTest.TargetType.weights(), word_size)

        @classmethod
        def threshold_bit(cls):
            return cls(FaultTest.TargetType.thresholds(), 1)

        @classmethod
        def threshold_word(cls, word_size=8):
            return cls(FaultTest.TargetType.thresholds(), word_size)

    def __init__(self, fault_test):
        self.fault_test = fault_test
        self.control = None

    def _run_control(self):
        if not self.fault_test.quiet:
            print("Running", self.fault_test.network + "-" + self.fault_test.dataset, "control test")
        _, _, accuracy = self.fault_test.run_test(num_runs=1, num_flips=0)
        self.control = accuracy[0]

    def _raw(self, name, num_runs, num_flips, layers, accuracies):
        return {"network": self.fault_test.network, "dataset": self.fault_test.dataset, "run count": num_runs, "flips": num_flips,
                "control": self.control, "layers": layers, "results": {name: accuracies}}

    def _stats(self, merged):
        """Per test type: all runs, the runs whose accuracy differs from the control ('effective'), min / max / averages and the
        deltas against the control (keys as in the reference's result files)."""
        out = dict(merged)
        out["results"] = {}
        control = merged["control"]
        for name, runs in merged["results"].items():
            effective = [a for a in runs if a != control]
            e = {"runs": {"all": runs, "effective": effective}, "effective count": len(effective), "min accuracy": min(runs),
                 "max accuracy": max(runs)}
            if effective:
                e["avg accuracy"] = sum(runs) / len(runs)
                e["avg effective accuracy"] = sum(effective) / len(effective)
                e["accuracy delta"] = control - e["avg accuracy"]
                e["effective accuracy delta"] = control - e["avg effective accuracy"]
            else:
                e["avg accuracy"] = control
            out["results"][name] = e
        return out

    def _run_tests(self, folder, num_runs, num_flips, test_types, target_layers):
        raws = []
        for test in test_types:
            _, _, accuracies = self.fault_test.run_test(num_runs, num_flips, test.word_size, test.target_type, target_layers)
            raw = self._raw(test.name, num_runs, num_flips, target_layers, accuracies)
            raws.append(raw)
            util.write_dict_to_file(os.path.join(folder, "temp", self.fault_test.network + "_results_" + test.name.replace(" ", "-") + ".json"), raw)
        return self._stats(util.dict_of_dicts_merge(*raws))

    def test_network(self, output_folder, num_runs, flip_counts, test_types, target_layers=[]):
        root = os.path.join(output_folder, self.fault_test.network, self.fault_test.dataset)
        if self.control is None:
            self._run_control()
        written = []
        for num_flips in flip_counts:
            folder = os.path.join(root, str(num_flips) + "flips")
            stats = self._run_tests(folder, num_runs, num_flips, test_types, target_layers)
            name = self.fault_test.network + "_" + self.fault_test.dataset
            name += "_stats_layer{}.json".format(target_layers) if len(target_layers) > 0 else "_stats.json"
            util.write_dict_to_file(os.path.join(folder, name), stats)
            written.append(os.path.join(folder, name))
        return written

    def comprehensive_test(self, output_folder, num_runs, flip_counts, target_layers=[]):
        """All six target x upset-size combinations (the reference's comprehensive_test builds this list and stops there,
        bnn/faults/faults.py:256-261; here it also runs it)."""
        T = NetworkTest.TestType
        types = [T.any_bit(), T.any_word(), T.weight_bit(), T.weight_word(), T.threshold_bit(), T.threshold_word()]
        return self.test_network(output_folder, num_runs, flip_counts, types, target_layers)
