"""Shared helpers for the tests: the oracle is only ever used here as the checker."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from oracle import keras_resnet50_oracle as KO  # noqa: E402

# bf16 mode against the fp32 oracle: north_star bounds |dlogit| by 1e-2; d(prob) / prob ~ d(logit), so the bound on
# probabilities is RELATIVE to the row's largest probability (an absolute 1e-2 on numbers of size 1e-3 binds nothing)
PROB_RTOL_BF16 = 1e-2
PROB_RTOL_FP32 = 1e-4


def oracle_for(model):
    """Oracle evaluating the SAME weight tensors as `model` (the product's Keras stand-in)."""
    weights = {l.name: l.get_weights() for l in model.layers if l.get_weights()}
    return KO.ResNet50Oracle(weights=weights)


def rel_err(got, ref):
    got = np.asarray(got, np.float64)
    ref = np.asarray(ref, np.float64)
    return float(np.abs(got - ref).max() / max(np.abs(ref).max(), 1e-30))


def distinct_images(batch, seed, index, kind="normal"):
    """Synthetic micro-batch number `index` of a stream.  Random-init ResNet-50 maps N(0,1) images to nearly the same
    probabilities (they differ by ~1e-5), so every image is also scaled by a factor in [1, 3) that changes from image
    to image and from micro-batch to micro-batch: the network is positively homogeneous (zero biases), the logits
    scale with the image and the probabilities of neighbouring micro-batches differ by percents.  A result
    delivered for the wrong input is then a value error even at the oracle's tolerance."""
    x = KO.synthetic_images(batch, seed=seed + index, kind=kind)
    scale = 1.0 + ((index * 5 + np.arange(batch) * 3) % 16).astype(np.float32) / 8.0
    return np.ascontiguousarray(x * scale[:, None, None, None], dtype=np.float32)


def fused_boundary_cuts(part_at):
    """True when every cut is at the end of a fused kernel group (a residual block's `_out`, pool1_pool, avg_pool):
    the partitioned pipeline then runs the same kernels on the same inputs as the un-partitioned stage, bit for bit
    (tests/test_stage_gpu.py::test_partition_chain_matches_oracle_and_unpartitioned proves that identity)."""
    return all(c.endswith("_out") or c in ("pool1_pool", "avg_pool") for c in part_at)


class StreamChecker:
    """Value check of what a pipeline delivered, micro-batch by micro-batch, against
      (1) the un-partitioned model on ONE GPU stage (`LocalPredictor(model).run(x)` of THAT input): bit for bit when
          every cut is a fused-group boundary, else within bf16 re-rounding (relative 3e-2 of the row maximum);
      (2) the oracle: same arg-max, max |dprob| <= rtol * max prob.
    Also requires the expected outputs of different micro-batches to differ, so that a swapped, stale or
    duplicated micro-batch cannot pass (1)."""

    def __init__(self, model, part_at, mode="bf16", max_batch=4, input_format="float32", preprocess=None, device=0,
                 oracle_tol="relative"):
        from adapt_b200.executor import LocalPredictor

        self.model = model
        self.exact = fused_boundary_cuts(part_at)
        self.mode = mode
        self.preprocess = preprocess
        self.local = LocalPredictor(model, mode=mode, max_batch=max_batch, input_format=input_format, device=device)
        self.oracle = oracle_for(model)
        # "relative": N(0,1)-type inputs (logits O(0.1..1)): max |dprob| <= 1e-2 (bf16) / 1e-4 (fp32) of the row's
        # largest probability, the image of north_star's bound on logits.  "absolute": 0..255 caffe-type inputs, whose
        # logits are O(10) and whose bf16 logit error (~0.09) no longer fits 1e-2: the bound is then stated on the
        # probabilities themselves, max |dprob| <= 1e-2 / 1e-4.
        assert oracle_tol in ("relative", "absolute")
        self.oracle_tol = oracle_tol
        self.rtol = PROB_RTOL_BF16 if mode == "bf16" else PROB_RTOL_FP32

    def expected(self, inputs):
        return [self.local.run(np.asarray(x)) for x in inputs]

    def check(self, inputs, outs, oracle_every=1):
        assert len(outs) == len(inputs)
        want = self.expected(inputs)
        for i in range(len(want)):
            for j in range(i + 1, len(want)):
                if want[i].shape == want[j].shape and not np.array_equal(np.asarray(inputs[i]), np.asarray(inputs[j])):
                    assert not np.array_equal(want[i], want[j]), f"micro-batches {i} and {j} are indistinguishable"
        for i, (x, y, w) in enumerate(zip(inputs, outs, want)):
            assert y.shape == w.shape and y.dtype == np.float32, (i, y.shape, w.shape)
            if self.exact:
                np.testing.assert_array_equal(y, w, err_msg=f"micro-batch {i} differs from the un-partitioned model")
            else:
                assert rel_err(y, w) <= 3e-2, f"micro-batch {i}: {rel_err(y, w)} vs the un-partitioned model"
            if oracle_every and i % oracle_every == 0:
                xf = np.asarray(x)
                if self.preprocess is not None:
                    xf = self.preprocess(xf)
                ref = self.oracle.predict(xf)[0]
                assert y.argmax(-1).tolist() == ref.argmax(-1).tolist(), f"micro-batch {i}: arg-max differs from the oracle"
                err = np.abs(y - ref).max(-1) / (ref.max(-1) if self.oracle_tol == "relative" else 1.0)
                assert err.max() <= self.rtol, f"micro-batch {i}: max |dprob| ({self.oracle_tol}) = {err.max():.3e} > {self.rtol}"

    def close(self):
        self.local.close()
