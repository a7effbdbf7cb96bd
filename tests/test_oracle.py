"""CPU: pin the oracle on everything that can be pinned without TensorFlow."""
import json
import os

import numpy as np
import pytest

from _util import KO, oracle_for

GOLD = os.path.join(os.path.dirname(__file__), "golden")


def test_topology_constants_match_keras_resnet50():
    L = KO.build_resnet50_layers()
    assert len(L) == 177                                   # Keras model.layers
    kinds = {}
    for l in L:
        kinds[l.kind] = kinds.get(l.kind, 0) + 1
    assert kinds == {"InputLayer": 1, "ZeroPadding2D": 2, "Conv2D": 53, "BatchNormalization": 53, "Activation": 49,
                     "MaxPooling2D": 1, "Add": 16, "GlobalAveragePooling2D": 1, "Dense": 1}
    assert KO.total_params(L) == 25_636_712                # Keras ResNet50(include_top=True).count_params()
    assert len(KO.weight_shapes(L)) == 320                 # len(model.get_weights())
    assert KO.total_macs(L) == 3_857_973_248
    by = {l.name: l for l in L}
    assert by["conv1_conv"].out_shape == (112, 112, 64)
    assert by["pool1_pool"].out_shape == (56, 56, 64)
    assert by["conv3_block1_1_conv"].stride == 2 and by["conv3_block1_2_conv"].stride == 1   # v1: stride on the 1x1
    assert by["conv3_block4_out"].out_shape == (28, 28, 512)
    assert by["conv4_block6_out"].out_shape == (14, 14, 1024)
    assert by["conv5_block3_out"].out_shape == (7, 7, 2048)


def test_product_layer_table_agrees_with_oracle(model50):
    L = KO.build_resnet50_layers()
    assert [l.name for l in model50.layers] == [l.name for l in L]
    for a, b in zip(model50.layers, L):
        assert a.class_name == b.kind
        assert tuple(a.output_shape[1:]) == tuple(b.out_shape)
        assert a.inbound_names == b.inbound
    ws = KO.weight_shapes(L)
    assert [tuple(w.shape) for w in model50.get_weights()] == [s for _, _, s in ws]


@pytest.mark.parametrize("name,depths,n_layers,n_params", [("ResNet101", (3, 4, 23, 3), 347, 44_707_176),
                                                           ("ResNet152", (3, 8, 36, 3), 517, 60_419_944)])
def test_deeper_variants_agree_with_oracle_topology(pkg, name, depths, n_layers, n_params):
    """Keras' ResNet101 / ResNet152 (v1) are the same builder with other stack depths: published layer and
    parameter counts, and the product's stand-in lists the same layers, shapes and edges as the oracle."""
    import adapt_b200

    L = KO.build_resnet50_layers(depths)
    assert len(L) == n_layers and KO.total_params(L) == n_params
    m = getattr(adapt_b200, name)(weights=None)
    assert [l.name for l in m.layers] == [l.name for l in L]
    for a, b in zip(m.layers, L):
        assert a.class_name == b.kind and tuple(a.output_shape[1:]) == tuple(b.out_shape) and a.inbound_names == b.inbound
    assert len(oracle_for(m).layers) == n_layers       # depth inferred from the weight names


def test_conv_against_independent_formulation():
    torch = pytest.importorskip("torch")
    F = torch.nn.functional
    rng = np.random.default_rng(3)
    for (h, cin, cout, k, s, padding) in [(15, 8, 16, 3, 1, "same"), (16, 4, 8, 1, 2, "valid"), (20, 3, 8, 7, 2, "valid"),
                                          (9, 16, 8, 3, 1, "same")]:
        x = rng.standard_normal((2, h, h, cin)).astype(np.float32)
        w = rng.standard_normal((k, k, cin, cout)).astype(np.float32)
        b = rng.standard_normal(cout).astype(np.float32)
        got = KO.conv2d(x, w, b, s, padding)
        ref = F.conv2d(torch.from_numpy(x).permute(0, 3, 1, 2), torch.from_numpy(w).permute(3, 2, 0, 1),
                       torch.from_numpy(b), stride=s, padding=(k // 2 if padding == "same" else 0))
        np.testing.assert_allclose(got, ref.permute(0, 2, 3, 1).numpy(), rtol=1e-4, atol=1e-4)


def test_layer_math_known_answers():
    x = np.arange(2 * 4 * 4 * 1, dtype=np.float32).reshape(2, 4, 4, 1) - 10
    assert KO.zero_pad(x, 1).shape == (2, 6, 6, 1) and KO.zero_pad(x, 1)[0, 0, 0, 0] == 0
    mp = KO.max_pool(KO.zero_pad(x, 1), 3, 2)
    assert mp.shape == (2, 2, 2, 1)
    assert mp[0, 0, 0, 0] == 0.0          # window covers padding zeros and negatives only -> 0
    assert mp[1, 1, 1, 0] == x[1].max()
    g = KO.global_avg_pool(x)
    np.testing.assert_allclose(g[:, 0], [x[0].mean(), x[1].mean()])
    z = np.array([[1.0, 2.0, 3.0]], np.float32)
    np.testing.assert_allclose(KO.softmax(z).sum(), 1.0, rtol=1e-6)
    bn = KO.batch_norm(np.ones((1, 1, 1, 2), np.float32), np.array([2, 1], np.float32), np.array([0.5, 0], np.float32),
                       np.array([1, 0], np.float32), np.array([1, 4], np.float32))
    np.testing.assert_allclose(bn.ravel(), [0.5, 1 / np.sqrt(4 + 1.001e-5)], rtol=1e-6)
    r = KO.bf16_round(np.array([1.0, 1.00390625, 1.01171875, -3.1415927], np.float32))
    np.testing.assert_array_equal(r, np.array([1.0, 1.0, 1.015625, -3.140625], np.float32))


def test_pipeline_equals_unpartitioned(model50):
    o = oracle_for(model50)
    x = KO.synthetic_images(1, seed=5, kind="normal")
    full, _ = o.predict(x)
    for part_at in (["conv3_block4_out"], ["conv2_block3_out", "conv3_block4_out", "conv4_block6_out"],
                    ["pool1_pad", "conv2_block1_add", "avg_pool"]):
        np.testing.assert_array_equal(o.pipeline(x, part_at), full)
    with pytest.raises(RuntimeError):
        o.layers_between("conv2_block1_1_relu", "predictions")


def test_golden_vectors(model50):
    """Fixtures generated by tests/golden/make_golden.py from this oracle + the product's weight
    generator: guards both against silent drift (they are NOT reference outputs: parity unpinned)."""
    path = os.path.join(GOLD, "resnet50_seed1234_golden.json")
    gold = json.load(open(path))
    o = oracle_for(model50)
    x = KO.synthetic_images(gold["batch"], seed=gold["input_seed"], kind=gold["input_kind"])
    assert abs(float(np.float64(x).sum()) - gold["input_sum"]) < 1e-3 * abs(gold["input_sum"]) + 1e-3
    probs, taps = o.predict(x, taps=list(gold["tap_mean_abs"].keys()))
    assert probs.argmax(-1).tolist() == gold["argmax"]
    np.testing.assert_allclose(np.sort(probs, -1)[:, -5:], np.array(gold["top5_probs"]), rtol=2e-4)
    for name, v in gold["tap_mean_abs"].items():
        assert abs(float(np.abs(taps[name]).mean()) - v) <= 2e-4 * abs(v) + 1e-7, name
    np.testing.assert_allclose(o.logits[:, :8], np.array(gold["logits_head"]), rtol=2e-4, atol=2e-4)


def test_timed_cpu_stage_equals_the_oracle(model50):
    """`oracle/cpu_pipeline.py::TorchCpuStage` is what bench.py's cpu_baseline and `--impl reference` TIME (torch CPU
    kernels); it must compute what the numpy oracle computes, on every partition BASELINE.json names."""
    from oracle.cpu_pipeline import TorchCpuStage

    weights = {l.name: l.get_weights() for l in model50.layers if l.get_weights()}
    o = oracle_for(model50)
    x = KO.synthetic_images(1, seed=9, kind="caffe")
    full = o.predict(x)[0]
    for part_at in ([], ["conv3_block4_out"], ["conv2_block3_out", "conv3_block4_out", "conv4_block6_out"]):
        bounds = ["input_1"] + part_at + ["predictions"]
        cur, cur_ref = x, x
        for a, b in zip(bounds[:-1], bounds[1:]):
            cur = TorchCpuStage(weights, a, b, threads=4)(cur)
            cur_ref = o.run_range(cur_ref, a, b)[0]
            assert cur.shape == cur_ref.shape
            assert np.abs(cur - cur_ref).max() <= 1e-5 * max(1.0, float(np.abs(cur_ref).max())), (a, b)
        assert np.abs(cur - full).max() <= 1e-5 and cur.argmax(-1).tolist() == full.argmax(-1).tolist()


def test_tf_crosscheck_fixture_is_current(model50):
    """tests/golden/resnet50_seed1234_tf_crosscheck.npz (tools/export_for_tf.py) is the hand-over artefact a maintainer
    with TensorFlow runs tools/tf_check.py against.  Here: it still describes today's weight generator and oracle
    (they are NOT reference outputs -- parity stays unpinned until tf_check.py has been run somewhere)."""
    import sys

    sys.path.insert(0, os.path.join(os.path.dirname(GOLD), "..", "tools"))
    from export_for_tf import make_inputs, weights_digest

    g = np.load(os.path.join(GOLD, "resnet50_seed1234_tf_crosscheck.npz"))
    assert int(g["seed"]) == 1234 and int(g["n_weight_arrays"]) == 320
    assert weights_digest(model50.get_weights()) == str(g["weights_sha256"])
    x = make_inputs()
    assert abs(float(np.float64(x).sum()) - float(g["inputs_sum"])) < 1e-3
    o = oracle_for(model50)
    names = [str(n) for n in g["tap_names"]]
    probs, taps = o.predict(x[1:2], taps=names)          # (one of the two images keeps the CPU suite short)
    np.testing.assert_allclose(probs, g["probs"][1:2], rtol=1e-4, atol=1e-7)
    np.testing.assert_allclose(o.logits, g["logits"][1:2], rtol=1e-4, atol=1e-5)
    for n in names:
        s = taps[n].reshape(1, -1)[:, ::int(g["tap_stride"])]
        np.testing.assert_allclose(s, g["tap_" + n][1:2], rtol=1e-4, atol=1e-5, err_msg=n)
