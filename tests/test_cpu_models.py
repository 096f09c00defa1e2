"""CPU unit tests: model topology against the Keras numbers recorded in SURVEY.md section 4 / Appendix A."""
import pytest
import torch

from idc_models_b200.models import spec as S
from idc_models_b200.models.applications import build_transfer_model


@pytest.mark.parametrize("arch,layers,params,out_shape", [
    ("densenet121", 427, 7037504, (1, 1, 1024)),
    ("densenet169", 595, 12642880, (1, 1, 1664)),
    ("densenet201", 707, 18321984, (1, 1, 1920)),
    ("mobilenet_v2", 155, 2257984, (2, 2, 1280)),
    ("vgg16", 19, 14714688, (1, 1, 512)),
])
def test_layer_and_param_counts(arch, layers, params, out_shape):
    sp = S.build_spec(arch, (50, 50, 3))
    assert len(sp.layers) == layers
    assert sp.param_count() == params
    assert sp.output.out_shape == out_shape


def test_small_cnn_matches_reference_shapes():
    sp = S.build_spec("small_cnn")
    shapes = [tuple(p.shape) for l in sp.layers for p in l.params]
    # reference (Keras layouts): (3,3,3,32) (32,) (128,8) (8,) (8,1) (1,)  -- secure_fed_model.py:73-78
    assert shapes == [(32, 3, 3, 3), (32,), (8, 128), (8,), (1, 8), (1,)]
    assert sp.param_count() == 1937


@pytest.mark.parametrize("arch,ft,n_tensors,n_floats", [
    ("vgg16", 15, 8, 7079937),
    ("mobilenet_v2", 100, 58, 1863873),
    ("densenet121", 150, 239, 5453248 + 1025),
])
def test_fine_tune_at_trainable_sets(arch, ft, n_tensors, n_floats):
    m = build_transfer_model(arch, device="cpu")
    m.base.trainable = True
    for l in m.base.layers[:ft]:
        l.trainable = False
    tw = m.trainable_weights
    assert len(tw) == n_tensors
    assert sum(t.numel() for t in tw) == n_floats
    # the trainable set is ONE contiguous range of the flat master buffer (packed all-reduce bucket)
    assert len(m.store.ranges(m.trainable_layer_indices())) == 1


def test_keras_layer_indices():
    m = S.build_spec("mobilenet_v2")
    assert m.layers[99].name == "block_11_expand" and m.layers[100].name == "block_11_expand_BN"
    assert m.layers[152].name == "Conv_1" and m.layers[154].name == "out_relu"
    d = S.build_spec("densenet121")
    assert d.layers[150].name == "conv4_block2_1_conv" and d.layers[149].name == "conv4_block2_0_relu"
    v = S.build_spec("vgg16")
    assert [v.layers[i].name for i in (15, 16, 17, 18)] == ["block5_conv1", "block5_conv2", "block5_conv3", "block5_pool"]


def test_correct_pad_asymmetry():
    assert S.correct_pad(50) == (0, 1) and S.correct_pad(25) == (1, 1)


def test_densenet201_on_cifar_shape():
    sp = S.build_spec("densenet201", (32, 32, 3))
    assert sp.output.out_shape == (1, 1, 1920)


def test_frozen_bn_runs_in_inference_mode_and_trainable_bn_updates_moving_stats():
    m = build_transfer_model("mobilenet_v2", device="cpu")
    m.compile(optimizer="rmsprop", loss="bce", metrics=["accuracy"])
    x = torch.rand(4, 50, 50, 3)
    y = torch.tensor([0., 1., 0., 1.])
    before = m.store.stats.clone()
    m.fit([(x, y)], epochs=1, verbose=0)
    assert torch.equal(before, m.store.stats), "a frozen base must not touch BN moving statistics"
    m.base.trainable = True
    m.compile(optimizer="rmsprop", loss="bce", metrics=["accuracy"])
    m.fit([(x, y)], epochs=1, verbose=0)
    assert not torch.equal(before, m.store.stats)


def test_get_set_weights_keras_layouts_roundtrip():
    m = build_transfer_model("vgg16", device="cpu")
    w = m.get_weights()
    assert w[0].shape == (3, 3, 3, 64) and w[1].shape == (64,)        # HWIO kernel, bias
    assert w[-2].shape == (512, 1) and w[-1].shape == (1,)            # Dense (in, out)
    m2 = build_transfer_model("vgg16", device="cpu", seed=3)
    m2.set_weights(w)
    assert all((a == b).all() for a, b in zip(w, m2.get_weights()))
    # channel-padded stem storage keeps its pad lanes at zero
    assert m2.store.storage_view("block1_conv1/kernel", "master")[..., 3:].abs().max() == 0


def test_densenet201_cifar_fine_tune_set_matches_survey():
    """dist_model_tf_dense.py: DenseNet201, 32x32x3, 10 classes, fine_tune_at = 150 -> 479 trainable tensors,
    16,611,530 floats in the packed gradient all-reduce."""
    m = build_transfer_model("densenet201", (32, 32, 3), 10, device="cpu")
    m.base.trainable = True
    for l in m.base.layers[:150]:
        l.trainable = False
    tw = m.trainable_weights
    assert len(tw) == 479
    assert sum(t.numel() for t in tw) == 16611530


def test_incremental_sequential_add_builds_the_same_model():
    """``model = Sequential(); model.add(...)`` (the form of /root/reference/secure_fed_model.py:85-94) builds the
    network of the list form, lazily at first use; adding afterwards is refused"""
    import numpy as np
    import pytest
    import torch
    from idc_models_b200.models import applications as apps
    from idc_models_b200.models.network import clone_model

    def layers():
        return [apps.Conv2D(32, kernel_size=(3, 3), activation="relu", input_shape=(10, 10, 3), strides=2),
                apps.MaxPooling2D(pool_size=(2, 2)), apps.Dropout(0.25), apps.Flatten(),
                apps.Dense(8, activation="relu"), apps.Dropout(0.5), apps.Dense(1)]
    m = apps.Sequential(device="cpu", seed=3)
    for l in layers():
        m.add(l)
    ref = apps.Sequential(layers(), device="cpu", seed=3)
    m.compile(optimizer="rmsprop", loss="bce", metrics=["accuracy"])
    assert m.count_params() == ref.count_params() == 1937
    assert all(np.array_equal(a, b) for a, b in zip(m.get_weights(), ref.get_weights()))
    x, y = torch.rand(8, 10, 10, 3), (torch.rand(8) > 0.5).float()
    h = m.fit([(x, y)] * 2, epochs=1, verbose=0)
    assert len(h.history["loss"]) == 1
    with pytest.raises(RuntimeError, match="after the model has been built"):
        m.add(apps.Dense(1))
    c = clone_model(m.build())
    assert c.count_params() == 1937
    with pytest.raises(ValueError, match="no layers"):
        apps.Sequential(device="cpu").compile(optimizer="sgd", loss="bce")
