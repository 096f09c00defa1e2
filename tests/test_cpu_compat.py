"""idc_models_b200.compat: a script written against the reference's imports (tf.keras / tf.distribute / tf.data /
tf.image, tff.learning / tff.simulation, phe.paillier, sklearn's roc_auc_score) runs after changing the import lines
only (/root/reference/dist_model_tf_mobile.py:105-170, fed_model.py:186-236, secure_fed_model.py:85-121)."""
import io

import numpy as np
import pytest
import torch

from idc_models_b200.compat import paillier, roc_auc_score, tf, tff


def test_transfer_script_shape_with_reference_names():
    strategy = tf.distribute.MirroredStrategy()
    assert strategy.num_replicas_in_sync == 1
    IMG_SHAPE = (50, 50, 3)
    with strategy.scope():
        base_model = tf.keras.applications.MobileNetV2(input_shape=IMG_SHAPE, include_top=False, weights=None)
        base_model.trainable = False
        model = tf.keras.Sequential([base_model, tf.keras.layers.GlobalAveragePooling2D(), tf.keras.layers.Dense(1)],
                                    device="cpu")
        model.compile(optimizer=tf.keras.optimizers.RMSprop(lr=1e-4),
                      loss=tf.keras.losses.BinaryCrossentropy(from_logits=True), metrics=["accuracy"])
    assert len(base_model.layers) == 155 and len(model.trainable_weights) == 2
    x, y = torch.rand(4, 50, 50, 3), (torch.rand(4) > 0.5).float()
    ds = tf.data.Dataset.from_tensor_slices((x.numpy(), y.numpy())).batch(2).prefetch(tf.data.experimental.AUTOTUNE)
    history = model.fit(ds, epochs=1, validation_data=ds, verbose=0)
    assert set(history.history) >= {"loss", "accuracy", "val_loss", "val_accuracy"}
    # phase 2 of the scripts: unfreeze, re-freeze below fine_tune_at, re-compile
    base_model.trainable = True
    for layer in base_model.layers[:100]:
        layer.trainable = False
    model.compile(optimizer=tf.keras.optimizers.RMSprop(lr=1e-5),
                  loss=tf.keras.losses.BinaryCrossentropy(from_logits=True), metrics=["accuracy"])
    assert len(model.trainable_weights) == 58
    loss0, acc0 = model.evaluate(ds, steps=2)
    assert np.isfinite(loss0) and 0.0 <= acc0 <= 1.0


def test_secure_script_model_and_keys_with_reference_names():
    model = tf.keras.models.Sequential(device="cpu")
    model.add(tf.keras.layers.Conv2D(32, kernel_size=(3, 3), activation="relu", input_shape=(10, 10, 3), strides=2))
    model.add(tf.keras.layers.MaxPooling2D(pool_size=(2, 2)))
    model.add(tf.keras.layers.Dropout(0.25))
    model.add(tf.keras.layers.Flatten())
    model.add(tf.keras.layers.Dense(8, activation="relu"))
    model.add(tf.keras.layers.Dropout(0.5))
    model.add(tf.keras.layers.Dense(1))
    model.compile(optimizer="rmsprop", loss=tf.keras.losses.BinaryCrossentropy(from_logits=True), metrics=["accuracy"])
    weights = model.get_weights()
    assert [w.shape for w in weights] == [(3, 3, 3, 32), (32,), (128, 8), (8,), (8, 1), (1,)]
    pubkey, privkey = paillier.generate_paillier_keypair(n_length=256)
    enc = [pubkey.encrypt(float(v)) for v in weights[-1].reshape(-1)]          # Client.encrypt_weights, one scalar
    s = enc[0] + enc[0]
    assert abs(privkey.decrypt(s) / 2 - float(weights[-1].reshape(-1)[0])) < 1e-6
    model.set_weights(weights)
    clone = tf.keras.models.clone_model(model)
    assert clone.count_params() == model.count_params() == 1937
    assert roc_auc_score(np.array([0, 0, 1, 1]), np.array([0.1, 0.4, 0.35, 0.8])) == pytest.approx(0.75)


def test_federated_script_shape_with_reference_names():
    def create_tf_dataset_for_client_fn(client_id):
        g = torch.Generator().manual_seed(int(client_id))
        x, y = torch.rand(8, 10, 10, 3, generator=g), (torch.rand(8, generator=g) > 0.5).float()
        return tf.data.Dataset.from_tensor_slices((x.numpy(), y.numpy())).batch(4)

    client_data = tff.simulation.ClientData.from_clients_and_fn(list(range(3)), create_tf_dataset_for_client_fn)
    federated_train_data = [client_data.create_tf_dataset_for_client(c) for c in client_data.client_ids]

    def model_fn():
        m = tf.keras.Sequential([tf.keras.layers.Conv2D(8, (3, 3), activation="relu", input_shape=(10, 10, 3), strides=2),
                                 tf.keras.layers.Flatten(), tf.keras.layers.Dense(1)], device="cpu")
        return tff.learning.from_keras_model(m, loss=tf.keras.losses.BinaryCrossentropy(from_logits=True),
                                             metrics=["accuracy"])

    process = tff.learning.build_federated_averaging_process(
        model_fn, client_optimizer_fn=lambda: tf.keras.optimizers.SGD(0.02))
    state = process.initialize()
    before = state.master.clone()
    state, metrics = process.next(state, federated_train_data)
    assert not torch.equal(before, state.master) and len(metrics) == 2
    evaluation = tff.learning.build_federated_evaluation(model_fn)
    out = evaluation(state.model, federated_train_data)
    assert len(out) == 2


def test_image_functions_match_the_dataset_pipeline(tmp_path):
    Image = pytest.importorskip("PIL.Image")
    from idc_models_b200.data.dataset import decode_img
    rng = np.random.default_rng(0)
    a = rng.integers(0, 256, (61, 47, 3), dtype=np.uint8)
    p = tmp_path / "x.png"
    Image.fromarray(a, "RGB").save(p)
    img = tf.io.read_file(str(p))
    img = tf.image.decode_png(img, channels=3)
    img = tf.image.convert_image_dtype(img, np.float32)
    assert img.shape == (61, 47, 3) and np.array_equal(img, a.astype(np.float32) / 255.0)
    out = tf.image.resize(img, [50, 50])
    assert np.abs(out - decode_img(str(p), (50, 50))).max() < 5e-6


def test_remaining_reference_names():
    """everything else the five scripts call on tf / tff / tfds (grep of /root/reference/*.py)"""
    import os
    from idc_models_b200.compat import tfds
    tf.compat.v1.enable_v2_behavior()
    parts = tf.strings.split(os.path.join("data", "1", "patch.png"), os.path.sep)
    assert (parts[-2] == "1") is True
    assert tf.cast(np.arange(3), tf.float32).dtype == np.float32 and tf.cast(torch.ones(2), tf.double).dtype == torch.float64
    assert tf.nest.map_structure(lambda a, b: a + b, {"x": [1, 2], "y": (3,)}, {"x": [10, 20], "y": (30,)}) == \
        {"x": [11, 22], "y": (33,)}
    assert tf.compat.v1.py_func(lambda a, b: a * b, [3, 4], tf.double) == 12
    assert tf.keras.applications.densenet.DenseNet201 is tf.keras.applications.DenseNet201
    datasets, info = tfds.load(name="cifar10", with_info=True, as_supervised=True, n_synthetic=40)
    image, label = next(iter(datasets["train"]))
    assert np.asarray(image).shape == (32, 32, 3) and 0 <= int(label) < 10
    assert info.splits["train"].num_examples == 40 and info.source == "synthetic"
    cd = tff.simulation.ClientData.from_clients_and_fn(list(range(10)), lambda c: c)
    train, test = tff.simulation.ClientData.train_test_client_split(cd, 2)
    assert len(train.client_ids) == 8 and len(test.client_ids) == 2
