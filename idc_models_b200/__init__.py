"""idc_models_b200 — a B200-native (sm_100a) framework with the capabilities of
``jamesnguyen123/idc_models``: DenseNet / MobileNetV2 / VGG16 IDC patch classifiers, two-phase
transfer learning under synchronous data parallelism, FedAvg and secure-aggregation FedAvg.

Layout::

    models/    DenseNet-121/169/201, MobileNetV2, VGG16, the 10x10 secure-FL CNN, Keras layer tables
    nn/        NHWC layers + freeze semantics on top of ops/
    ops/       op front-ends: a plain PyTorch fp32 implementation (CPU + oracle) and the ctypes
               bindings to the hand-written sm_100a kernels in csrc/ (the GPU path)
    optim/     flat-buffer RMSprop / SGD / Adam
    parallel/  World, Mirrored / CentralStorage strategies, gradient bucketing, fused comm kernels
    train/     compile / fit / evaluate / History, metrics, two-phase fine-tuning driver
    data/      folder PNG reader, synthetic IDC / CIFAR generators, tf.data-like pipeline
    fed/       ClientData, FedAvg process, secure (pairwise-mask) aggregation
    io/        Keras .h5 import/export (own HDF5 subset), checkpoints
    utils/     Timer, logging, plotting
    compat     the reference's import surface (tf / tff / tfds / paillier names) over the modules above
    run_reference   executes a reference-style script, incl. the upstream ones unmodified, on this framework
"""
__version__ = "0.2.0"
