"""The five workloads behind the same-named entry scripts at the repository root.

Each ``main_*`` reproduces the control flow, hyper-parameters and printed lines of
its reference script and takes the same positional arguments; optional ``--flag value`` overrides are
layered on top. When the dataset folders the reference globs
(``<path>/data/IDC_regular_ps50_idx5/*/*/*``, ``<path>/data/balanced_IDC_30k/*/*``) are absent — there is no
network here — a synthetic dataset of the same shape and size is used instead (``--synthetic`` forces it).

Known reference defects are implemented as the evident intent: ``os.path.exists`` instead
of ``sys.path.exists`` in fed_model, sparse (integer-label) softmax cross-entropy in the dense script, and
file listings that are shuffled ONCE so the take/skip splits are disjoint.
"""
from __future__ import annotations

import argparse
import glob
import os
from dataclasses import dataclass

import numpy as np
import torch

from .data import synthetic
from .data.dataset import Dataset, make_process_path, prepare_for_training
from .fed import fedavg as tff_learning
from .fed.client_data import ClientData
from .models import applications as apps
from .optim.optimizers import RMSprop
from .parallel.strategy import CentralStorageStrategy, FederatedStrategy, MirroredStrategy
from .parallel.world import World
from .train.engine import ModelCheckpoint
from .train.losses import BinaryCrossentropy, CategoricalCrossentropy
from .train.metrics import BinaryAccuracy, auroc
from .utils.plot import log
from .utils.timer import Timer


# ---------------------------------------------------------------------------------------------------------
# configuration (module-level constants of the reference, Appendix B)
# ---------------------------------------------------------------------------------------------------------
@dataclass
class TransferConfig:
    script: str
    arch: str
    img_shape: tuple = (50, 50, 3)
    units: int = 1
    batch_size: int = 32              # GLOBAL batch for mobile/vgg; per-replica for dense
    batch_per_replica: bool = False
    dataset_size: int = 24257
    dataset_glob: str = ""
    initial_epochs: int = 10
    validation_steps: int = 20
    fine_tune_epochs: int = 10
    base_learning_rate: float = 1e-4
    fine_tune_at: int = 100
    acc_ylim: tuple | None = (0.8, 1.0)
    loss_ylim: tuple | None = (0.0, 1.0)


MOBILE = TransferConfig("dist_model_tf_mobile", "mobilenet_v2", dataset_glob="data/IDC_regular_ps50_idx5/*/*/*",
                        base_learning_rate=1e-4, fine_tune_at=100)
VGG = TransferConfig("dist_model_tf_vgg", "vgg16", dataset_glob="data/balanced_IDC_30k/*/*",
                     base_learning_rate=1e-3, fine_tune_at=15, acc_ylim=None, loss_ylim=None)
DENSE = TransferConfig("dist_model_tf_dense", "densenet201", img_shape=(32, 32, 3), units=10, batch_size=256,
                       batch_per_replica=True, dataset_size=50000, base_learning_rate=1e-4, fine_tune_at=150)


def _common_flags(ap: argparse.ArgumentParser) -> None:
    ap.add_argument("--synthetic", action="store_true", help="use synthetic data even if <path>/data exists")
    ap.add_argument("--arch", default=None)
    ap.add_argument("--device", default=None, help="cpu | cuda (default: cuda when available)")
    ap.add_argument("--weights", default=None, help="Keras .h5 weight file for the base model")
    ap.add_argument("--dataset-size", type=int, default=None)
    ap.add_argument("--batch-size", type=int, default=None)
    ap.add_argument("--lr", type=float, default=None)
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--steps-per-epoch", type=int, default=None, help="cap the steps of every epoch (smoke runs)")
    ap.add_argument("--graphs", action="store_true", help="replay the fused step as a CUDA graph")
    ap.add_argument("--log-jsonl", default=None, help="append one JSON record per epoch / round to this file")
    ap.add_argument("--dtype", default="bf16", choices=["bf16", "fp32"],
                    help="compute dtype: bf16 = the fused tcgen05 pipeline on CUDA (bf16-emulating reference executor on "
                         "CPU); fp32 = the fp32 reference executor on either device")
    ap.add_argument("--gpus", type=int, default=None,
                    help="expected number of replicas/processes (one per GPU, launched by torchrun); checked, not spawned")
    ap.add_argument("--shuffle-buffer", type=int, default=1000, help="tf.data shuffle buffer (BUFFER_SIZE / shuffle(1000))")
    ap.add_argument("--train-split", type=float, default=0.8, help="fraction of the listing used for training")
    ap.add_argument("--val-split", type=float, default=None,
                    help="validation fraction (default: 0.1 for the transfer scripts, 0.2 for the federated ones)")
    ap.add_argument("--native-loader", action="store_true",
                    help="serve the cached training set from the C++ loader (threads + pinned ring + async H2D)")


def _device(args) -> str:
    if args.device:
        return args.device
    return "cuda" if torch.cuda.is_available() else "cpu"


def _executor_kind(args, device: str) -> str:
    if args.dtype == "fp32":
        return "reference"
    return "auto"        # CUDA: fused bf16 kernels; CPU: the PyTorch reference executor


def _check_gpus(args, world) -> None:
    if args.gpus is not None and args.gpus != world.size:
        raise SystemExit(f"--gpus {args.gpus}: {world.size} process(es) are running; launch one per GPU with "
                         f"`python -m torch.distributed.run --nproc-per-node {args.gpus} --master-addr 127.0.0.1 <script> ...`")


def _make_strategy(kind: str, device: str, comm: str):
    world = World.from_env(device)
    if kind == "central":
        return CentralStorageStrategy(world=world, comm=comm)
    return MirroredStrategy(world=world, comm=comm)


def _idc_dataset(path: str, pattern: str, size: int, img_shape, synthetic_only: bool, seed: int, label_as_int=False,
                 order="shuffled"):
    """The labelled patch dataset behind every script. ``order="shuffled"`` is the IID listing
    (``list_files(<pattern>, shuffle=True)``); ``order="class_sorted"`` is fed_model.py's non-IID listing
    (/root/reference/fed_model.py:161-164): the class ``1/`` files, shuffled among themselves, followed by
    the class ``0/`` files, so consecutive ``skip/take`` client shards are single-class."""
    files = [] if synthetic_only else glob.glob(os.path.join(path, pattern))
    if files:
        if order == "class_sorted":
            root = os.path.join(path, pattern)
            assert root.endswith("/*/*"), "the class-sorted listing expects a <class>/<file> layout"
            root = root[:-len("/*/*")]
            ones = list(Dataset.list_files(os.path.join(root, "1", "*"), shuffle=True, seed=seed))
            zeros = list(Dataset.list_files(os.path.join(root, "0", "*"), shuffle=True, seed=seed + 1))
            frozen = ones + zeros
        else:
            list_ds = Dataset.list_files(os.path.join(path, pattern), shuffle=True, seed=seed)
            # shuffled once, then frozen: the reference re-shuffles every epoch, which leaks train into validation
            frozen = list(list_ds)
        ds = Dataset.from_generator(lambda: frozen, len(frozen))
        return ds.map(make_process_path(img_shape[:2], label_as_int), num_parallel_calls=-1), len(frozen)
    return synthetic.synthetic_idc(size, img_shape, seed, order, label_as_int), size


def _cifar10_datasets(path: str, n_train: int, synthetic_only: bool, seed: int):
    """``tfds.load('cifar10')`` of /root/reference/dist_model_tf_dense.py:120-121: an on-disk copy under
    ``<path>/data`` (binary or python version, data/cifar.py) when there is one, else the synthetic stand-in."""
    from .data import cifar
    if not synthetic_only:
        found = cifar.find_cifar10(path)
        if found is not None:
            (xtr, ytr), (xte, yte) = cifar.load_cifar10(found)
            print(f"CIFAR-10 from {found}: {len(xtr)} train / {len(xte)} test images")
            return (Dataset.from_tensor_slices((xtr, ytr)), Dataset.from_tensor_slices((xte, yte)), len(xtr))
    n_test = max(n_train // 5, 1)
    return synthetic.synthetic_cifar10(n_train, seed), synthetic.synthetic_cifar10(n_test, seed + 1), n_train


# ---------------------------------------------------------------------------------------------------------
# W1: two-phase transfer learning under synchronous data parallelism
# ---------------------------------------------------------------------------------------------------------
def main_transfer(cfg: TransferConfig, argv: list) -> dict:
    ap = argparse.ArgumentParser(prog=cfg.script + ".py")
    ap.add_argument("path")
    _common_flags(ap)
    ap.add_argument("--initial-epochs", type=int, default=cfg.initial_epochs)
    ap.add_argument("--fine-tune-epochs", type=int, default=cfg.fine_tune_epochs)
    ap.add_argument("--fine-tune-at", type=int, default=cfg.fine_tune_at)
    ap.add_argument("--validation-steps", type=int, default=cfg.validation_steps)
    ap.add_argument("--comm", default="auto", help="auto | fused | nccl | gloo")
    ap.add_argument("--central-storage", action="store_true",
                    help="CentralStorageStrategy (use_mirror = False in the dense script)")
    ap.add_argument("--gpu-to-use", type=int, default=None, help="informational: replicas = launched processes")
    ap.add_argument("--dataset", default=None, help="dense script: cifar10 (reference) | idc (north star)")
    args = ap.parse_args(argv)
    path = args.path
    arch = args.arch or cfg.arch
    img_shape, units = cfg.img_shape, cfg.units
    dataset_kind = args.dataset or ("cifar10" if cfg.script == "dist_model_tf_dense" else "idc")
    if cfg.script == "dist_model_tf_dense" and dataset_kind == "idc":
        img_shape, units = (50, 50, 3), 1
    device = _device(args)
    strategy = _make_strategy("central" if args.central_storage else "mirrored", device, args.comm)
    _check_gpus(args, strategy.world)
    num_devices = strategy.num_replicas_in_sync
    batch = args.batch_size or cfg.batch_size
    BATCH_SIZE = batch * num_devices if cfg.batch_per_replica else batch
    DATASET_SIZE = args.dataset_size or cfg.dataset_size
    initial_epochs, fine_tune_epochs = args.initial_epochs, args.fine_tune_epochs
    total_epochs = initial_epochs + fine_tune_epochs

    if dataset_kind == "cifar10":
        # tfds.load('cifar10') stand-in; train and "validation"(=test split) each .repeat(2)
        train, test, DATASET_SIZE = _cifar10_datasets(path, DATASET_SIZE, args.synthetic, args.seed)
        train, test = train.map(synthetic.scale), test.map(synthetic.scale)
        train_batches = train.cache().shuffle(args.shuffle_buffer, seed=args.seed).batch(BATCH_SIZE).repeat(2)
        validation_batches = test.cache().shuffle(args.shuffle_buffer, seed=args.seed).batch(BATCH_SIZE).repeat(2)
    else:
        labeled_ds, DATASET_SIZE = _idc_dataset(path, cfg.dataset_glob or MOBILE.dataset_glob, DATASET_SIZE, img_shape,
                                                args.synthetic, args.seed)
        TRAIN_SIZE = int(args.train_split * DATASET_SIZE)
        VALIDATION_SIZE = int((0.1 if args.val_split is None else args.val_split) * DATASET_SIZE)
        TEST_SIZE = DATASET_SIZE - TRAIN_SIZE - VALIDATION_SIZE
        train_ds = labeled_ds.take(TRAIN_SIZE)
        validation_ds = labeled_ds.skip(TRAIN_SIZE).take(VALIDATION_SIZE)
        test_ds = labeled_ds.skip(TRAIN_SIZE + VALIDATION_SIZE).take(TEST_SIZE)
        if args.native_loader:
            from .data.loader import NativeLoader
            train_batches = NativeLoader.from_dataset(train_ds, BATCH_SIZE, args.shuffle_buffer, seed=args.seed,
                                                      device=device if device.startswith("cuda") else None)
        else:
            train_batches = prepare_for_training(train_ds, BATCH_SIZE, seed=args.seed, shuffle_buffer_size=args.shuffle_buffer)
        validation_batches = prepare_for_training(validation_ds, BATCH_SIZE, seed=args.seed, shuffle_buffer_size=args.shuffle_buffer)
        test_batches = prepare_for_training(test_ds, BATCH_SIZE, seed=args.seed,
                                            shuffle_buffer_size=args.shuffle_buffer)  # built, never used (as upstream)
        del test_batches

    base_learning_rate = args.lr or cfg.base_learning_rate
    loss = (lambda: BinaryCrossentropy(from_logits=True)) if units == 1 else (lambda: CategoricalCrossentropy(from_logits=True))
    with strategy.scope():
        base_model = apps.BaseSpec(arch, img_shape, args.weights or None)
        base_model.trainable = False
        global_average_layer = apps.GlobalAveragePooling2D()
        prediction_layer = apps.Dense(units)
        model = apps.Sequential([base_model, global_average_layer, prediction_layer], device=device, seed=args.seed,
                                executor=_executor_kind(args, device))
        model.compile(optimizer=RMSprop(lr=base_learning_rate), loss=loss(), metrics=["accuracy"])
    if args.graphs and device.startswith("cuda"):
        model.executor.use_graphs = True
    model.log_jsonl = args.log_jsonl

    loss0, accuracy0 = model.evaluate(validation_batches, steps=args.validation_steps)
    with Timer("Pre-training with " + str(num_devices) + " devices"):
        history = model.fit(train_batches, epochs=initial_epochs, validation_data=validation_batches, verbose=0,
                            steps_per_epoch=args.steps_per_epoch, validation_steps=args.steps_per_epoch)

    base_model.trainable = True
    # Let's take a look to see how many layers are in the base model
    print("Number of layers in the base model: ", len(base_model.layers))
    fine_tune_at = args.fine_tune_at
    with strategy.scope():
        for layer in base_model.layers[:fine_tune_at]:
            layer.trainable = False
        model.compile(loss=loss(), optimizer=RMSprop(lr=base_learning_rate / 10), metrics=["accuracy"])
    if args.graphs and device.startswith("cuda"):
        model.executor.use_graphs = True

    with Timer("Fine-tuning with " + str(num_devices) + " devices"):
        history_fine = model.fit(train_batches, epochs=total_epochs, initial_epoch=history.epoch[-1],
                                 validation_data=validation_batches, verbose=0,
                                 steps_per_epoch=args.steps_per_epoch, validation_steps=args.steps_per_epoch)
    if strategy.world.rank == 0:
        log(path, history, history_fine, num_devices, initial_epochs, cfg.acc_ylim, cfg.loss_ylim)
    return {"history": history, "history_fine": history_fine, "model": model, "loss0": loss0, "accuracy0": accuracy0}


# ---------------------------------------------------------------------------------------------------------
# W2: FedAvg (fed_model.py)
# ---------------------------------------------------------------------------------------------------------
def main_fed(argv: list) -> dict:
    ap = argparse.ArgumentParser(prog="fed_model.py")
    ap.add_argument("path")
    ap.add_argument("NUM_ROUNDS")
    ap.add_argument("iid", help="'iid' for an IID listing, anything else for the class-sorted (non-IID) listing")
    _common_flags(ap)
    ap.add_argument("--clients", type=int, default=10)
    ap.add_argument("--train-fraction", type=float, default=0.8)
    ap.add_argument("--client-size", type=int, default=None)
    ap.add_argument("--pretrain-epochs", type=int, default=10)
    ap.add_argument("--fine-tune-at", type=int, default=15)
    ap.add_argument("--placement", default="round_robin", choices=["round_robin", "balanced"],
                    help="client -> GPU placement under torchrun (balanced: LPT on client sizes, fed/scheduler.py)")
    ap.add_argument("--validation-steps", type=int, default=20, help="batches of the pre-training evaluate() call")
    ap.add_argument("--fed-comm", default="fused", choices=["fused", "nccl"],
                    help="aggregation path on GPUs: one fused kernel over NVLink peer memory, or the library baseline")
    args = ap.parse_args(argv)
    np.random.seed(0)
    path_data = args.path
    NUM_ROUNDS = int(args.NUM_ROUNDS)
    is_iid = args.iid == "iid"
    device = _device(args)
    arch = args.arch or "vgg16"
    NUM_CLIENTS = args.clients
    NUM_TRAIN_CLIENTS = int(NUM_CLIENTS * args.train_fraction)
    NUM_TEST_CLIENTS = NUM_CLIENTS - NUM_TRAIN_CLIENTS
    DATASET_SIZE = args.dataset_size or 30000
    CLIENT_SIZE = args.client_size or int(DATASET_SIZE / NUM_CLIENTS)
    TRAIN_SIZE = int(args.train_split * DATASET_SIZE)
    VALIDATION_SIZE = int((0.2 if args.val_split is None else args.val_split) * DATASET_SIZE)
    BATCH_SIZE = args.batch_size or 32
    IMG_SHAPE = (50, 50, 3)
    base_learning_rate = args.lr or 0.001
    fine_tune_at = args.fine_tune_at

    labeled_ds, _ = _idc_dataset(path_data, "data/balanced_IDC_30k/*/*", DATASET_SIZE, IMG_SHAPE, args.synthetic,
                                 args.seed, label_as_int=True, order="shuffled" if is_iid else "class_sorted")
    strategy = FederatedStrategy(World.from_env(device))
    _check_gpus(args, strategy.world)
    ckpt = os.path.join(path_data, "pretrained", "cp.ckpt")
    toTrain = not os.path.exists(ckpt)                       # intent of fed_model.py:175

    def pretrained(labeled_ds, path, train=False):
        checkpoint_path = os.path.join(path, "pretrained", "cp.ckpt")
        cp_callback = ModelCheckpoint(filepath=checkpoint_path, save_weights_only=True, verbose=1)
        train_ds = labeled_ds.take(TRAIN_SIZE)
        validation_ds = labeled_ds.skip(TRAIN_SIZE).take(VALIDATION_SIZE)
        train_batches = prepare_for_training(train_ds, BATCH_SIZE, seed=args.seed, shuffle_buffer_size=args.shuffle_buffer)
        validation_batches = prepare_for_training(validation_ds, BATCH_SIZE, seed=args.seed,
                                                  shuffle_buffer_size=args.shuffle_buffer)
        with strategy.scope():
            base_model = apps.BaseSpec(arch, IMG_SHAPE, args.weights or None)
            base_model.trainable = False
            pretrained_model = apps.Sequential([base_model, apps.GlobalAveragePooling2D(), apps.Dense(1)],
                                               device=device, seed=args.seed, executor=_executor_kind(args, device))
        validation_steps = args.validation_steps
        if train:
            pretrained_model.compile(optimizer=RMSprop(lr=base_learning_rate), loss=BinaryCrossentropy(from_logits=True),
                                     metrics=[BinaryAccuracy()])
            loss0, accuracy0 = pretrained_model.evaluate(validation_batches, steps=validation_steps)
            with Timer("Pre-training"):
                pretrained_model.fit(train_batches, epochs=args.pretrain_epochs, validation_data=validation_batches,
                                     callbacks=[cp_callback], steps_per_epoch=args.steps_per_epoch,
                                     validation_steps=args.steps_per_epoch)
        else:
            print("Loading pretrained model")
            pretrained_model.load_weights(checkpoint_path)
        base_model.trainable = True
        for layer in base_model.layers[:fine_tune_at]:
            layer.trainable = False
        return pretrained_model

    pretrained_model = pretrained(labeled_ds, path_data, train=toTrain)
    # every rank pre-trained (or loaded) on its own: the weight-gradient kernels accumulate with atomics, so the copies
    # differ in the last bits. The federation starts from ONE model: rank 0's.
    strategy.sync_variables(pretrained_model.store)

    def create_tf_dataset_for_client_fn(client_id):
        return labeled_ds.skip(client_id * CLIENT_SIZE).take(CLIENT_SIZE)

    client_ids = np.arange(NUM_CLIENTS).tolist()
    client_data = ClientData.from_clients_and_fn(client_ids, create_tf_dataset_for_client_fn)
    client_train, client_test = ClientData.train_test_client_split(client_data, NUM_TEST_CLIENTS, seed=0)

    def make_federated_data(cd, ids):
        return [prepare_for_training(cd.create_tf_dataset_for_client(x), BATCH_SIZE, seed=args.seed,
                                     shuffle_buffer_size=args.shuffle_buffer) for x in ids]

    federated_train_data = make_federated_data(client_train, client_train.client_ids)
    federated_test_data = make_federated_data(client_test, client_test.client_ids)
    sample_batch = next(iter(federated_train_data[0]))

    def model_fn():
        from .models.network import clone_model
        with strategy.scope():
            model = clone_model(pretrained_model)
        return tff_learning.from_keras_model(model, dummy_batch=sample_batch, loss=BinaryCrossentropy(from_logits=True),
                                             metrics=[BinaryAccuracy()])

    fed_avg = tff_learning.build_federated_averaging_process(
        model_fn, client_optimizer_fn=lambda: RMSprop(lr=base_learning_rate / 10), comm=args.fed_comm,
        placement=args.placement)
    evaluation = tff_learning.build_federated_evaluation(model_fn, placement=args.placement)

    print("Starting federated training")
    rounds = []
    with Timer("Federated training"):
        state = fed_avg.initialize()
        state = tff_learning.state_with_new_model_weights(
            state, trainable_weights=[v for v in pretrained_model.trainable_weights],
            non_trainable_weights=[v for v in pretrained_model.non_trainable_weights])
        init_metrics = evaluation(state.model, federated_test_data)
        if strategy.world.rank == 0:
            print("Initial model: {0:f} \n".format(init_metrics[0]))
        for round_num in range(int(NUM_ROUNDS)):
            state, train_metrics = fed_avg.next(state, federated_train_data)
            test_metrics = evaluation(state.model, federated_test_data)
            rounds.append((train_metrics, test_metrics))
            if strategy.world.rank == 0:
                print("{0:2d}, {1:f}, {2:f}, {3:f}, {4:f} \n".format(round_num, train_metrics[0], train_metrics[1],
                                                                    test_metrics[0], test_metrics[1]))
    return {"state": state, "rounds": rounds, "init_metrics": init_metrics}


# ---------------------------------------------------------------------------------------------------------
# W3: secure-aggregation FedAvg (secure_fed_model.py)
# ---------------------------------------------------------------------------------------------------------
def create_small_cnn(device, seed=0, lr=0.001, executor="auto"):
    """secure_fed_model.py:84-98"""
    model = apps.Sequential([
        apps.Conv2D(32, kernel_size=(3, 3), activation="relu", input_shape=(10, 10, 3), strides=2),
        apps.MaxPooling2D(pool_size=(2, 2)),
        apps.Dropout(0.25),
        apps.Flatten(),
        apps.Dense(8, activation="relu"),
        apps.Dropout(0.5),
        apps.Dense(1),
    ], device=device, seed=seed, executor=executor)
    model.compile(optimizer=RMSprop(lr=lr), loss=BinaryCrossentropy(from_logits=True), metrics=[BinaryAccuracy(), auroc])
    return model


def main_secure(argv: list) -> dict:
    from .fed.secure import Client, Server
    ap = argparse.ArgumentParser(prog="secure_fed_model.py")
    ap.add_argument("path")
    ap.add_argument("NUM_ROUNDS", type=int)
    ap.add_argument("percent", type=float)
    _common_flags(ap)
    ap.add_argument("--clients", type=int, default=2)
    ap.add_argument("--epochs", type=int, default=5)
    ap.add_argument("--verbose", type=int, default=1)
    ap.add_argument("--protect", default="mask", choices=["mask", "paillier"],
                    help="mask: pairwise Philox masks in a fixed-point ring (default); paillier: the reference's "
                         "element-wise Paillier ciphertexts (parity path, slow)")
    ap.add_argument("--paillier-bits", type=int, default=2048)
    ap.add_argument("--fused-clients", action="store_true",
                    help="map the clients onto the GPUs (torchrun: clients % GPUs == 0): all resident clients train in one "
                         "launch per step and the masked sum is one kernel across GPUs (fed/secure_fused.py)")
    args = ap.parse_args(argv)
    np.random.seed(0)
    device = _device(args)
    NUM_CLIENTS = args.clients
    keypair = None
    if args.protect == "paillier":
        from .fed.paillier import generate_paillier_keypair
        keypair = generate_paillier_keypair(args.paillier_bits)      # one pair shared by all clients (upstream :79)
    DATASET_SIZE = args.dataset_size or 30000
    TRAIN_SIZE = int(args.train_split * DATASET_SIZE)
    TEST_SIZE = int((0.2 if args.val_split is None else args.val_split) * DATASET_SIZE)
    CLIENT_SIZE = int(TRAIN_SIZE / NUM_CLIENTS)
    CLIENT_TRAIN_SIZE = int(0.8 * CLIENT_SIZE)
    CLIENT_TEST_SIZE = int(0.2 * CLIENT_SIZE)
    BATCH_SIZE = args.batch_size or 32
    IMG_SHAPE = (10, 10, 3)
    lr = args.lr or 0.001
    labeled_ds, _ = _idc_dataset(args.path, "data/balanced_IDC_30k/*/*", DATASET_SIZE, IMG_SHAPE, args.synthetic, args.seed,
                                 label_as_int=True)
    client_data = labeled_ds.take(TRAIN_SIZE)
    test_data = prepare_for_training(labeled_ds.skip(TRAIN_SIZE).take(TEST_SIZE), BATCH_SIZE, seed=args.seed,
                                     shuffle_buffer_size=args.shuffle_buffer)
    # CPU: PyTorch reference executor; CUDA: the single-kernel fused pipeline (csrc/small_cnn.cu)
    model_fn = lambda: create_small_cnn(device, args.seed, lr, executor=_executor_kind(args, device))  # noqa: E731

    def create_clients(all_data, client_ids, percent):
        return [Client(all_data.shard(NUM_CLIENTS, int(i)), int(i), percent, model_fn=model_fn, n_clients=NUM_CLIENTS,
                       batch_size=BATCH_SIZE, train_size=CLIENT_TRAIN_SIZE, test_size=CLIENT_TEST_SIZE,
                       session_seed=args.seed, protect=args.protect, keypair=keypair) for i in client_ids]

    if args.fused_clients:
        return _main_secure_fused(args, client_data, labeled_ds, NUM_CLIENTS, CLIENT_TRAIN_SIZE, TRAIN_SIZE, TEST_SIZE,
                                  BATCH_SIZE, lr, device)
    clients = create_clients(client_data, np.arange(NUM_CLIENTS), args.percent)
    server = Server(NUM_CLIENTS)
    results = []
    with Timer("Secure fed model"):
        for _ in np.arange(args.NUM_ROUNDS):
            weight_updates = []
            for c in clients:
                weights, history = c.client_fit(args.epochs, verbose=args.verbose)
                weight_updates.append(weights)
            ave_weights = server.aggregate(weight_updates)
            for c in clients:
                c.client_update(ave_weights)
            loss, acc, auc = clients[0].evaluate(test_data)
            print(loss, acc, auc)
            results.append((loss, acc, auc))
    return {"clients": clients, "server": server, "results": results}


def _materialise(ds):
    xs, ys = [], []
    for x, y in ds:
        xs.append(np.asarray(x, dtype=np.float32))
        ys.append(float(np.asarray(y).reshape(-1)[0]))
    return torch.from_numpy(np.stack(xs)), torch.tensor(ys, dtype=torch.float32)


def _main_secure_fused(args, client_data, labeled_ds, n_clients, client_train_size, train_size, test_size, batch_size,
                       lr, device) -> dict:
    """secure_fed_model.py's round loop (212-236) with the clients resident on the GPUs: same shards
    (``shard(NUM_CLIENTS, i)`` then take), same epochs per round, same printed ``loss acc auc`` line of client 0."""
    from .fed.secure_fused import FusedSecureFederation
    if not str(device).startswith("cuda"):
        raise SystemExit("--fused-clients needs CUDA devices")
    world = World.from_env(device)
    template = create_small_cnn(world.device, args.seed, lr)
    fed = FusedSecureFederation(world, n_clients, args.percent, template=template, lr=lr, batch_size=batch_size,
                                seed=args.seed, comm="fused")
    for j, cid in enumerate(fed.client_ids()):
        x, y = _materialise(client_data.shard(n_clients, int(cid)).take(client_train_size))
        fed.set_client_data(j, x, y)
    xt, yt = _materialise(labeled_ds.skip(train_size).take(min(test_size, 20 * batch_size)))
    results = []
    with Timer("Secure fed model"):
        for _ in range(args.NUM_ROUNDS):
            with Timer("Training for " + str(fed.k_local) + " resident clients"):
                fed.client_fit(args.epochs)
                torch.cuda.synchronize()
            with Timer("Masked aggregation"):
                fed.aggregate()
                fed.client_update()
                torch.cuda.synchronize()
            loss, acc, auc = fed.evaluate(xt, yt, steps=20, with_auc=True)
            if world.rank == 0:
                print(loss, acc, auc)
            results.append((loss, acc, auc))
    return {"federation": fed, "results": results}
