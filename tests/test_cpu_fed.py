"""CPU unit tests: FedAvg algebra, client data split, secure masked sum."""
import numpy as np
import pytest
import torch
from hypothesis import given, settings, strategies as st

from idc_models_b200.fed import fedavg, secure
from idc_models_b200.fed.client_data import ClientData
from idc_models_b200.ops import _lib


def test_weighted_delta_mean_equals_weighted_model_average():
    torch.manual_seed(0)
    w0 = torch.randn(64)
    clients = torch.randn(5, 64)
    n = torch.tensor([3000., 1000., 2000., 500., 3500.])
    got = fedavg.weighted_delta_mean(w0, clients, n, server_lr=1.0)
    want = (clients * n[:, None]).sum(0) / n.sum()
    assert torch.allclose(got, want, atol=1e-5)


def test_non_finite_client_is_dropped_and_non_trainable_untouched():
    w0 = torch.zeros(32)
    clients = torch.ones(3, 32)
    clients[1, 5] = float("nan")
    clients[2] *= 3
    n = torch.tensor([1., 5., 1.])
    got = fedavg.weighted_delta_mean(w0, clients, n, ranges=[(0, 16)])
    assert torch.allclose(got[:16], torch.full((16,), 2.0))     # mean of clients 0 and 2 only
    assert torch.equal(got[16:], w0[16:])                       # outside the trainable range


def test_server_lr_scales_the_update():
    w0 = torch.zeros(8)
    clients = torch.ones(2, 8)
    got = fedavg.weighted_delta_mean(w0, clients, torch.tensor([1., 1.]), server_lr=0.5)
    assert torch.allclose(got, torch.full((8,), 0.5))


def test_client_data_split_is_disjoint_and_deterministic():
    cd = ClientData.from_clients_and_fn(list(range(10)), lambda i: [i])
    tr, te = ClientData.train_test_client_split(cd, 2, seed=0)
    assert len(tr.client_ids) == 8 and len(te.client_ids) == 2
    assert set(tr.client_ids).isdisjoint(te.client_ids)
    tr2, te2 = ClientData.train_test_client_split(cd, 2, seed=0)
    assert te.client_ids == te2.client_ids
    assert tr.create_tf_dataset_for_client(tr.client_ids[0]) == [tr.client_ids[0]]
    with pytest.raises(KeyError):
        tr.create_tf_dataset_for_client(te.client_ids[0])


def test_philox_known_answer():
    """Random123 known-answer test vectors for philox4x32-10."""
    import ctypes as C
    lib = _lib.host_lib()
    out = (C.c_uint32 * 4)()
    lib.idcbh_philox(C.c_uint32(0), C.c_uint32(0), C.c_uint32(0), C.c_uint32(0), C.c_uint32(0), C.c_uint32(0), out)
    assert [hex(v) for v in out] == ["0x6627e8d5", "0xe169c58d", "0xbc57ac4c", "0x9b00dbd8"]
    lib.idcbh_philox(C.c_uint32(0xffffffff), C.c_uint32(0xffffffff), C.c_uint32(0xffffffff), C.c_uint32(0xffffffff),
                     C.c_uint32(0xffffffff), C.c_uint32(0xffffffff), out)
    assert [hex(v) for v in out] == ["0x408f276d", "0x41c83b0e", "0xa20bc7c6", "0x6d5451fd"]


@settings(max_examples=25, deadline=None)
@given(st.integers(2, 9), st.integers(1, 40), st.integers(0, 2 ** 40))
def test_masks_cancel_exactly_in_the_ring(k, n, seed):
    rng = np.random.RandomState(seed % (2 ** 31))
    w = (rng.randn(k, n) * 3).astype(np.float32)
    y = secure.mask_vectors(w, seed)
    s = secure.ring_sum(list(y))
    q = np.rint(w.astype(np.float64) * (1 << secure.FRAC_BITS)).astype(np.int64).sum(0)
    assert np.array_equal(s.view(np.int32).astype(np.int64), q)           # bit-exact cancellation
    mean = secure.dequantise_mean(s, k)
    assert np.abs(mean - w.astype(np.float64).mean(0)).max() <= 0.5 / (1 << secure.FRAC_BITS) + 1e-7
    # an individual masked vector reveals nothing obvious: it is far from the quantised weights
    if n >= 8:
        assert np.mean(y[0].view(np.int32) == np.rint(w[0] * (1 << secure.FRAC_BITS)).astype(np.int32)) < 0.5


def test_percent_selects_the_tensor_prefix_and_round_trip_matches_float_mean():
    from idc_models_b200.workloads import create_small_cnn
    model_fn = lambda: create_small_cnn("cpu", executor="reference")  # noqa: E731
    from idc_models_b200.data import synthetic
    ds = synthetic.synthetic_idc(64, (10, 10, 3), label_as_int=True)
    clients = [secure.Client(ds.shard(3, i), i, 0.5, model_fn=model_fn, n_clients=3, session_seed=7) for i in range(3)]
    for i, c in enumerate(clients):
        for j, t in enumerate(c.model.trainable_weights):
            t.fill_(0.1 * (i + 1) + 0.01 * j)
    sent = [c.enc_model(c.model.get_weights()) for c in clients]
    assert clients[0].n_protected(6) == 3
    assert [isinstance(t, secure.MaskedTensor) for t in sent[0]] == [True, True, True, False, False, False]
    ave = secure.Server(3).aggregate(sent)
    dec = clients[0].dec_model(ave)
    want = [np.mean([c.model.get_weights()[t] for c in clients], axis=0) for t in range(6)]
    for a, b in zip(dec, want):
        assert np.abs(a - b).max() < 1e-4
    with pytest.raises(RuntimeError):
        secure.Server(3).aggregate(sent[:2])          # a missing client must abort the round


def test_paillier_homomorphic_mean_matches_float_mean():
    """the reference's protection mechanism (secure_fed_model.py:79,109-121,160-168) restated on Python ints"""
    import numpy as np
    from idc_models_b200.fed.paillier import generate_paillier_keypair
    pk, sk = generate_paillier_keypair(512, seed=1)
    a, b, c = 0.37251, -1.25, 3.0e-4
    ea, eb, ec = pk.encrypt(a), pk.encrypt(b), pk.encrypt(c)
    assert abs(sk.decrypt(ea) - a) < 1e-9 and abs(sk.decrypt(eb) - b) < 1e-9
    assert abs(sk.decrypt(ea + eb) - (a + b)) < 1e-9
    assert abs(sk.decrypt((ea + eb + ec) / 3) - (a + b + c) / 3) < 1e-9
    assert abs(sk.decrypt(ea * -2.5) - a * -2.5) < 1e-9
    assert ea.c != pk.encrypt(a).c                                  # probabilistic encryption
    # object arrays, as Client.enc_model / Server.aggregate use them
    w1, w2 = np.random.RandomState(0).randn(3, 2), np.random.RandomState(1).randn(3, 2)
    enc = np.vectorize(pk.encrypt, otypes=[object])
    dec = np.vectorize(sk.decrypt, otypes=[np.float64])
    assert np.allclose(dec((enc(w1) + enc(w2)) / 2.0), (w1 + w2) / 2, atol=1e-9)


def test_secure_clients_paillier_path_equals_mask_path():
    import numpy as np
    from idc_models_b200.fed.paillier import generate_paillier_keypair
    from idc_models_b200.fed.secure import Client, Server
    from idc_models_b200.data.dataset import Dataset
    kp = generate_paillier_keypair(512, seed=2)
    rs = np.random.RandomState(0)
    weights = [[rs.randn(2, 3).astype(np.float32), rs.randn(3).astype(np.float32), rs.randn(4).astype(np.float32)]
               for _ in range(3)]
    ds = Dataset.from_tensor_slices((np.zeros((4, 2, 2, 3), np.float32), np.zeros(4, np.float32)))
    results = {}
    for protect in ("mask", "paillier"):
        clients = [Client(ds, i, 0.67, model_fn=lambda: None, n_clients=3, protect=protect, keypair=kp) for i in range(3)]
        assert clients[0].n_protected(3) == 2
        updates = [c.enc_model(weights[i]) for i, c in enumerate(clients)]
        assert not isinstance(updates[0][2], np.ndarray) or updates[0][2].dtype != object     # third tensor in the clear
        ave = Server(3).aggregate(updates)
        results[protect] = clients[1].dec_model(ave)
    want = [np.mean([w[i] for w in weights], axis=0) for i in range(3)]
    for got in results.values():
        for g, w in zip(got, want):
            assert np.allclose(g, w, atol=2e-5)


def test_client_scheduler_balances_and_is_deterministic():
    from idc_models_b200.fed.scheduler import assign_clients, clients_of_rank
    assert assign_clients([100] * 8, 4) == [0, 1, 2, 3, 0, 1, 2, 3]          # equal sizes: round-robin
    costs = [900, 100, 100, 100, 400, 400, 50, 50]
    plan = assign_clients(costs, 2)
    loads = [sum(c for c, r in zip(costs, plan) if r == k) for k in range(2)]
    assert sorted(loads) == [1050, 1050] or max(loads) <= 4 / 3 * (sum(costs) / 2)
    assert plan == assign_clients(costs, 2)
    assert sorted(clients_of_rank(costs, 2, 0) + clients_of_rank(costs, 2, 1)) == list(range(8))
    assert assign_clients([5, 3], 1) == [0, 0]


def test_paillier_crt_decryption_equals_the_textbook_form():
    """the Chinese-remainder decryption (two half-length exponentiations) against L(c^lambda mod n^2) * mu mod n"""
    import random
    from idc_models_b200.fed import paillier
    for bits in (256, 512, 1024):
        pk, sk = paillier.generate_paillier_keypair(n_length=bits, seed=bits)
        rng = random.Random(bits)
        for _ in range(12):
            m = rng.randrange(0, pk.n)
            c = pk.raw_encrypt(m)
            assert sk.raw_decrypt(c) == sk.raw_decrypt_plain(c) == m
        for edge in (0, 1, pk.n - 1):
            assert sk.raw_decrypt(pk.raw_encrypt(edge)) == edge
        mean = (pk.encrypt(3.0) + pk.encrypt(-5.5) + pk.encrypt(0.25)) / 3
        assert abs(sk.decrypt(mean) - (3.0 - 5.5 + 0.25) / 3) < 1e-6
