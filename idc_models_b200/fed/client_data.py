"""Simulated federated client populations — the slice of ``tff.simulation.ClientData`` the reference uses
(/root/reference/fed_model.py:178-189): an id -> dataset factory and a
random train/test split over CLIENTS (not examples)."""
from __future__ import annotations

from typing import Callable, Sequence

import numpy as np


class ClientData:
    def __init__(self, client_ids: Sequence, fn: Callable):
        self.client_ids = list(client_ids)
        self._fn = fn

    @classmethod
    def from_clients_and_fn(cls, client_ids: Sequence, create_tf_dataset_for_client_fn: Callable) -> "ClientData":
        return cls(client_ids, create_tf_dataset_for_client_fn)

    # The TFF release the reference was written for handed out a dataset for ANY id (its fed_model.py asks the train
    # split for ids[0:8] although the random split may have put some of them in the test split,
    # /root/reference/fed_model.py:186-189). run_reference switches the check off to reproduce that.
    strict_ids = True

    def create_tf_dataset_for_client(self, client_id):
        if self.strict_ids and client_id not in self.client_ids:
            raise KeyError(f"unknown client {client_id!r}")
        return self._fn(client_id)

    create_dataset_for_client = create_tf_dataset_for_client

    @staticmethod
    def train_test_client_split(client_data: "ClientData", num_test_clients: int, seed: int | None = None):
        """Pick ``num_test_clients`` client ids at random as the test population (TFF semantics)."""
        ids = list(client_data.client_ids)
        if not 0 < num_test_clients < len(ids):
            raise ValueError("num_test_clients must be in (0, n_clients)")
        rng = np.random.RandomState(seed) if seed is not None else np.random
        test = set(np.asarray(ids, dtype=object)[rng.choice(len(ids), num_test_clients, replace=False)].tolist())
        train_ids = [i for i in ids if i not in test]
        test_ids = [i for i in ids if i in test]
        return ClientData(train_ids, client_data._fn), ClientData(test_ids, client_data._fn)
