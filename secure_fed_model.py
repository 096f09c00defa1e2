#!/usr/bin/env python
"""Same-named entry point as the reference's `secure_fed_model.py`:
`python secure_fed_model.py <path> <NUM_ROUNDS> <percent>`.

FedAvg over `Client`/`Server` objects where the first `int(n_tensors * percent)` weight tensors travel
protected (pairwise-mask secure aggregation instead of the reference's Paillier encryption); prints the
per-client Timer lines and `loss acc auc` per round. Reference: /root/reference/secure_fed_model.py."""
import sys

from idc_models_b200 import workloads


def main():
    workloads.main_secure(sys.argv[1:])


if __name__ == "__main__":
    main()
