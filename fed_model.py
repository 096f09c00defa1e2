#!/usr/bin/env python
"""Same-named entry point as the reference's `fed_model.py`: `python fed_model.py <path> <NUM_ROUNDS> <iid|other>`.

Centralised head pre-training (checkpointed to `<path>/pretrained/cp.ckpt`, reloaded when present), then
federated averaging over simulated clients (8 train / 2 test by default) with per-round train/test metrics
printed in the reference's format. Reference: /root/reference/fed_model.py."""
import sys

from idc_models_b200 import workloads


def main():
    workloads.main_fed(sys.argv[1:])


if __name__ == "__main__":
    main()
