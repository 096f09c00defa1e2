#!/usr/bin/env python
"""Same-named entry point as the reference's `dist_model_tf_mobile.py`: `python dist_model_tf_mobile.py <path>`.

Two-phase transfer learning (10 epochs head-only, then fine-tuning from layer `fine_tune_at` at lr/10) under a
mirrored data-parallel strategy; prints the reference's Timer lines, the base-model layer count and both
History dicts and writes `<path>/logs/plot_dev{N}.png`. Multi-GPU: launch with
`python -m torch.distributed.run --nproc-per-node N dist_model_tf_mobile.py <path>`. Extra `--flags` are
documented in idc_models_b200/workloads.py (reference: /root/reference/dist_model_tf_mobile.py)."""
import sys

from idc_models_b200 import workloads


def main():
    workloads.main_transfer(workloads.MOBILE, sys.argv[1:])


if __name__ == "__main__":
    main()
