"""``python -m idc_models_b200`` — what is installed and how to drive it.

    python -m idc_models_b200            # environment report: versions, devices, native libraries, kernels
    python -m idc_models_b200 build      # = python -m idc_models_b200.build
    python -m idc_models_b200 run <reference-style script> [args]   # = python -m idc_models_b200.run_reference
"""
from __future__ import annotations

import os
import sys


def report() -> int:
    import torch
    from . import __version__
    from .ops import _lib
    print(f"idc_models_b200 {__version__}  (python {sys.version.split()[0]}, torch {torch.__version__})")
    if torch.cuda.is_available():
        for i in range(torch.cuda.device_count()):
            p = torch.cuda.get_device_properties(i)
            print(f"  cuda:{i}  {p.name}  sm_{p.major}{p.minor}  {p.multi_processor_count} SMs  {p.total_memory / 2**30:.0f} GiB")
    else:
        print("  no CUDA device: models run on the PyTorch fp32 executor (CPU); the native libraries still load")
    native = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_native")
    for name, loader, prefix in (("libidcb200.so", _lib.cuda_lib, "idcb_"), ("libidcb200_host.so", _lib.host_lib, "idcbh_")):
        path = os.path.join(native, name)
        try:
            loader()
            size = os.path.getsize(path) / 2**20
            print(f"  {name:22s} loaded ({size:.1f} MiB)")
        except Exception as e:  # noqa: BLE001
            print(f"  {name:22s} NOT loadable: {e}")
    print("  entry points: dist_model_tf_{dense,mobile,vgg}.py, fed_model.py, secure_fed_model.py (repository root);")
    print("                bench.py; python -m idc_models_b200.run_reference <script>; idc_models_b200.compat")
    return 0


def main() -> int:
    argv = sys.argv[1:]
    if not argv:
        return report()
    if argv[0] == "build":
        from . import build
        build.build_all(force="--force" in argv, verbose="-v" in argv)
        print("built", build.CUDA_LIB, build.HOST_LIB)
        return 0
    if argv[0] == "run":
        from . import run_reference
        return run_reference.main(argv[1:])
    print(__doc__)
    return 2


if __name__ == "__main__":
    sys.exit(main())
