"""Run a script written against the reference's libraries — including the reference's own, UNMODIFIED scripts — on
this framework:

    python -m idc_models_b200.run_reference /root/reference/secure_fed_model.py <path> <NUM_ROUNDS> <percent>
    python -m idc_models_b200.run_reference /root/reference/dist_model_tf_mobile.py <path>
    IDCB_DEVICE=cpu python -m idc_models_b200.run_reference my_script.py ...

The modules those scripts import (``tensorflow``, ``tensorflow_federated``, ``tensorflow_datasets``, ``phe``,
``nest_asyncio``, ``keras.utils.np_utils``, and ``matplotlib.pyplot`` when matplotlib is absent) are registered in
``sys.modules`` as thin module objects over :mod:`idc_models_b200.compat` before the script is executed with
``runpy`` as ``__main__``; ``sklearn.metrics`` is the real one when installed. Nothing of the script is edited.

This is a convenience for users switching over, not the benchmark's reference arm: every model, kernel and
collective that runs is this framework's (``bench.py --impl reference`` stays "unavailable").
``fed_model.py`` calls ``sys.path.exists`` (an AttributeError upstream); the runner gives ``sys.path`` that method
(= ``os.path.exists``, what the author meant) so the script runs; ``secure_fed_model.py`` with ``percent > 0`` builds ragged
``np.array``s, which NumPy >= 1.24 refuses: inside this process that one error case gets the old object-array result.
``IDCB_PAILLIER_BITS`` shortens the key the script generates at import (pure-Python big integers are slow at 2048).
"""
from __future__ import annotations

import os
import runpy
import sys
import types


def _module(name: str, **attrs) -> types.ModuleType:
    m = types.ModuleType(name)
    m.__dict__.update(attrs)
    m.__file__ = __file__
    return m


def _from_namespace(name: str, ns) -> types.ModuleType:
    """a module whose attributes are those of a compat namespace (nested namespaces become sub-modules)"""
    m = _module(name)
    for k, v in vars(ns).items():
        if isinstance(v, types.SimpleNamespace):
            sub = _from_namespace(f"{name}.{k}", v)
            sys.modules[f"{name}.{k}"] = sub
            setattr(m, k, sub)
        else:
            setattr(m, k, v)
    return m


class _Plt:
    """The handful of pyplot calls the scripts' ``log()`` makes, recorded and written by utils/plot.py's own PNG writer
    (used only when matplotlib is not installed)."""

    def __init__(self):
        self.panels, self.cur = [], None

    def figure(self, *a, **kw):
        self.panels, self.cur = [], None

    def subplot(self, *a, **kw):
        self.cur = {"lines": [], "title": "", "ylim": None}
        self.panels.append(self.cur)

    def _panel(self):
        if self.cur is None:
            self.subplot()
        return self.cur

    def plot(self, *args, label=None, **kw):
        ys = args[-1] if len(args) >= 1 else []
        xs = args[0] if len(args) >= 2 else list(range(len(ys)))
        self._panel()["lines"].append((list(xs), [float(v) for v in ys], label))

    def ylim(self, *args):
        p = self._panel()
        if args:
            p["ylim"] = tuple(args[0])
        return p["ylim"] or (0.0, 1.0)

    def legend(self, *a, **kw):
        pass

    def xlabel(self, *a, **kw):
        pass

    def ylabel(self, *a, **kw):
        pass

    def title(self, t, *a, **kw):
        self._panel()["title"] = str(t)

    def savefig(self, path, *a, **kw):
        from .utils import plot as P
        os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
        P.write_panels_png(path, self.panels)


def _legacy_ragged_arrays() -> None:
    """``np.array(list_of_arrays_with_different_shapes)`` gave a 1-D object array (with a deprecation warning) in the
    NumPy the reference ran on; NumPy >= 1.24 raises. secure_fed_model.py relies on the old behaviour
    (``np.array(self.model.get_weights())``, ``np.array(ave_weights)``), so within a run_reference process the old
    result is restored for exactly that error case."""
    import numpy as np
    real = np.array
    if getattr(real, "_idcb_legacy", False):
        return

    def array(obj, *args, **kw):
        try:
            return real(obj, *args, **kw)
        except ValueError as e:
            if "inhomogeneous" not in str(e) or args or "dtype" in kw:
                raise
            return real(obj, dtype=object, **kw)
    array._idcb_legacy = True
    np.array = array


def install_shims(device: str | None = None) -> None:
    """Register the stand-in modules. ``device`` (or ``IDCB_DEVICE``) picks where models are built (default: cuda when
    available)."""
    from . import compat
    from .models import applications as apps
    dev = device or os.environ.get("IDCB_DEVICE")
    if dev:
        # the scripts never name a device: give the model constructors the requested default
        seq = apps.Sequential

        def Sequential(items=None, **kw):
            kw.setdefault("device", dev)
            return seq(items, **kw)
        for ns in (compat.tf.keras, compat.tf.keras.models):
            ns.Sequential = Sequential
    # fed_model.py asks ``sys.path.exists(...)`` where it means os.path.exists (an AttributeError upstream): give the
    # module search path that method, so that the script does what its author meant
    if not hasattr(sys.path, "exists"):
        class _SearchPath(list):
            exists = staticmethod(os.path.exists)
        sys.path = _SearchPath(sys.path)
    from .fed.client_data import ClientData
    ClientData.strict_ids = False        # the TFF of the reference's time did not check ids (see client_data.py)
    _legacy_ragged_arrays()
    bits = os.environ.get("IDCB_PAILLIER_BITS")
    if bits:
        # ``paillier.generate_paillier_keypair()`` is called without arguments at import time by secure_fed_model.py;
        # a smaller modulus makes a demonstration run of the encrypted path practical on pure-Python big integers
        gen = compat.paillier.generate_paillier_keypair
        compat.paillier.generate_paillier_keypair = lambda n_length=int(bits), **kw: gen(n_length=n_length, **kw)
    tf = _from_namespace("tensorflow", compat.tf)
    tf.__version__ = "2.1.0 (idc_models_b200.compat)"
    sys.modules["tensorflow"] = tf
    sys.modules["tensorflow_federated"] = _from_namespace("tensorflow_federated", compat.tff)
    sys.modules["tensorflow_datasets"] = _from_namespace("tensorflow_datasets", compat.tfds)
    phe = _module("phe")
    phe.paillier = _from_namespace("phe.paillier", compat.paillier)
    sys.modules["phe"], sys.modules["phe.paillier"] = phe, phe.paillier
    sys.modules["nest_asyncio"] = _module("nest_asyncio", apply=lambda *a, **kw: None)

    def to_categorical(y, num_classes=None, dtype="float32"):
        import numpy as np
        y = np.asarray(y, dtype=np.int64).reshape(-1)
        n = int(num_classes or (y.max() + 1 if y.size else 0))
        out = np.zeros((y.size, n), dtype=dtype)
        out[np.arange(y.size), y] = 1
        return out
    np_utils = _module("keras.utils.np_utils", to_categorical=to_categorical)
    utils = _module("keras.utils", np_utils=np_utils, to_categorical=to_categorical)
    sys.modules.setdefault("keras", _module("keras", utils=utils))
    sys.modules["keras.utils"], sys.modules["keras.utils.np_utils"] = utils, np_utils
    sys.modules["keras"].utils = utils
    try:
        import sklearn.metrics  # noqa: F401
    except Exception:  # noqa: BLE001
        sk = _module("sklearn")
        sk.metrics = _module("sklearn.metrics", roc_auc_score=compat.roc_auc_score)
        sys.modules["sklearn"], sys.modules["sklearn.metrics"] = sk, sk.metrics
    try:
        import matplotlib
        matplotlib.use("Agg")
        import matplotlib.pyplot  # noqa: F401
    except Exception:  # noqa: BLE001
        plt = _Plt()
        pyplot = _module("matplotlib.pyplot", **{n: getattr(plt, n) for n in dir(plt) if not n.startswith("_")})
        mpl = _module("matplotlib", pyplot=pyplot, use=lambda *a, **kw: None)
        sys.modules["matplotlib"], sys.modules["matplotlib.pyplot"] = mpl, pyplot


def main(argv: list | None = None) -> int:
    argv = list(sys.argv[1:] if argv is None else argv)
    if not argv:
        print(__doc__)
        return 2
    script = argv[0]
    install_shims()
    sys.argv = [script] + argv[1:]
    runpy.run_path(script, run_name="__main__")
    return 0


if __name__ == "__main__":
    sys.exit(main())
