"""TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).  BUILD-CONTAINER ONLY.

Imports the UNMODIFIED reference Python from ``/root/reference`` so that its own code can
produce golden vectors (``oracle/make_golden.py``) and validate the oracle restatement.
``/root/reference`` does not exist on the GPU box: nothing that runs there (``-m gpu`` tests,
``smoke()``, ``bench.py``) may call :func:`load_reference`; tests that do are skipped when the
directory is absent.

What is stubbed, and why (nothing in the reference tree is touched):

* ``norse``       -> ``oracle.norse_restated`` (package absent; see that file's header)
* ``lightning``   -> two empty base classes with ``save_hyperparameters`` (package absent;
                     only used as a base class, ``models/soda.py:18``, ``utils/datasets.py:16``)
* ``torchmetrics``-> a no-op ``MeanAveragePrecision`` (absent; constructed at ``models/soda.py:89``)
* ``utils.plotter``, ``tqdm`` -> placeholders (matplotlib absent; display code is out of scope)
* ``parse_header`` is wrapped to return Python ints: on numpy >= 2 the reference's
  ``np.uint8`` event size overflows at ``prophesee_toolbox/src/io/psee_loader.py:57``
  (its pinned numpy is 1.26.4, ``environment.yml``).
"""

from __future__ import annotations

import argparse
import importlib
import os
import sys
import types

REFERENCE_ROOT = "/root/reference"

_loaded = None


def reference_available() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "models"))


class _AttrDict(dict):
    __getattr__ = dict.__getitem__
    __setattr__ = dict.__setitem__


def _make_lightning_stub():
    import inspect

    import torch

    L = types.ModuleType("lightning")

    def save_hyperparameters(self, *args, ignore=None, **kwargs):
        frame = inspect.currentframe().f_back
        init_args = {}
        code = frame.f_code
        for name in code.co_varnames[1 : code.co_argcount + code.co_kwonlyargcount]:
            if name in frame.f_locals:
                init_args[name] = frame.f_locals[name]
        for name in ignore or []:
            init_args.pop(name, None)
        self._hparams = _AttrDict(init_args)

    class LightningModule(torch.nn.Module):
        @property
        def hparams(self):
            return self._hparams

        def log(self, *a, **k):
            pass

        def log_dict(self, *a, **k):
            pass

    class LightningDataModule:
        def __init__(self):
            pass

        @property
        def hparams(self):
            return self._hparams

    LightningModule.save_hyperparameters = save_hyperparameters
    LightningDataModule.save_hyperparameters = save_hyperparameters
    L.LightningModule = LightningModule
    L.LightningDataModule = LightningDataModule
    return L


def _make_norse_stub():
    from oracle import norse_restated as nr

    norse = types.ModuleType("norse")
    nt = types.ModuleType("norse.torch")
    nt.LIFCell, nt.LICell = nr.LIFCell, nr.LICell
    nt.LIFParameters, nt.LIParameters = nr.LIFParameters, nr.LIParameters
    mod = types.ModuleType("norse.torch.module")
    snn = types.ModuleType("norse.torch.module.snn")
    snn.SNNCell, snn.SNN, snn._merge_states = nr.SNNCell, nr.SNN, nr._merge_states
    utils = types.ModuleType("norse.torch.utils")
    state = types.ModuleType("norse.torch.utils.state")
    state._is_module_stateful = nr._is_module_stateful
    norse.torch, nt.module, mod.snn, nt.utils, utils.state = nt, mod, snn, utils, state
    return {
        "norse": norse,
        "norse.torch": nt,
        "norse.torch.module": mod,
        "norse.torch.module.snn": snn,
        "norse.torch.utils": utils,
        "norse.torch.utils.state": state,
    }


def load_reference() -> argparse.Namespace:
    """Returns a namespace with the reference's own modules (models, datasets, box, ...)."""
    global _loaded
    if _loaded is not None:
        return _loaded
    if not reference_available():
        raise RuntimeError("reference tree not present (expected only in the build container)")

    stubs = {"lightning": _make_lightning_stub()}
    stubs.update(_make_norse_stub())

    tm = types.ModuleType("torchmetrics")
    tmd = types.ModuleType("torchmetrics.detection")

    class MeanAveragePrecision:
        def __init__(self, *a, **k):
            pass

    tmd.MeanAveragePrecision = MeanAveragePrecision
    tm.detection = tmd
    stubs["torchmetrics"], stubs["torchmetrics.detection"] = tm, tmd

    try:
        import tqdm  # noqa: F401
    except Exception:  # pragma: no cover
        tq = types.ModuleType("tqdm")
        tq.tqdm = type("tqdm", (), {"write": staticmethod(print)})
        stubs["tqdm"] = tq

    saved = {k: sys.modules.get(k) for k in list(stubs) + ["utils", "models"]}
    sys.modules.update(stubs)
    sys.path.insert(0, REFERENCE_ROOT)
    try:
        # ``utils/__init__.py`` imports the plotter (matplotlib); pre-seed a bare package instead
        utils_pkg = types.ModuleType("utils")
        utils_pkg.__path__ = [os.path.join(REFERENCE_ROOT, "utils")]
        sys.modules["utils"] = utils_pkg
        plot = types.ModuleType("utils.plotter")
        plot.Plotter = type("Plotter", (), {})
        sys.modules["utils.plotter"] = plot
        utils_pkg.plotter = plot

        dat = importlib.import_module("prophesee_toolbox.src.io.dat_events_tools")
        _orig_parse_header = dat.parse_header

        def parse_header_py_ints(f):
            bod, ev_type, ev_size, size = _orig_parse_header(f)
            return int(bod), int(ev_type), int(ev_size), size

        dat.parse_header = parse_header_py_ints
        psee = importlib.import_module("prophesee_toolbox.src.io.psee_loader")
        datasets = importlib.import_module("utils.datasets")
        box = importlib.import_module("utils.box")
        roi = importlib.import_module("utils.roi")
        anchors = importlib.import_module("utils.anchors")
        models = importlib.import_module("models")
        generator = importlib.import_module("models.generator")
        modules = importlib.import_module("models.modules")
    finally:
        sys.path.remove(REFERENCE_ROOT)

    _loaded = argparse.Namespace(
        models=models,
        generator=generator,
        modules=modules,
        datasets=datasets,
        box=box,
        roi=roi,
        anchors=anchors,
        psee=psee,
        dat=dat,
    )
    return _loaded
