"""
TEST INFRASTRUCTURE ONLY.  Imports the UNMODIFIED reference package from /root/reference/src inside the
build container so that golden vectors can be generated from the reference's own code.
/root/reference does not exist on the GPU box: nothing under tests/ -m gpu, bench.py or smoke() calls this;
only oracle/make_golden.py and the `not gpu` cross-check tests (skipped when the tree is absent) do.

Third-party modules that the reference imports at module level but that are absent here are replaced with
inert mocks; `dgl` is replaced by oracle/dgl_standin (a pure-PyTorch restatement, see its docstring).
"""
import os
import sys
import types
from unittest import mock

REFERENCE_SRC = os.environ.get("GRAPPA_REFERENCE_SRC", "/root/reference/src")
_HERE = os.path.dirname(os.path.abspath(__file__))


def reference_available() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_SRC, "grappa"))


def _install_mocks():
    for name in ["matplotlib", "matplotlib.pyplot", "matplotlib.colors", "matplotlib.ticker", "matplotlib.cm",
                 "matplotlib.patches", "matplotlib.lines", "matplotlib.figure", "matplotlib.axes",
                 "mpl_toolkits", "mpl_toolkits.axes_grid1", "seaborn", "hydra", "hydra.utils", "omegaconf", "wandb",
                 "pytorch_lightning.loggers", "pytorch_lightning.loggers.wandb", "pytorch_lightning.trainer",
                 "pytorch_lightning.callbacks", "pytorch_lightning.callbacks.model_checkpoint",
                 "pytorch_lightning.utilities", "pytorch_lightning.utilities.types"]:
        if name not in sys.modules:
            try:
                __import__(name)
            except Exception:
                sys.modules[name] = mock.MagicMock(name=name)
    if "pytorch_lightning" not in sys.modules or isinstance(sys.modules["pytorch_lightning"], mock.MagicMock):
        try:
            import pytorch_lightning  # noqa: F401
        except Exception:
            import torch
            pl = types.ModuleType("pytorch_lightning")

            class LightningModule(torch.nn.Module):
                pass

            class LightningDataModule:
                pass

            pl.LightningModule = LightningModule
            pl.LightningDataModule = LightningDataModule
            pl.Trainer = mock.MagicMock(name="Trainer")
            pl.seed_everything = lambda *a, **k: None
            for sub in ["loggers", "trainer", "callbacks", "utilities"]:
                setattr(pl, sub, sys.modules.get(f"pytorch_lightning.{sub}", mock.MagicMock()))
            pl.__path__ = []
            sys.modules["pytorch_lightning"] = pl


_loaded = None


def load_reference():
    """Returns the imported reference package `grappa` (cached)."""
    global _loaded
    if _loaded is not None:
        return _loaded
    if not reference_available():
        raise RuntimeError(f"reference tree not found at {REFERENCE_SRC}")
    standin = os.path.join(_HERE, "dgl_standin")
    if standin not in sys.path:
        sys.path.insert(0, standin)
    import dgl  # noqa: F401  (the stand-in)
    assert "standin" in dgl.__version__
    _install_mocks()
    if REFERENCE_SRC not in sys.path:
        sys.path.insert(0, REFERENCE_SRC)
    import grappa  # noqa
    import grappa.models.grappa  # noqa
    import grappa.models.energy  # noqa
    import grappa.utils.dgl_utils  # noqa
    import grappa.utils.tuple_indices  # noqa
    _loaded = grappa
    return grappa


def default_model_config():
    """The reference's configs/model/default.yaml as a dict (BASELINE cfg 1 means the YAML)."""
    import yaml
    path = os.path.join(os.path.dirname(REFERENCE_SRC), "configs", "model", "default.yaml")
    with open(path) as f:
        return yaml.safe_load(f)
